// Host-side filter design and scan tables for kernel (b).  Extended precision (x87 long double)
// is used for every matrix power so that the tables are correctly rounded doubles.
#include "tables.h"
#include <cmath>
#include <complex>
#include <cstring>
#include <limits>

namespace cbl {

// ------------------------------------------------------------------ Butterworth design
// Same construction as iir1 so that the coefficients are the reference's to the last bit:
// analog prototype poles on the unit circle, prewarped bilinear transform, zeros at z = -1,
// a0-normalised biquads, whole-cascade DC gain folded into stage 0's numerator.
namespace {
typedef std::complex<double> cd;
const double PI = 3.1415926535897932384626433832795028841971;     // iir1 MathSupplement.h:51
const double PI_2 = 1.5707963267948966192313216916397514420986;   // iir1 MathSupplement.h:52

struct Bq { double b0, b1, b2, a1, a2; };

cd dc_response(const Bq* q, int n)
{
    // Cascade::response(0): numerator and denominator accumulated separately, then divided
    const cd z1 = std::polar(1., -0.), z2 = std::polar(1., -0.);
    cd num(1), den(1);
    for (int i = 0; i < n; i++) {
        cd t(q[i].b0), b(1);
        t = cd(t.real() + q[i].b1 * z1.real(), t.imag() + q[i].b1 * z1.imag());
        t = cd(t.real() + q[i].b2 * z2.real(), t.imag() + q[i].b2 * z2.imag());
        b = cd(b.real() + q[i].a1 * z1.real(), b.imag() + q[i].a1 * z1.imag());
        b = cd(b.real() + q[i].a2 * z2.real(), b.imag() + q[i].a2 * z2.imag());
        num *= t;
        den *= b;
    }
    return num / den;
}
}  // namespace

int butter_lowpass_sos(int order, double fs, double fc_hz, double* sos)
{
    if (order < 1 || order > 2 * kMaxK) return -1;
    const double fc = fc_hz / fs;
    if (!(fc < 0.5) || fc < 0.0) return -1;                 // PoleFilter.cpp:62-63
    const double warp = std::tan(PI * fc);
    Bq q[kMaxK];
    int n = 0;
    for (int i = 0; i < order / 2; i++) {
        const cd sp = std::polar(1., PI_2 + (2 * i + 1) * PI / (2.0 * order));   // Butterworth.cpp:61-66
        const cd fs_ = warp * sp;
        const cd zp = (1. + fs_) / (1. - fs_);                                 // PoleFilter.cpp:47-57
        // conjugate pole pair + double zero at -1 (Biquad.cpp:156-201)
        Bq b;
        b.a1 = -2 * zp.real();
        b.a2 = std::norm(zp);
        b.b0 = 1; b.b1 = -(-1.0 + -1.0); b.b2 = (-1.0) * (-1.0);
        q[n++] = b;
    }
    if (order & 1) {
        const cd fs_ = warp * cd(-1, 0);
        const cd zp = (1. + fs_) / (1. - fs_);
        Bq b;                                                                  // Biquad.cpp:141-154
        b.a1 = -zp.real(); b.a2 = 0; b.b0 = 1; b.b1 = 1; b.b2 = 0;
        q[n++] = b;
    }
    const double scale = 1.0 / std::abs(dc_response(q, n));                    // Cascade.cpp:116-117
    q[0].b0 *= scale; q[0].b1 *= scale; q[0].b2 *= scale;
    for (int i = 0; i < n; i++) {
        sos[6 * i + 0] = q[i].b0; sos[6 * i + 1] = q[i].b1; sos[6 * i + 2] = q[i].b2;
        sos[6 * i + 3] = 1.0;     sos[6 * i + 4] = q[i].a1; sos[6 * i + 5] = q[i].a2;
    }
    return n;
}

// ------------------------------------------------------------------ small dense algebra (long double)
namespace {
typedef long double ld;
struct Mat { ld a[kMaxS][kMaxS]; };

Mat mat_mul(const Mat& x, const Mat& y, int S)
{
    Mat r;
    for (int i = 0; i < S; i++)
        for (int j = 0; j < S; j++) {
            ld acc = 0;
            for (int k = 0; k < S; k++) acc += x.a[i][k] * y.a[k][j];
            r.a[i][j] = acc;
        }
    return r;
}
Mat mat_eye(int S)
{
    Mat r;
    for (int i = 0; i < S; i++)
        for (int j = 0; j < S; j++) r.a[i][j] = (i == j);
    return r;
}
Mat mat_pow(Mat b, uint64_t n, int S)
{
    Mat r = mat_eye(S);
    while (n) {
        if (n & 1) r = mat_mul(r, b, S);
        n >>= 1;
        if (n) b = mat_mul(b, b, S);
    }
    return r;
}
void mat_store(const Mat& m, int S, double* out)
{
    for (int i = 0; i < S; i++)
        for (int j = 0; j < S; j++) out[i * S + j] = (double)m.a[i][j];
}
// scaled max-norm: max_ij |m_ij| * scale_j / scale_i  (how much of a typical state survives)
ld scaled_norm(const Mat& m, int S, const std::vector<ld>& sc)
{
    ld v = 0;
    for (int i = 0; i < S; i++)
        for (int j = 0; j < S; j++) {
            const ld e = fabsl(m.a[i][j]) * sc[j] / sc[i];
            if (e > v) v = e;
        }
    return v;
}
}  // namespace

bool make_filter_plan(int K, const double (*sos)[6], int D, FilterPlan* P, const char** msg)
{
    static const char* e_k = "n_sos must be 1..4";
    static const char* e_d = "decimation must be one of 1,2,4,8,16,32";
    static const char* e_c = "SOS coefficients must be finite with a0 != 0 and b0 != 0";
    if (K < 1 || K > kMaxK) { *msg = e_k; return false; }
    if (!(D == 1 || D == 2 || D == 4 || D == 8 || D == 16 || D == 32)) { *msg = e_d; return false; }
    const int S = 2 * K;
    P->K = K; P->D = D; P->S = S;
    P->LT = iir_run_len(D);
    P->NK = P->LT / D;
    std::memset(&P->tables, 0, sizeof(P->tables));
    ld a1[kMaxK], a2[kMaxK], c1[kMaxK], c2[kMaxK];
    P->g[0] = 1.0;
    for (int k = 0; k < K; k++) {
        // Biquad::setCoefficients (Biquad.cpp:123-139): everything divided by a0, in double
        const double a0 = sos[k][3];
        const double b0 = sos[k][0] / a0, b1 = sos[k][1] / a0, b2 = sos[k][2] / a0;
        const double A1 = sos[k][4] / a0, A2 = sos[k][5] / a0;
        if (!std::isfinite(b0) || !std::isfinite(b1) || !std::isfinite(b2) || !std::isfinite(A1) ||
            !std::isfinite(A2) || a0 == 0.0 || b0 == 0.0) { *msg = e_c; return false; }
        a1[k] = A1; a2[k] = A2;
        c1[k] = (ld)b1 / (ld)b0; c2[k] = (ld)b2 / (ld)b0;
        P->tables.a1[k] = A1; P->tables.a2[k] = A2;
        P->tables.c1[k] = (double)c1[k]; P->tables.c2[k] = (double)c2[k];
        P->g[k + 1] = (double)((ld)P->g[k] * (ld)b0);
    }
    P->tables.gain = P->g[K];
    // use the rounded (double) c's in the model too, so model and kernel agree
    for (int k = 0; k < K; k++) { c1[k] = P->tables.c1[k]; c2[k] = P->tables.c2[k]; }

    // one-sample transition in kernel coordinates: columns = images of unit states with x = 0;
    // Cu = in'_K as a function of the state; Bv = image of x = 1 from zero state
    Mat A;
    ld Cu[kMaxS], Bv[kMaxS];
    for (int j = 0; j <= S; j++) {
        ld s[kMaxS];
        for (int i = 0; i < S; i++) s[i] = (i == j);
        ld in = (j == S) ? 1 : 0;
        ld ns[kMaxS];
        for (int k = 0; k < K; k++) {
            const ld w = in - a1[k] * s[2 * k] - a2[k] * s[2 * k + 1];
            in = w + c1[k] * s[2 * k] + c2[k] * s[2 * k + 1];
            ns[2 * k] = w; ns[2 * k + 1] = s[2 * k];
        }
        if (j < S) { for (int i = 0; i < S; i++) A.a[i][j] = ns[i]; Cu[j] = in; }
        else       { for (int i = 0; i < S; i++) Bv[i] = ns[i]; }
    }
    for (int i = 0; i < S; i++) for (int j = 0; j < S; j++) P->A[i][j] = A.a[i][j];

    // typical magnitude of each state component: peak over frequency of |(zI - A)^-1 B|
    P->scale.assign(S, 0);
    {
        const int NF = 2048;
        for (int f = 0; f <= NF; f++) {
            const ld w = 3.14159265358979323846264338327950288L * f / NF;
            const std::complex<ld> z(cosl(w), sinl(w));
            // solve (zI - A) x = B by Gaussian elimination with partial pivoting
            std::complex<ld> m[kMaxS][kMaxS + 1];
            for (int i = 0; i < S; i++) {
                for (int j = 0; j < S; j++) m[i][j] = std::complex<ld>((i == j), 0) * z - std::complex<ld>(A.a[i][j], 0);
                m[i][S] = Bv[i];
            }
            bool ok = true;
            for (int c = 0; c < S && ok; c++) {
                int piv = c;
                for (int r = c + 1; r < S; r++) if (std::abs(m[r][c]) > std::abs(m[piv][c])) piv = r;
                if (std::abs(m[piv][c]) == 0) { ok = false; break; }
                if (piv != c) for (int j = 0; j <= S; j++) std::swap(m[piv][j], m[c][j]);
                for (int r = 0; r < S; r++) {
                    if (r == c) continue;
                    const std::complex<ld> fct = m[r][c] / m[c][c];
                    for (int j = c; j <= S; j++) m[r][j] -= fct * m[c][j];
                }
            }
            if (!ok) continue;
            for (int i = 0; i < S; i++) {
                const ld v = std::abs(m[i][S] / m[i][i]);
                if (std::isfinite((double)v) && v > P->scale[i]) P->scale[i] = v;
            }
        }
        for (int i = 0; i < S; i++) if (!(P->scale[i] > 0)) P->scale[i] = 1;
    }

    // R[m] = Cu A^(mD): walk the row vector through the run
    {
        ld r[kMaxS];
        for (int i = 0; i < S; i++) r[i] = Cu[i];
        for (int n = 0; n < P->LT; n++) {
            if (n % D == 0) for (int i = 0; i < S; i++) P->tables.R[n / D][i] = (double)r[i];
            ld t[kMaxS];
            for (int j = 0; j < S; j++) {
                ld acc = 0;
                for (int i = 0; i < S; i++) acc += r[i] * A.a[i][j];
                t[j] = acc;
            }
            for (int j = 0; j < S; j++) r[j] = t[j];
        }
    }
    // M = A^LT and its powers
    const Mat M = mat_pow(A, (uint64_t)P->LT, S);
    {
        Mat p = M;
        for (int j = 0; j < 5; j++) {
            // Mp is stored with the kernel's fixed leading dimension S (row*S + col)
            for (int i = 0; i < S; i++) for (int c = 0; c < S; c++) P->tables.Mp[j][i * S + c] = (double)p.a[i][c];
            p = mat_mul(p, p, S);
        }
    }
    P->plane.assign((size_t)S * S * 33, 0.0);
    {
        Mat p = mat_eye(S);
        for (int t = 0; t <= 32; t++) {
            for (int i = 0; i < S; i++) for (int c = 0; c < S; c++) P->plane[(size_t)(i * S + c) * 33 + t] = (double)p.a[i][c];
            p = mat_mul(p, M, S);
        }
    }
    // look-back window: smallest J with scaled |M_tile^J| < 2^-64
    const Mat Mtile = mat_pow(M, 32, S);
    const int JMAX = 16;
    const ld tol = ldexpl(1.0L, -64);
    P->window = 0; P->chained = true;
    {
        Mat p = Mtile;
        for (int j = 1; j <= JMAX; j++) {
            if (scaled_norm(p, S, P->scale) < tol) { P->window = j; P->chained = false; break; }
            p = mat_mul(p, Mtile, S);
        }
    }
    if (P->chained) P->window = 1;
    P->mtile_pow.assign((size_t)(P->window + 1) * S * S, 0.0);
    {
        Mat p = mat_eye(S);
        for (int j = 0; j <= P->window; j++) {
            mat_store(p, S, &P->mtile_pow[(size_t)j * S * S]);
            p = mat_mul(p, Mtile, S);
        }
    }
    // decay length in samples (for the multi-GPU fix-up): first n (multiple of 64) such that the
    // scaled |A^n| stays < 2^-60 ... checked on a 64-sample grid and through the tile powers
    P->decay_samples = 0;
    if (!P->chained) {
        const Mat A64 = mat_pow(A, 64, S);
        Mat p = A64;
        const ld tol60 = ldexpl(1.0L, -60);
        uint64_t n = 64, last_bad = 0;
        const uint64_t limit = (uint64_t)P->window * 32 * P->LT + 64;
        while (n <= limit) {
            if (!(scaled_norm(p, S, P->scale) < tol60)) last_bad = n;
            p = mat_mul(p, A64, S);
            n += 64;
        }
        P->decay_samples = last_bad + 64;
    }
    return true;
}

void plan_shard_matrix(const FilterPlan& P, uint64_t n, double* m)
{
    const int S = P.S;
    Mat A;
    for (int i = 0; i < S; i++) for (int j = 0; j < S; j++) A.a[i][j] = P.A[i][j];
    const Mat p = mat_pow(A, n, S);
    // reference coordinates: s_ref[i] = g[stage(i)] * s_kernel[i]
    for (int i = 0; i < S; i++)
        for (int j = 0; j < S; j++)
            m[i * S + j] = (double)(p.a[i][j] * (ld)P.g[i / 2] / (ld)P.g[j / 2]);
}

}  // namespace cbl
