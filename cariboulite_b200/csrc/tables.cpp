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
typedef std::complex<ld> cl;
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
    for (int i = 0; i < kMaxS; i++)
        for (int j = 0; j < kMaxS; j++) r.a[i][j] = (i == j && i < S);
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
// Gauss-Jordan with partial pivoting on an (n x n | m) augmented system, in place.  false if singular.
bool solve_inplace(ld* a, int n, int m)
{
    const int w = n + m;
    for (int c = 0; c < n; c++) {
        int piv = c;
        for (int r = c + 1; r < n; r++) if (fabsl(a[r * w + c]) > fabsl(a[piv * w + c])) piv = r;
        if (!(fabsl(a[piv * w + c]) > 0)) return false;
        if (piv != c) for (int j = 0; j < w; j++) std::swap(a[piv * w + j], a[c * w + j]);
        const ld inv = 1 / a[c * w + c];
        for (int j = c; j < w; j++) a[c * w + j] *= inv;
        for (int r = 0; r < n; r++) {
            if (r == c) continue;
            const ld f = a[r * w + c];
            if (f == 0) continue;
            for (int j = c; j < w; j++) a[r * w + j] -= f * a[c * w + j];
        }
    }
    return true;
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

// ------------------------------------------------------------------ the two realisations
struct Form {
    int K; bool par;
    ld a1[kMaxK], a2[kMaxK], c1[kMaxK], c2[kMaxK];    // values the kernel uses (already rounded to double)
};
// one sample: ns = A s + B x, out = Cu s + x (this is y / G)
void form_step(const Form& F, const ld* s, ld x, ld* ns, ld* out)
{
    if (F.par) {
        ld o = x;
        for (int k = 0; k < F.K; k++) {
            o += F.c1[k] * s[2 * k] + F.c2[k] * s[2 * k + 1];
            ns[2 * k] = x - F.a1[k] * s[2 * k] - F.a2[k] * s[2 * k + 1];
            ns[2 * k + 1] = s[2 * k];
        }
        *out = o;
    } else {
        ld in = x;
        for (int k = 0; k < F.K; k++) {
            const ld w = in - F.a1[k] * s[2 * k] - F.a2[k] * s[2 * k + 1];
            in = w + F.c1[k] * s[2 * k] + F.c2[k] * s[2 * k + 1];
            ns[2 * k] = w; ns[2 * k + 1] = s[2 * k];
        }
        *out = in;
    }
}
struct Realisation { Mat A; ld B[kMaxS], Cu[kMaxS]; };
Realisation realise(const Form& F)
{
    Realisation R;
    const int S = 2 * F.K;
    R.A = mat_eye(0);
    for (int j = 0; j <= S; j++) {
        ld s[kMaxS], ns[kMaxS], out;
        for (int i = 0; i < S; i++) s[i] = (i == j);
        form_step(F, s, j == S ? 1 : 0, ns, &out);
        if (j < S) { for (int i = 0; i < S; i++) R.A.a[i][j] = ns[i]; R.Cu[j] = out; }
        else       { for (int i = 0; i < S; i++) R.B[i] = ns[i]; }
    }
    return R;
}
// typical magnitude of each state component: peak over frequency of |(zI - A)^-1 B|
std::vector<ld> state_scale(const Realisation& R, int S)
{
    std::vector<ld> scale(S, 0);
    const int NF = 2048;
    for (int f = 0; f <= NF; f++) {
        const ld w = 3.14159265358979323846264338327950288L * f / NF;
        const cl z(cosl(w), sinl(w));
        cl m[kMaxS][kMaxS + 1];
        for (int i = 0; i < S; i++) {
            for (int j = 0; j < S; j++) m[i][j] = cl((i == j), 0) * z - cl(R.A.a[i][j], 0);
            m[i][S] = R.B[i];
        }
        bool ok = true;
        for (int c = 0; c < S && ok; c++) {
            int piv = c;
            for (int r = c + 1; r < S; r++) if (std::abs(m[r][c]) > std::abs(m[piv][c])) piv = r;
            if (std::abs(m[piv][c]) == 0) { ok = false; break; }
            if (piv != c) for (int j = 0; j <= S; j++) std::swap(m[piv][j], m[c][j]);
            for (int r = 0; r < S; r++) {
                if (r == c) continue;
                const cl fct = m[r][c] / m[c][c];
                for (int j = c; j <= S; j++) m[r][j] -= fct * m[c][j];
            }
        }
        if (!ok) continue;
        for (int i = 0; i < S; i++) {
            const ld v = std::abs(m[i][S] / m[i][i]);
            if (std::isfinite((double)v) && v > scale[i]) scale[i] = v;
        }
    }
    for (int i = 0; i < S; i++) if (!(scale[i] > 0)) scale[i] = 1;
    return scale;
}

// Partial fractions of Hn(q) = prod (1 + c1 q + c2 q^2) / prod (1 + a1 q + a2 q^2), q = z^-1:
//   Hn(q) = 1 + sum_k (alpha_k q + beta_k q^2) / D_k(q)
// by evaluating (Hn(q) - 1) / q * D_k(q) = Nn(q) / (q prod_{j != k} D_j(q)) at the roots of D_k.
bool partial_fractions(int K, const ld* a1, const ld* a2, const ld* c1, const ld* c2, ld* alpha, ld* beta, const char** why)
{
    int degN = 0, degD = 0;
    for (int k = 0; k < K; k++) {
        degN += c2[k] != 0 ? 2 : (c1[k] != 0 ? 1 : 0);
        degD += a2[k] != 0 ? 2 : (a1[k] != 0 ? 1 : 0);
        if (a1[k] == 0 && a2[k] == 0) { *why = "a section without poles"; return false; }
    }
    if (degN > degD) { *why = "more zeros than poles"; return false; }
    for (int k = 0; k < K; k++) {
        cl q[2];
        int nr;
        if (a2[k] == 0) { q[0] = cl(-1 / a1[k], 0); nr = 1; }
        else {
            // poles z: z^2 + a1 z + a2 = 0; q = 1 / z
            const cl disc = std::sqrt(cl(a1[k] * a1[k] - 4 * a2[k], 0));
            const cl z0 = (-a1[k] + disc) / cl(2, 0), z1 = (-a1[k] - disc) / cl(2, 0);
            if (std::abs(z0 - z1) < 1e-9L * std::abs(z0)) { *why = "a double pole inside one section"; return false; }
            q[0] = cl(1, 0) / z0; q[1] = cl(1, 0) / z1; nr = 2;
        }
        cl rho[2];
        for (int r = 0; r < nr; r++) {
            cl num(1, 0), den = q[r];
            for (int j = 0; j < K; j++) {
                num *= cl(1, 0) + c1[j] * q[r] + c2[j] * q[r] * q[r];
                if (j != k) den *= cl(1, 0) + a1[j] * q[r] + a2[j] * q[r] * q[r];
            }
            if (!(std::abs(den) > 0)) { *why = "two sections share a pole"; return false; }
            rho[r] = num / den;
        }
        if (nr == 1) { alpha[k] = rho[0].real(); beta[k] = 0; }
        else if (fabsl(q[0].imag()) > 0) {            // complex pair: alpha + beta q = rho
            beta[k] = rho[0].imag() / q[0].imag();
            alpha[k] = rho[0].real() - beta[k] * q[0].real();
        } else {                                      // two real poles
            beta[k] = ((rho[0] - rho[1]) / (q[0] - q[1])).real();
            alpha[k] = (rho[0] - beta[k] * q[0]).real();
        }
        if (!std::isfinite((double)alpha[k]) || !std::isfinite((double)beta[k])) { *why = "partial fractions overflow"; return false; }
    }
    return true;
}

// T with T A_c = A_p T and T B_c = B_p (the state map between two realisations driven by the same
// input), solved in coordinates normalised by the typical state magnitudes.
bool state_map(const Realisation& Rc, const std::vector<ld>& sc, const Realisation& Rp, const std::vector<ld>& sp,
               int S, Mat* T)
{
    const int nu = S * S, ne = S * S + S;
    std::vector<ld> a((size_t)ne * (nu + 1), 0);
    auto Ac = [&](int i, int j) { return Rc.A.a[i][j] * sc[j] / sc[i]; };
    auto Ap = [&](int i, int j) { return Rp.A.a[i][j] * sp[j] / sp[i]; };
    const int w = nu + 1;
    int row = 0;
    for (int i = 0; i < S; i++)
        for (int j = 0; j < S; j++, row++) {
            for (int k = 0; k < S; k++) {
                a[(size_t)row * w + i * S + k] += Ac(k, j);       // (T Ac)[i][j]
                a[(size_t)row * w + k * S + j] -= Ap(i, k);       // (Ap T)[i][j]
            }
        }
    for (int i = 0; i < S; i++, row++) {
        for (int k = 0; k < S; k++) a[(size_t)row * w + i * S + k] = Rc.B[k] / sc[k];
        a[(size_t)row * w + nu] = Rp.B[i] / sp[i];
    }
    // elimination over all ne rows (the system is consistent and has full column rank)
    std::vector<int> prow(nu, -1);
    std::vector<char> used(ne, 0);
    for (int c = 0; c < nu; c++) {
        int piv = -1;
        ld best = 0;
        for (int r = 0; r < ne; r++) if (!used[r] && fabsl(a[(size_t)r * w + c]) > best) { best = fabsl(a[(size_t)r * w + c]); piv = r; }
        if (piv < 0 || !(best > 1e-300L)) return false;
        used[piv] = 1; prow[c] = piv;
        const ld inv = 1 / a[(size_t)piv * w + c];
        for (int j = 0; j < w; j++) a[(size_t)piv * w + j] *= inv;
        for (int r = 0; r < ne; r++) {
            if (r == piv) continue;
            const ld f = a[(size_t)r * w + c];
            if (f == 0) continue;
            for (int j = 0; j < w; j++) a[(size_t)r * w + j] -= f * a[(size_t)piv * w + j];
        }
    }
    for (int r = 0; r < ne; r++) if (!used[r] && fabsl(a[(size_t)r * w + nu]) > 1e-9L) return false;   // inconsistent
    for (int i = 0; i < S; i++)
        for (int j = 0; j < S; j++) T->a[i][j] = a[(size_t)prow[i * S + j] * w + nu] * sp[i] / sc[j];
    return true;
}

// ------------------------------------------------------------------ AP form (iir_rerun.cuh, phase 3)
// one sample of the all-pole cascade: state (v1,v2) per section, returns the last section's w
template <class F>
F ap_step(int K, const F* a1, const F* a2, F* s, F in)
{
    for (int k = 0; k < K; k++) {
        const F w = in - a1[k] * s[2 * k] - a2[k] * s[2 * k + 1];
        s[2 * k + 1] = s[2 * k];
        s[2 * k] = w;
        in = w;
    }
    return in;
}
// Tap with: all-pole cascade started from Tap s (and "all earlier samples zero" in the numerator) continues the
// parallel realisation started from s.  Both zero-input responses are sums of the same pole modes, so matching their
// first S samples matches them for good: O_ap Tap = O_par.  The observability matrices of the reference's narrow
// filters are conditioned like 1e12..1e16 - solved in binary128, the result is good to double precision.
bool ap_state_map(const Form& Fp, int K, double (*Tap)[kMaxS])
{
    typedef __float128 qf;
    const int S = 2 * K;
    qf a1[kMaxK], a2[kMaxK], c1[kMaxK], c2[kMaxK];
    for (int k = 0; k < K; k++) { a1[k] = (qf)Fp.a1[k]; a2[k] = (qf)Fp.a2[k]; c1[k] = (qf)Fp.c1[k]; c2[k] = (qf)Fp.c2[k]; }
    qf M[kMaxS][2 * kMaxS];
    for (int j = 0; j < S; j++) {
        qf sp[kMaxS], sa[kMaxS];
        for (int i = 0; i < S; i++) sp[i] = sa[i] = (i == j) ? (qf)1 : (qf)0;
        for (int n = 0; n < S; n++) {
            // parallel form, zero input: out = sum alpha v1 + beta v2, then the state moves on
            qf o = 0, ns[kMaxS];
            for (int k = 0; k < K; k++) {
                o += c1[k] * sp[2 * k] + c2[k] * sp[2 * k + 1];
                ns[2 * k] = -a1[k] * sp[2 * k] - a2[k] * sp[2 * k + 1];
                ns[2 * k + 1] = sp[2 * k];
            }
            for (int i = 0; i < S; i++) sp[i] = ns[i];
            M[n][S + j] = o;
            M[n][j] = ap_step<qf>(K, a1, a2, sa, (qf)0);
        }
    }
    auto qabs = [](qf v) { return v < 0 ? -v : v; };
    for (int c = 0; c < S; c++) {
        int piv = c;
        for (int r = c + 1; r < S; r++) if (qabs(M[r][c]) > qabs(M[piv][c])) piv = r;
        if (!(qabs(M[piv][c]) > 0)) return false;
        if (piv != c) for (int j = 0; j < 2 * S; j++) std::swap(M[piv][j], M[c][j]);
        const qf inv = (qf)1 / M[c][c];
        for (int j = 0; j < 2 * S; j++) M[c][j] *= inv;
        for (int r = 0; r < S; r++) {
            if (r == c) continue;
            const qf f = M[r][c];
            if (f == 0) continue;
            for (int j = 0; j < 2 * S; j++) M[r][j] -= f * M[c][j];
        }
    }
    for (int i = 0; i < S; i++)
        for (int j = 0; j < S; j++) {
            Tap[i][j] = (double)M[i][S + j];
            if (!std::isfinite(Tap[i][j])) return false;
        }
    return true;
}
}  // namespace

bool make_filter_plan(int K, const double (*sos)[6], int D, bool force_cascade, bool want_y32, FilterPlan* P, const char** msg,
                      int rerun_lt)
{
    static const char* e_k = "n_sos must be 1..4";
    static const char* e_d = "decimation must be one of 1,2,4,8,16,32";
    static const char* e_c = "SOS coefficients must be finite with a0 != 0 and b0 != 0";
    if (K < 1 || K > kMaxK) { *msg = e_k; return false; }
    if (!(D == 1 || D == 2 || D == 4 || D == 8 || D == 16 || D == 32)) { *msg = e_d; return false; }
    const int S = 2 * K;
    P->K = K; P->D = D; P->S = S;
    std::memset(&P->tables, 0, sizeof(P->tables));
    Form Fc;
    Fc.K = K; Fc.par = false;
    P->g[0] = 1.0;
    for (int k = 0; k < K; k++) {
        // Biquad::setCoefficients (Biquad.cpp:123-139): everything divided by a0, in double
        const double a0 = sos[k][3];
        const double b0 = sos[k][0] / a0, b1 = sos[k][1] / a0, b2 = sos[k][2] / a0;
        const double A1 = sos[k][4] / a0, A2 = sos[k][5] / a0;
        if (!std::isfinite(b0) || !std::isfinite(b1) || !std::isfinite(b2) || !std::isfinite(A1) ||
            !std::isfinite(A2) || a0 == 0.0 || b0 == 0.0) { *msg = e_c; return false; }
        Fc.a1[k] = A1; Fc.a2[k] = A2;
        // the kernel uses the rounded (double) c's: the model does too, so that model and kernel agree
        Fc.c1[k] = (double)((ld)b1 / (ld)b0); Fc.c2[k] = (double)((ld)b2 / (ld)b0);
        P->g[k + 1] = (double)((ld)P->g[k] * (ld)b0);
    }
    const Realisation Rc = realise(Fc);
    const std::vector<ld> sc = state_scale(Rc, S);

    // ---- try the parallel realisation
    Form F = Fc;
    Realisation R = Rc;
    std::vector<ld> scale = sc;
    Mat Tp = mat_eye(S);          // kernel <- normalised cascade coordinates
    P->par = false; P->par_err = 0; P->form_note = "";
    if (force_cascade) P->form_note = "cascade form requested";
    else {
        Form Fp = Fc;
        Fp.par = true;
        ld al[kMaxK], be[kMaxK];
        const char* why = "";
        bool ok = partial_fractions(K, Fc.a1, Fc.a2, Fc.c1, Fc.c2, al, be, &why);
        Realisation Rp;
        std::vector<ld> sp;
        Mat T;
        if (ok) {
            for (int k = 0; k < K; k++) { Fp.c1[k] = (double)al[k]; Fp.c2[k] = (double)be[k]; }
            Rp = realise(Fp);
            sp = state_scale(Rp, S);
            ok = state_map(Rc, sc, Rp, sp, S, &T);
            if (!ok) why = "no state map between the realisations (pole-zero cancellation?)";
        }
        if (ok) {
            // both realisations side by side on a deterministic full-scale test signal
            ld s1[kMaxS] = {0}, s2[kMaxS] = {0}, n1[kMaxS], n2[kMaxS], o1, o2, e2 = 0, p2 = 0;
            uint64_t lcg = 0x9E3779B97F4A7C15ull;
            const int N = 1 << 15;
            for (int n = 0; n < N; n++) {
                lcg = lcg * 6364136223846793005ull + 1442695040888963407ull;
                const ld x = (ld)((int)((lcg >> 40) & 8191) - 4096) + 3000 * cosl(0.0078539816339744830962L * n);
                form_step(Fc, s1, x, n1, &o1);
                form_step(Fp, s2, x, n2, &o2);
                for (int i = 0; i < S; i++) { s1[i] = n1[i]; s2[i] = n2[i]; }
                e2 += (o1 - o2) * (o1 - o2); p2 += o1 * o1;
            }
            P->par_err = (double)sqrtl(e2 / (p2 > 0 ? p2 : 1));
            ld worst = 0;
            for (int i = 0; i < S; i++) {
                ld acc = 0;
                for (int j = 0; j < S; j++) acc += T.a[i][j] * s1[j];
                const ld e = fabsl(acc - s2[i]) / sp[i];
                if (e > worst) worst = e;
            }
            if (!(P->par_err < 1e-11)) { ok = false; why = "parallel form loses too much to cancellation"; }
            else if (!(worst < 1e-9L * 8192)) { ok = false; why = "state map inaccurate"; }
        }
        // constant-input steady state must exist (no pole at z = 1) for the mantissa-injection input path
        ld sst[kMaxS];
        if (ok) {
            std::vector<ld> a((size_t)S * (S + 1));
            for (int i = 0; i < S; i++) {
                for (int j = 0; j < S; j++) a[(size_t)i * (S + 1) + j] = (ld)(i == j) - Rp.A.a[i][j];
                a[(size_t)i * (S + 1) + S] = Rp.B[i];
            }
            ok = solve_inplace(a.data(), S, 1);
            for (int i = 0; i < S && ok; i++) {
                sst[i] = a[(size_t)i * (S + 1) + S];
                if (!std::isfinite((double)sst[i]) || fabsl(sst[i]) > 4 * sp[i]) ok = false;
            }
            if (!ok) why = "pole at z = 1";
        }
        if (ok) {
            P->par = true;
            F = Fp; R = Rp; scale = sp; Tp = T;
            ld h = 1;
            for (int i = 0; i < S; i++) { P->sstar[i] = (double)sst[i]; h += Rp.Cu[i] * sst[i]; }
            P->hn1 = (double)h;
        } else {
            P->form_note = why;
        }
    }
    if (!P->par) { for (int i = 0; i < kMaxS; i++) P->sstar[i] = 0; P->hn1 = 1; }
    // the float-staged kernels exist for the preferred realisation only; they afford longer runs
    // (the rerun kernel of full-rate streams stages nothing: its run length comes from the input tile)
    P->rerun = iir_is_rerun(K, D, P->par);
    P->y32 = !P->rerun && want_y32 && iir_has_y32(K, P->par);
    P->LT = P->rerun ? (rerun_lt > 0 ? rerun_lt : iir_rerun_len(K)) : iir_run_len(K, D, P->y32);
    P->NK = P->LT / D;
    for (int k = 0; k < K; k++) {
        P->tables.a1[k] = (double)F.a1[k]; P->tables.a2[k] = (double)F.a2[k];
        P->tables.c1[k] = (double)F.c1[k]; P->tables.c2[k] = (double)F.c2[k];
    }
    P->tables.gain = P->g[K];
    // state maps: reference DF-II state = g[stage] * normalised cascade state
    {
        const int w = 2 * S;
        std::vector<ld> a((size_t)S * w);
        for (int i = 0; i < S; i++)
            for (int j = 0; j < S; j++) { a[(size_t)i * w + j] = Tp.a[i][j]; a[(size_t)i * w + S + j] = (i == j); }
        if (!solve_inplace(a.data(), S, S)) { *msg = "state map is singular"; return false; }
        for (int i = 0; i < S; i++)
            for (int j = 0; j < S; j++) {
                P->Tkr[i][j] = Tp.a[i][j] / (ld)P->g[j / 2];
                P->Trk[i][j] = (ld)P->g[i / 2] * a[(size_t)i * w + S + j];
            }
    }
    // ---- AP form of the rerun kernel's phase 3: zeros all at z = -1, poles away from z = 1
    P->ap = false; P->ap_stages = 0; P->ap_err = 0;
    if (P->rerun && P->par) {
        bool ok = true;
        int stages = 0;
        for (int k = 0; k < K && ok; k++) {
            if (Fc.c1[k] == 2 && Fc.c2[k] == 1 && Fc.a2[k] != 0) stages += 2;
            else ok = false;
        }
        ok = ok && stages == 2 * K && ap_state_map(F, K, P->Tap);
        ld prod = 1;
        for (int k = 0; k < K && ok; k++) {
            const ld d = 1 + F.a1[k] + F.a2[k];
            if (!(fabsl(d) > 0)) { ok = false; break; }
            prod /= d;
            P->sstar_ap[2 * k] = P->sstar_ap[2 * k + 1] = prod;
        }
        if (ok) {
            // the parallel form runs a deterministic full-scale signal, its state is mapped, and both forms go on:
            // the all-pole cascade on the binomially summed samples (earlier samples zero), all in long double
            ld sp[kMaxS] = {0}, np_[kMaxS], op, e2 = 0, p2 = 0, a1l[kMaxK], a2l[kMaxK];
            for (int k = 0; k < K; k++) { a1l[k] = F.a1[k]; a2l[k] = F.a2[k]; }
            uint64_t lcg = 0xD1B54A32D192ED03ull;
            auto sig = [&](int n) {
                lcg = lcg * 6364136223846793005ull + 1442695040888963407ull;
                return (ld)((int)((lcg >> 40) & 8191) - 4096) + 3000 * cosl(0.0078539816339744830962L * n);
            };
            const int N1 = 6000, N2 = 3000;
            for (int n = 0; n < N1; n++) {
                form_step(F, sp, sig(n), np_, &op);
                for (int i = 0; i < S; i++) sp[i] = np_[i];
            }
            ld sa[kMaxS];
            for (int i = 0; i < S; i++) {
                ld acc = 0;
                for (int j = 0; j < S; j++) acc += (ld)P->Tap[i][j] * sp[j];
                sa[i] = acc;
            }
            ld hist[2 * kMaxK + 1] = {0};
            for (int n = 0; n < N2; n++) {
                const ld x = sig(N1 + n);
                form_step(F, sp, x, np_, &op);
                for (int i = 0; i < S; i++) sp[i] = np_[i];
                ld t = x;                               // (1 + q)^stages with zero history: stage i keeps its last input
                for (int i = 0; i < stages; i++) { const ld u = t + hist[i]; hist[i] = t; t = u; }
                const ld oa = ap_step<ld>(K, a1l, a2l, sa, t);
                e2 += (op - oa) * (op - oa); p2 += op * op;
            }
            P->ap_err = (double)sqrtl(e2 / (p2 > 0 ? p2 : 1));
            ok = P->ap_err < 1e-11;
        }
        P->ap = ok;
        P->ap_stages = ok ? stages : 0;
    }
    const Mat& A = R.A;
    const ld* Cu = R.Cu;
    for (int i = 0; i < S; i++) for (int j = 0; j < S; j++) P->A[i][j] = A.a[i][j];
    P->scale = scale;

    // l1_zs = sum_{k < LT} |hn[k]|, hn = impulse response of y / G: bounds a run's zero-state outputs by
    // l1_zs * max|x| (the fixed-point staging of the float-output kernels is sized with it)
    {
        ld s[kMaxS] = {0}, ns[kMaxS], out, acc = 0;
        for (int n = 0; n < P->LT; n++) {
            form_step(F, s, n == 0 ? 1 : 0, ns, &out);
            acc += fabsl(out);
            for (int i = 0; i < S; i++) s[i] = ns[i];
        }
        P->l1_zs = (double)acc;
    }
    // R[m] = Cu A^(mD): walk the row vector through the run
    {
        ld r[kMaxS];
        for (int i = 0; i < S; i++) r[i] = Cu[i];
        for (int n = 0; n < P->LT; n++) {
            if (n % D == 0 && n / D < kMaxNK) for (int i = 0; i < S; i++) P->tables.R[n / D][i] = (double)r[i];
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
    P->scan_rounds = 5;
    {
        Mat p = M;
        for (int j = 0; j < 5; j++) {
            // Mp is stored with the kernel's fixed leading dimension S (row*S + col)
            for (int i = 0; i < S; i++) for (int c = 0; c < S; c++) P->tables.Mp[j][i * S + c] = (double)p.a[i][c];
            // Kogge-Stone round j adds M^(2^j) x (the value 2^j lanes below): once that matrix is below 2^-64 of
            // a typical state, this round and all later ones add nothing representable
            if (P->scan_rounds == 5 && scaled_norm(p, S, scale) < ldexpl(1.0L, -64)) P->scan_rounds = j;
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

void plan_ap_launch(const FilterPlan& P, bool words, double gain, ApLaunch* L)
{
    const int SG = P.ap_stages;
    const double beta = words ? 4096.0 : 32768.0;
    const double tmax = ldexp(2 * beta - 1, SG);                        // largest sum
    int bits = 0;
    frexp(tmax, &bits);                                                 // tmax < 2^bits
    const int j = bits - 20 > 1 ? bits - 20 : 1;
    const int emax = (1 << (12 - j)) - 1;                               // biased exponent that still fits the 32-bit word
    const int E = 1023 + 20 + j < emax ? 1023 + 20 + j : emax;
    const double kappa = ldexp(1.0, E - 1023 - (20 + j));
    const ld Cb = ldexpl(1.0L, 20 + j) + (ld)ldexp(beta, SG);
    L->exp = (uint32_t)E << (20 + j);
    L->mul = 1u << (32 - j);
    L->xorw = words ? 0x10001000u : 0x80008000u;
    for (int i = 0; i < P.S; i++) {
        for (int k = 0; k < P.S; k++) L->T[i][k] = kappa * P.Tap[i][k];
        L->bias[i] = (double)((ld)kappa * Cb * P.sstar_ap[i]);
    }
    L->gain = gain / kappa;
    L->out_bias = (double)(-(ld)gain * Cb * P.sstar_ap[P.S - 2]);
}

void plan_power_kernel(const FilterPlan& P, uint64_t n, double* m)
{
    const int S = P.S;
    Mat A = mat_eye(0);
    for (int i = 0; i < S; i++) for (int j = 0; j < S; j++) A.a[i][j] = P.A[i][j];
    mat_store(mat_pow(A, n, S), S, m);
}

void plan_shard_matrix(const FilterPlan& P, uint64_t n, double* m)
{
    const int S = P.S;
    Mat A = mat_eye(0), Tkr = mat_eye(0), Trk = mat_eye(0);
    for (int i = 0; i < S; i++)
        for (int j = 0; j < S; j++) { A.a[i][j] = P.A[i][j]; Tkr.a[i][j] = P.Tkr[i][j]; Trk.a[i][j] = P.Trk[i][j]; }
    // reference coordinates: s_ref = Trk s_kernel
    mat_store(mat_mul(mat_mul(Trk, mat_pow(A, n, S), S), Tkr, S), S, m);
}

void plan_state_to_kernel(const FilterPlan& P, const double* ref, double* ker)
{
    for (int i = 0; i < P.S; i++) {
        ld acc = 0;
        for (int j = 0; j < P.S; j++) acc += P.Tkr[i][j] * (ld)ref[j];
        ker[i] = (double)acc;
    }
}
void plan_state_to_ref(const FilterPlan& P, const double* ker, double* ref)
{
    for (int i = 0; i < P.S; i++) {
        ld acc = 0;
        for (int j = 0; j < P.S; j++) acc += P.Trk[i][j] * (ld)ker[j];
        ref[i] = (double)acc;
    }
}

}  // namespace cbl
