// oracle/oracle_iir.cpp - floating-point half of the CPU oracle (see oracle.h header note).
// TEST INFRASTRUCTURE ONLY - never part of the product path.
//
// Restates the part of the vendored iir1 library (berndporr/iir1 v1.9.1, LC/iir/) that the
// reference instantiates: Iir::Butterworth::LowPass<N, DirectFormII>.
#include "oracle.h"
#include <cmath>
#include <complex>
#include <limits>
#include <new>
#include <vector>

typedef std::complex<double> cplx;
static const double kPi   = 3.1415926535897932384626433832795028841971;   // MathSupplement.h:51
static const double kPi_2 = 1.5707963267948966192313216916397514420986;   // MathSupplement.h:52

namespace {

struct Sos { double b0, b1, b2, a0, a1, a2; };

// Biquad::setCoefficients, Biquad.cpp:123-139
Sos set_coefficients(double a0, double a1, double a2, double b0, double b1, double b2)
{
    Sos s;
    s.a0 = a0; s.a1 = a1 / a0; s.a2 = a2 / a0;
    s.b0 = b0 / a0; s.b1 = b1 / a0; s.b2 = b2 / a0;
    return s;
}

// Biquad::setTwoPole, Biquad.cpp:156-201
Sos set_two_pole(cplx pole1, cplx zero1, cplx pole2, cplx zero2)
{
    double a1, a2, b1, b2;
    if (pole1.imag() != 0) { a1 = -2 * pole1.real(); a2 = std::norm(pole1); }
    else                   { a1 = -(pole1.real() + pole2.real()); a2 = pole1.real() * pole2.real(); }
    if (zero1.imag() != 0) { b1 = -2 * zero1.real(); b2 = std::norm(zero1); }
    else                   { b1 = -(zero1.real() + zero2.real()); b2 = zero1.real() * zero2.real(); }
    return set_coefficients(1, a1, a2, 1, b1, b2);
}

// Biquad::setOnePole, Biquad.cpp:141-154
Sos set_one_pole(cplx pole, cplx zero)
{
    return set_coefficients(1, -pole.real(), 0, -zero.real(), 1, 0);
}

// MathSupplement.h:85-91
cplx addmul(const cplx& c, double v, const cplx& c1)
{
    return cplx(c.real() + v * c1.real(), c.imag() + v * c1.imag());
}

// Cascade::response, Cascade.cpp:55-77
cplx cascade_response(const std::vector<Sos>& st, double nf)
{
    double w = 2 * kPi * nf;
    const cplx czn1 = std::polar(1., -w);
    const cplx czn2 = std::polar(1., -2 * w);
    cplx ch(1), cbot(1);
    for (size_t i = 0; i < st.size(); i++) {
        const Sos& s = st[i];
        cplx cb(1);
        cplx ct(s.b0 / s.a0);
        ct = addmul(ct, s.b1 / s.a0, czn1);
        ct = addmul(ct, s.b2 / s.a0, czn2);
        cb = addmul(cb, s.a1 / s.a0, czn1);
        cb = addmul(cb, s.a2 / s.a0, czn2);
        ch *= ct;
        cbot *= cb;
    }
    return ch / cbot;
}

}  // namespace

extern "C" int orc_butter_lowpass_sos(int order, double fs, double fc_hz, double* sos)
{
    if (order < 1 || order > 16) return -1;
    const double fc = fc_hz / fs;                        // Butterworth.h:146-158
    if (!(fc < 0.5) || fc < 0.0) return -1;              // PoleFilter.cpp:62-63
    const cplx inf(std::numeric_limits<double>::infinity());
    // AnalogLowPass::design, Butterworth.cpp:49-68
    const double n2 = 2 * order;
    const int pairs = order / 2;
    std::vector<cplx> apoles;
    for (int i = 0; i < pairs; ++i) apoles.push_back(std::polar(1., kPi_2 + (2 * i + 1) * kPi / n2));
    // LowPassTransform, PoleFilter.cpp:47-90
    const double f = std::tan(kPi * fc);
    auto transform = [&](cplx c) -> cplx {
        if (c == inf) return cplx(-1, 0);
        c = f * c;
        return (1. + c) / (1. - c);
    };
    std::vector<Sos> st;
    for (int i = 0; i < pairs; ++i) {
        cplx p = transform(apoles[i]);
        cplx z = transform(inf);
        // addPoleZeroConjugatePairs (Layout.h:108-124) then setPoleZeroPair (Biquad.h:135-142)
        st.push_back(set_two_pole(p, z, std::conj(p), std::conj(z)));
    }
    if (order & 1) {
        cplx p = transform(cplx(-1, 0));
        cplx z = transform(inf);
        st.push_back(set_one_pole(p, z));
    }
    // Cascade::setLayout, Cascade.cpp:94-118: normalise DC gain by scaling stage 0's b only
    const double scale = 1.0 / std::abs(cascade_response(st, 0.0 / (2 * kPi)));
    st[0].b0 *= scale; st[0].b1 *= scale; st[0].b2 *= scale;       // Biquad::applyScale
    for (size_t i = 0; i < st.size(); i++) {
        sos[6 * i + 0] = st[i].b0; sos[6 * i + 1] = st[i].b1; sos[6 * i + 2] = st[i].b2;
        sos[6 * i + 3] = st[i].a0; sos[6 * i + 4] = st[i].a1; sos[6 * i + 5] = st[i].a2;
    }
    return (int)st.size();
}

// ------------------------------------------------------------------ run time
struct orc_cascade {
    int k;
    std::vector<Sos> st;
    std::vector<double> v1, v2;    // DirectFormII state, State.h:105-114
    // CascadeStages::filter, Cascade.h:161-169 + DirectFormII::filter, State.h:116-126
    inline double step(double in)
    {
        double out = in;
        for (int s = 0; s < k; s++) {
            const Sos& c = st[s];
            const double w = out - c.a1 * v1[s] - c.a2 * v2[s];
            const double o = c.b0 * w + c.b1 * v1[s] + c.b2 * v2[s];
            v2[s] = v1[s];
            v1[s] = w;
            out = o;
        }
        return out;
    }
};

extern "C" {

orc_cascade* orc_cascade_create(int nstages, const double* sos)
{
    if (nstages < 1) return nullptr;
    orc_cascade* c = new (std::nothrow) orc_cascade();
    if (!c) return nullptr;
    c->k = nstages;
    for (int i = 0; i < nstages; i++)   // CascadeStages::setup -> Biquad::setCoefficients(a0,a1,a2,b0,b1,b2)
        c->st.push_back(set_coefficients(sos[6 * i + 3], sos[6 * i + 4], sos[6 * i + 5],
                                         sos[6 * i + 0], sos[6 * i + 1], sos[6 * i + 2]));
    c->v1.assign(nstages, 0.0);
    c->v2.assign(nstages, 0.0);
    return c;
}
void orc_cascade_destroy(orc_cascade* c) { delete c; }
void orc_cascade_reset(orc_cascade* c) { c->v1.assign(c->k, 0.0); c->v2.assign(c->k, 0.0); }
void orc_cascade_get_state(const orc_cascade* c, double* v)
{
    for (int s = 0; s < c->k; s++) { v[2 * s] = c->v1[s]; v[2 * s + 1] = c->v2[s]; }
}
void orc_cascade_set_state(orc_cascade* c, const double* v)
{
    for (int s = 0; s < c->k; s++) { c->v1[s] = v[2 * s]; c->v2[s] = v[2 * s + 1]; }
}
void orc_cascade_run_f32(orc_cascade* c, const float* in, float* out, size_t n)
{
    for (size_t i = 0; i < n; i++) out[i] = static_cast<float>(c->step(in[i]));
}
void orc_cascade_run_f64(orc_cascade* c, const double* in, double* out, size_t n)
{
    for (size_t i = 0; i < n; i++) out[i] = c->step(in[i]);
}
void orc_cascade_run_cs16_inplace(orc_cascade* ci, orc_cascade* cq, int16_t* iq, size_t n)
{
    for (size_t k = 0; k < n; k++) {   // CaribouliteStream.cpp:291-298
        iq[2 * k + 0] = (int16_t)static_cast<float>(ci->step((float)iq[2 * k + 0]));
        iq[2 * k + 1] = (int16_t)static_cast<float>(cq->step((float)iq[2 * k + 1]));
    }
}


// Whole RX path of one in-memory capture (composition of the functions above, same order as
// oracle/ref_shim/ref_glue.cpp::ref_rx_path_f32): MTU-sized orc_smi_read requests -> CS16 ->
// (float)v / 4096.0f -> cascade per rail -> keep every D-th output.  Returns filtered samples or < 0.
long orc_rx_path_f32(const uint8_t* bytes, size_t nbytes, size_t native_batch_len, int channel,
                     orc_cascade* ci, orc_cascade* cq, int D, float* out, size_t n_samples)
{
    const size_t mtu = native_batch_len / ORC_BYTES_PER_SAMPLE;
    std::vector<int16_t> cs16((mtu + 1) * 2);
    std::vector<uint8_t> meta(mtu + 1);
    size_t cursor = 0, pos = 0, m = 0;
    while (pos < n_samples) {
        const size_t want = n_samples - pos < mtu ? n_samples - pos : mtu;
        const long got = orc_smi_read(bytes, nbytes, &cursor, native_batch_len, channel, cs16.data(), meta.data(), want);
        if (got < 0) return got;
        if (got == 0) break;
        for (long k = 0; k < got; k++) {
            const float yi = static_cast<float>(ci->step((float)cs16[2 * k + 0] / 4096.0f));
            const float yq = static_cast<float>(cq->step((float)cs16[2 * k + 1] / 4096.0f));
            if ((pos + (size_t)k) % (size_t)D == 0) { out[2 * m] = yi; out[2 * m + 1] = yq; m++; }
        }
        pos += (size_t)got;
    }
    return (long)m;
}

}  // extern "C"
