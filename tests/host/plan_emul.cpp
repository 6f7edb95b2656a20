// Host-side emulation of kernel (b)'s arithmetic from the FilterPlan tables (test helper, CPU only).
// Built by tests/test_host_plan.py together with cariboulite_b200/csrc/tables.cpp; checks that the
// tables (realisation, bias steady state, R, M powers, state maps) reproduce the sequential filter.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include "../../cariboulite_b200/csrc/tables.h"

using namespace cbl;

static void step(const FilterPlan& P, double* v, double x, double* out)
{
    const IirTables& T = P.tables;
    if (P.par) {
        double o = x;
        for (int k = 0; k < P.K; k++) {
            o = fma(T.c2[k], v[2 * k + 1], fma(T.c1[k], v[2 * k], o));
            const double w = fma(-T.a1[k], v[2 * k], fma(-T.a2[k], v[2 * k + 1], x));
            v[2 * k + 1] = v[2 * k]; v[2 * k] = w;
        }
        *out = o;
    } else {
        double in = x;
        for (int k = 0; k < P.K; k++) {
            const double t = fma(-T.a2[k], v[2 * k + 1], in);
            const double w = fma(-T.a1[k], v[2 * k], t);
            in = fma(T.c2[k], v[2 * k + 1], fma(T.c1[k], v[2 * k], w));
            v[2 * k + 1] = v[2 * k]; v[2 * k] = w;
        }
        *out = in;
    }
}

extern "C" {

// info[0]=par, [1]=window, [2]=chained, [3]=decay, [4]=par_err, [5]=hn1, [6]=run length, [7]=kept per run, [8]=y32, [9]=rerun
int emul_plan_info(int K, const double* sos, int D, int force_cascade, int want_y32, double* info)
{
    FilterPlan P; const char* msg = "";
    if (!make_filter_plan(K, (const double(*)[6])sos, D, force_cascade != 0, want_y32 != 0, &P, &msg)) return -1;
    info[0] = P.par; info[1] = P.window; info[2] = P.chained; info[3] = (double)P.decay_samples;
    info[4] = P.par_err; info[5] = P.hn1; info[6] = P.LT; info[7] = P.NK; info[8] = P.y32; info[9] = P.rerun;
    return 0;
}

// One rail: x[n] integer-valued samples, bias as the kernel would add it (0 = converted input).
// out[n/D] = gain * y at kept samples; state_ref_in/out in reference (DF-II) coordinates.
int emul_run(int K, const double* sos, int D, int force_cascade, int want_y32, double bias, double out_scale,
             const double* x, size_t n, const double* state_ref_in, double* out, double* state_ref_out)
{
    FilterPlan P; const char* msg = "";
    if (!make_filter_plan(K, (const double(*)[6])sos, D, force_cascade != 0, want_y32 != 0, &P, &msg)) return -1;
    if (!P.par) bias = 0;
    const int S = P.S, LT = P.LT, NK = P.NK;
    if (n % LT) return -2;
    const double gain = P.g[K] * out_scale;
    double sb[kMaxS], s0[kMaxS];
    for (int i = 0; i < S; i++) sb[i] = bias * P.sstar[i];
    const double out_bias = -bias * P.hn1 * gain;
    plan_state_to_kernel(P, state_ref_in, s0);
    std::vector<double> ys(NK);
    for (size_t r = 0; r < n / LT; r++) {
        double v[kMaxS];
        for (int i = 0; i < S; i++) v[i] = sb[i];
        for (int t = 0; t < LT; t++) {
            double y;
            step(P, v, x[r * LT + t] + bias, &y);
            if (t % D == 0) ys[t / D] = y;
        }
        if (P.rerun) {
            // iir_rerun.cuh: the run again from its true entry state (offset by the steady state of the bias)
            double w[kMaxS];
            for (int i = 0; i < S; i++) w[i] = sb[i] + s0[i];
            for (int t = 0; t < LT; t++) {
                double y;
                step(P, w, x[r * LT + t] + bias, &y);
                if (t % D == 0) out[r * NK + t / D] = fma(y, gain, out_bias);
            }
        } else
        for (int m = 0; m < NK; m++) {
            double y = ys[m];
            for (int c = 0; c < S; c++) y = fma(P.tables.R[m][c], s0[c], y);
            out[r * NK + m] = fma(y, gain, out_bias);
        }
        // s0 <- M s0 + f
        double ns[kMaxS];
        for (int i = 0; i < S; i++) {
            double acc = v[i] - sb[i];
            for (int c = 0; c < S; c++) acc = fma(P.plane[(size_t)(i * S + c) * 33 + 1], s0[c], acc);
            ns[i] = acc;
        }
        for (int i = 0; i < S; i++) s0[i] = ns[i];
    }
    plan_state_to_ref(P, s0, state_ref_out);
    return P.par ? 1 : 0;
}

// The AP form of the rerun kernel's phase 3 (iir_rerun.cuh), one rail of integer samples: every run starts from the
// sequentially carried true state (kernel coordinates), mapped to the all-pole state; the numerator is the integer
// two-tap cascade with "earlier samples zero", the sum enters the double through the same 32-bit word and multiply
// the kernel uses.  info[0] = ap, [1] = stages, [2] = ap_err.  Returns 1 if the plan takes the AP form, 0 if not.
int emul_run_ap(int K, const double* sos, int D, int words, double out_scale, const int32_t* x, size_t n,
                const double* state_ref_in, double* out, double* state_ref_out, double* info)
{
    FilterPlan P; const char* msg = "";
    if (!make_filter_plan(K, (const double(*)[6])sos, D, false, false, &P, &msg)) return -1;
    info[0] = P.ap; info[1] = P.ap_stages; info[2] = P.ap_err;
    if (!P.rerun || !P.ap) return 0;
    const int S = P.S, LT = P.LT, NK = P.NK, SG = P.ap_stages;
    if (n % LT) return -2;
    ApLaunch L;
    plan_ap_launch(P, words != 0, P.g[K] * out_scale, &L);
    const uint32_t beta = words ? 4096u : 32768u;
    double s0[kMaxS];
    plan_state_to_kernel(P, state_ref_in, s0);
    for (size_t r = 0; r < n / LT; r++) {
        double sg[kMaxS];
        for (int i = 0; i < S; i++) {
            double acc = L.bias[i];
            for (int c = 0; c < S; c++) acc = fma(L.T[i][c], s0[c], acc);
            sg[i] = acc;
        }
        uint32_t hist[2 * kMaxK];
        for (int i = 0; i < SG; i++) hist[i] = beta << i;
        double v[kMaxS];
        for (int i = 0; i < S; i++) v[i] = s0[i];
        for (int t = 0; t < LT; t++) {
            const int32_t xi = x[r * LT + t];
            uint32_t u = (uint32_t)(xi + (int32_t)beta);
            for (int i = 0; i < SG; i++) { const uint32_t s = u + hist[i]; hist[i] = u; u = s; }
            const unsigned long long bits = (unsigned long long)(u + L.exp) * L.mul;
            double in;
            memcpy(&in, &bits, 8);
            for (int k = 0; k < K; k++) {
                const double w = fma(-P.tables.a1[k], sg[2 * k], fma(-P.tables.a2[k], sg[2 * k + 1], in));
                sg[2 * k + 1] = sg[2 * k]; sg[2 * k] = w; in = w;
            }
            if (t % D == 0) out[r * NK + t / D] = fma(in, L.gain, L.out_bias);
            double y;
            step(P, v, (double)xi, &y);          // the true state carried on (the kernel gets it from the scan)
        }
        for (int i = 0; i < S; i++) s0[i] = v[i];
    }
    plan_state_to_ref(P, s0, state_ref_out);
    return 1;
}

// max |Mp[j] - M^(2^j)| relative to the largest entry of that power, and sparsity check of the pattern the kernel assumes
double emul_check_powers(int K, const double* sos, int D, int force_cascade, int want_y32)
{
    FilterPlan P; const char* msg = "";
    if (!make_filter_plan(K, (const double(*)[6])sos, D, force_cascade != 0, want_y32 != 0, &P, &msg)) return -1;
    const int S = P.S;
    double worst = 0;
    for (int j = 0; j < 5; j++) {
        double diff = 0, mx = 0;
        for (int i = 0; i < S; i++)
            for (int c = 0; c < S; c++) {
                const double a = P.tables.Mp[j][i * S + c], b = P.plane[(size_t)(i * S + c) * 33 + (1 << j)];
                diff = fmax(diff, fabs(a - b));
                mx = fmax(mx, fmax(fabs(a), fabs(b)));
                const bool in_pattern = P.par ? (c / 2 == i / 2) : (c < 2 * (i / 2 + 1));
                if (!in_pattern && (a != 0.0 || b != 0.0)) return 1e300;
            }
        if (mx > 0) worst = fmax(worst, diff / mx);
    }
    return worst;
}
}
