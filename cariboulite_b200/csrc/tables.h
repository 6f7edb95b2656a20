// Host-side filter design and scan tables for kernel (b).
#pragma once
#include <stdint.h>
#include <vector>
#include "iir.cuh"

namespace cbl {

// Butterworth low-pass as iir1 designs it (LC/iir/iir/Butterworth.cpp:49-68, PoleFilter.cpp:47-90,
// Biquad.cpp:123-201, Cascade.cpp:94-118).  sos rows (b0,b1,b2,a0,a1,a2). Returns stages or -1.
int butter_lowpass_sos(int order, double fs, double fc, double* sos);

// Everything the kernels need for one (cascade, decimation) pair.
struct FilterPlan {
    int K = 0, D = 1, LT = 32, NK = 32, S = 0;
    bool rerun = false;           // full-rate form: the rerun kernel (iir_rerun.cuh), no output stage, no R table
    bool y32 = false;             // float-staged kernels (plain CF32 output, preferred realisation)
    bool par = false;             // kernels run the parallel (partial-fraction) realisation, else the cascade
    const char* form_note = "";   // why the cascade was chosen, when it was
    double par_err = 0;           // parallel vs cascade, rel. rms over a long-double simulation (0 if not tried)
    double g[kMaxK + 1];          // g[k] = prod_{j<k} b0_j ; g[K] = overall gain G
    IirTables tables;             // constant-bank tables (gain left at G, sbias/out_bias at 0: the launch fills them)
    double sstar[kMaxS];          // steady state under a unit constant input, kernel coordinates
    double hn1 = 1;               // H(1) / G
    double l1_zs = 0;             // sum of |impulse response of y / G| over one run: zero-state outputs <= l1_zs * max|x|
    long double Tkr[kMaxS][kMaxS];// kernel state <- reference (DF-II v1,v2) state, per rail
    long double Trk[kMaxS][kMaxS];// reference state <- kernel state
    std::vector<double> plane;    // [S*S][33]  (M^t)[e] at plane[e*33 + t]
    std::vector<double> mtile_pow;// [window+1][S*S]
    int scan_rounds = 5;          // Kogge-Stone rounds of the warp scan that can contribute (M^(2^j) not yet negligible)
    int window = 1;               // look-back window J (tiles)
    bool chained = false;         // no useful contraction: serial inclusive chain
    long double A[kMaxS][kMaxS];  // one-sample transition, kernel coordinates
    uint64_t decay_samples = 0;   // n with |A^n| (scaled) < 2^-60 for all larger n (0 if chained)
    std::vector<long double> scale;  // typical magnitude of each state component
    // AP form of the rerun kernel's phase 3 (iir_rerun.cuh): every section's zeros are a double zero at z = -1 (or one
    // zero there for a first-order section), so y / G = (1 + q)^ap_stages applied to the samples, then an all-pole cascade
    bool ap = false;
    int ap_stages = 0;            // two-tap integer stages = number of zeros at z = -1
    double ap_err = 0;            // AP continuation against the parallel form from a mapped mid-stream state, rel. rms
    double Tap[kMaxS][kMaxS];     // all-pole state (v1,v2 per section; earlier samples taken as zero) <- kernel state
    long double sstar_ap[kMaxS];  // all-pole steady state under a unit constant input; [2K-2] is also its output
};

// false (with msg) if the cascade cannot be handled (b0 == 0, non-finite coefficients, ...)
// want_y32: the output is plain CF32 (float-staged kernels, longer runs) - plan->y32 says whether they will be used
// rerun_lt > 0: run length of a rerun plan (the library keeps a second, short-run plan for small submits)
bool make_filter_plan(int K, const double (*sos)[6], int D, bool force_cascade, bool want_y32, FilterPlan* plan, const char** msg,
                      int rerun_lt = 0);

// Launch constants of the AP form for one input kind (SMI words: 13-bit fields, offset 4096; CS16: offset 32768) and
// output gain.  The binomial sum t = sum_i C(stages, i) u[n-i] of the offset-binary samples u = x + beta is dropped into
// a double by  X = bits((t + exp) * mul)  = kappa (2^(20+j) + t) = kappa (Cb + N(q) x),  kappa a power of two.
struct ApLaunch {
    uint32_t exp, mul, xorw;      // exponent bits added in the last sum, 2^(32-j), offset-binary flip of both halves of a word
    double gain, out_bias;        // y = fma(w_K, gain, out_bias)
    double T[kMaxS][kMaxS];       // kappa * Tap
    double bias[kMaxS];           // all-pole state under the constant kappa * Cb
};
void plan_ap_launch(const FilterPlan& plan, bool words, double gain, ApLaunch* out);

// A^n in REFERENCE state coordinates (row-major S x S doubles)
void plan_shard_matrix(const FilterPlan& plan, uint64_t n, double* m);
// A^n in KERNEL state coordinates
void plan_power_kernel(const FilterPlan& plan, uint64_t n, double* m);
// per-rail state vectors (S doubles) between the two coordinate systems
void plan_state_to_kernel(const FilterPlan& plan, const double* ref, double* ker);
void plan_state_to_ref(const FilterPlan& plan, const double* ker, double* ref);

}  // namespace cbl
