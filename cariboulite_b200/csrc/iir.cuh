// Kernel (b): cascaded-biquad IIR as a block-parallel linear-recurrence scan with the
// decimation fused into the store - shared declarations.
//
// Math (per rail; the two rails I and Q share coefficients, never state):
//   reference stage k (LC/iir/iir/State.h:116-126):  w = in - a1 v1 - a2 v2 ; out = b0 w + b1 v1 + b2 v2
//   Two realisations of the same transfer function H(z) = G * prod_k (1 + c1 q + c2 q^2) / (1 + a1 q + a2 q^2),
//   q = z^-1, c = b/b0, G = prod b0 (tables.cpp picks one per filter):
//   CASCADE coordinates: every stage's state is divided by g_k = prod_{j<k} b0_j, so that
//       w' = in' - a1 v1' - a2 v2' ;  in'_{k+1} = w' + c1 v1' + c2 v2'
//     and y = G * in'_K.  The reference's arithmetic order minus one multiply per inner stage.
//   PARALLEL coordinates (partial fractions, needs distinct poles): all sections see the input itself,
//       w_k = x - a1_k v1_k - a2_k v2_k ;  y = G * (x + sum_k alpha_k v1_k + beta_k v2_k)
//     2K FMAs per sample on the recurrence and 2K more only where an output is kept; the K sections are
//     independent chains.  State converts to/from the reference's (v1,v2) by a constant S x S matrix.
//   Either way the 2K-vector s obeys s[n+1] = A s[n] + B x[n]; y/G = Cu s[n] + x[n].
//
// Input conversion without conversion instructions (integer inputs, PARALLEL form): the 13-bit (16-bit)
// two's-complement field is XOR-ed into offset binary and dropped into the mantissa of a double with a
// fixed exponent, which IS the double x + 2^20 + 2^12 (x + 2^21 + 2^15) - one LOP3 and one IMAD.WIDE
// (plus a shift for the field in the upper half-word), no I2F on the quarter-rate pipe.  A constant input offset c only shifts the whole state trajectory by the steady state
// s* = (I - A)^-1 B c when the run starts AT s*, and the output by c * H(1): runs start at s*, s* is
// subtracted from the run end state and c * H(1) * gain rides in the final multiply-add.
//
// Parallelisation: a *run* is LT consecutive samples owned by one thread; a *tile* is the 32
// consecutive runs of one warp.  Every thread filters its run from ZERO state (exactly the
// reference's arithmetic order), staging in'_K at the kept (decimated) positions.  Run end
// states are combined across the warp with shuffles (Kogge-Stone with the constant matrices
// M^(2^j), M = A^LT); tiles are chained through global memory with a decoupled look-back whose
// window is bounded by the filter's contraction (M_tile^J ~ 0).  Finally each kept output is
// corrected with y[m] = G * (in'_K,zs[m] + R[m] . s0) and stored - decimation costs no extra pass.
#pragma once
#include <cuda.h>
#include "common.cuh"

namespace cbl {

constexpr int kMaxK  = 4;
constexpr int kMaxS  = 2 * kMaxK;      // state dimension per rail
constexpr int kMaxNK = 32;             // kept outputs per run
constexpr int kMaxDevices = 64;        // device ordinals the per-device launch set-up keeps a slot for

enum { IN_WORDS = 0, IN_CS16 = 1, IN_CF32 = 2,
       IN_WORDS_HIF = 3 };   // internal: SMI words with the HiF field order (I = lo13), its own kernel instantiation
// epilogue modes
enum { OUT_F32 = 0,          // (float)y                              (parity F32)
       OUT_Q16_CS16 = 1,     // (int16)(float)y                       (parity Q16, CS16)
       OUT_Q16_CF32 = 2,     // (float)q16 / 4096                     (parity Q16, CF32)
       OUT_Q16_CF64 = 3,     // (double)q16 / 4096
       OUT_Q16_CS8 = 4,      // (int8)((q16 >> 5) & 0xFF)
       OUT_F64 = 5 };        // y as double (parity F32, CF64)

// Filter constants that sit in the kernel-parameter constant bank (operands of DFMA).
struct IirTables {
    double a1[kMaxK], a2[kMaxK];          // recurrence, both forms
    double c1[kMaxK], c2[kMaxK];          // cascade: b1/b0, b2/b0.  parallel: alpha_k, beta_k
    double gain;                          // G * output scale
    double out_bias;                      // added after the gain: -c * Hn(1) * gain (0 when the input carries no offset)
    double sbias[kMaxS];                  // s*: state every run starts from (0 when the input carries no offset)
    double Mp[5][kMaxS * kMaxS];          // M^(2^j), row-major S x S (S = 2K)
    double R[kMaxNK][kMaxS];              // R[m] = Cu A^(m D).  The rerun kernel has no use for it; its AP form keeps
                                          // Tap there (rows 0..S-1) and the all-pole bias state (row kApBiasRow)
    // fixed-point staging of the float-output kernels fed with integers: phase 1 stores the low word of
    // gain * y_zs + stage_mo (stage_mo = 1.5 * 2^(52-F) + 2^(31-F), folded into out_bias), phase 3 rebuilds the
    // double from that word and stage_hi and subtracts stage_mo again - no FP64<->FP32 conversion on the way
    double stage_mo;
    uint32_t stage_hi, stage_pad;
};

constexpr int kApBiasRow = kMaxS;         // IirTables::R row of the AP form's bias state

struct IirParams {
    uint32_t n_streams;
    uint32_t n_runs;            // valid runs per stream
    uint32_t tiles_per_stream;  // tiles processed per stream in this launch
    uint32_t total_tiles;       // n_streams * tiles_per_stream
    uint32_t last_tile_runs;    // valid runs in a stream's last tile (1..32); only meaningful when the
                                // launch covers the whole stream (write_state != 0)
    int      window;            // J: look-back window in tiles (aggregates only)
    int      chained;           // 1: no contraction - wait for the predecessor's inclusive state
    int      write_state;       // last tile of each stream writes the stream end state
    int      layout;            // IN_WORDS: 0 = S1G (I = hi13), 1 = HiF (I = lo13).  The kernel filters rail A = hi13
                                // and rail B = lo13 and swaps at its borders (state in/out, output store)
    int      out_mode;
    const double* mtile_pow;    // [window+1][S*S]  M_tile^j (row-major)
    const double* plane;        // [S*S][33]        plane[e*33 + t] = (M^t)[e]
    double*   desc_agg;         // [total_tiles][2S]  zero-state tile end states; every double validates itself:
    double*   desc_incl;        // [total_tiles][2S]  true tile end states         the arrays are pre-set to kDescEmpty
    uint32_t* ticket;           // word 0 zeroed per launch; words 32..63 are zeros the kernel reads (iir_kernel.cuh take_ticket)
    const double* state_in;     // [n_streams][2S] kernel coordinates
    double*   state_out;        // [n_streams][2S]
    void*     out;              // kept outputs, NK per run, streams out_stride elements apart
    uint64_t  out_stride;
    const int32_t* gate;        // optional: tiles of stream s are skipped when gate[s * gate_stride] != 0 (a capture that
    uint32_t  gate_stride;      // lost word alignment is redone on the slow path); stride 0 = one flag for the launch
    // mantissa injection (iir_kernel.cuh inject_f64): sign-flip | exponent << 1 word, and the multiplier 2^31
    uint32_t  inj_xor, inj_mul;
    int       scan_rounds;      // Kogge-Stone rounds of the warp scan that can still contribute (<= 5)
    int       out_vec;          // CF32 output may be stored 32 bytes at a time (base and stream stride 32-byte aligned)
    // rerun kernel, AP form of phase 3 (iir_rerun.cuh): exponent bits added in the last two-tap sum, the power of two
    // that spreads the sum over the double's register pair, the offset-binary flip of both halves of an input word
    int       ap;
    uint32_t  ap_exp, ap_mul, ap_xor;
    uint32_t  rr_boxes;         // rerun kernel: 4 KiB boxes per tile = run length / samples per 128-byte row piece (runtime:
                                // small submits take a plan with short runs so that they still spread over the GPU)
};

struct IirTailParams {
    uint32_t n_streams;
    const void* in; uint64_t in_stride;     // elements
    uint64_t first;  uint64_t n;            // samples [first, n) of every stream
    void* out; uint64_t out_stride;
    int in_kind, layout, out_mode, K, D, par;
    double a1[kMaxK], a2[kMaxK], c1[kMaxK], c2[kMaxK];
    double gain;
    double* state;                          // [n_streams][2S] kernel coordinates, updated in place ...
    const double* state_src;                // ... or, when set, read from here and written to `state` (the commit of a submit)
    // optional: the context's control block (device) copied into its page-locked host mirror by this launch, the last
    // one of a submit - instead of a D2H copy operation on the stream
    const uint32_t* mirror_src; uint32_t* mirror_dst; uint32_t mirror_words;
    const int32_t* gate;                    // optional: stream s is skipped when gate[s * gate_stride] != 0
    uint32_t gate_stride;
};

// Run length LT (samples one thread filters sequentially) for a cascade of K sections and a decimation factor D.
// Longer runs mean less per-tile work (scan, look-back, ticket) per sample, but a warp stages two tiles' kept
// outputs (LT / D slots each) in shared memory, and the CTA wants all the warps the register file allows
// (iir_kernel.cuh: 12 for the float-staged kernels of K <= 2, else 8): LT is the longest multiple of 32 whose
// stage still lets that many warps fit, at most kMaxNK kept outputs per run (profiles/r1_sweeps.txt, sweeps 17-19).
#ifndef CBL_IIR_LT_D4
#define CBL_IIR_LT_D4 128        // measured: 128 at 9 warps beats the formula's 64 at 12 (sweep 19)
#endif
#ifndef CBL_IIR_K3_WARPS
#define CBL_IIR_K3_WARPS 12      // float-staged parallel-form kernels of 3 sections: 8 (255 registers) or 12 (168, small spills)
#endif
// Look-back and phase 3 of a tile run one tile late, after phases 1-2 of the warp's next tile, out of the other half of
// a double-buffered output stage (CBL_IIR_DEFER = 1, the default).  The alternative (0) runs them at once: one
// stage per warp, so runs of up to kMaxNK kept outputs fit (256 samples at D = 8 instead of 160), tickets taken one
// at a time.  Measured (profiles/r2_sweeps.txt, sweep 31): wide filters at D <= 4 and from three sections on gain
// 3-7 %, the C4 kernel loses 3 % (1.092 against 1.056 ms) and filters whose look-back reads two tiles lose 8-25 % -
// the wait for a neighbour's aggregate is real time there.  With the two-ahead ticket pipeline of the deferred form
// it is far worse (1.52 ms): a warp then holds two consecutive tiles and processes them an iteration apart.
#ifndef CBL_IIR_DEFER
#define CBL_IIR_DEFER 1
#endif
constexpr bool kIirDefer = CBL_IIR_DEFER != 0;
constexpr int kIirStageBufs = kIirDefer ? 2 : 1;
__host__ __device__ constexpr int iir_target_warps(int K, bool y32) { return (K <= 2 && y32) ? 12 : ((K == 3 && y32) ? CBL_IIR_K3_WARPS : 8); }
__host__ __device__ constexpr int iir_run_len(int K, int D, bool y32)
{
    if (D == 1) return 32;
    if (D == 2) return 64;
    if (D == 4 && CBL_IIR_LT_D4 != 0) return CBL_IIR_LT_D4;
    const int slot = y32 ? 256 : 512;                                    // bytes per kept output of a warp
    const int per_warp = (232448 - 1024) / iir_target_warps(K, y32) - 2 * 4096 - 2 * 2 * kMaxK * 8 - 16;   // minus ring, scratch, barriers
    int nk = per_warp / (kIirStageBufs * slot);
    if (nk > kMaxNK) nk = kMaxNK;
    nk &= ~3;                                                             // phase 3 works in chunks of 4 kept outputs
    int lt = (nk * D / 32) * 32;
    return lt < 32 ? 32 : lt;
}
// Full-rate and lightly decimated streams of the parallel realisation take the rerun kernel (iir_rerun.cuh): the
// input tile stays in shared memory and is filtered a second time from the true entry state, so the run length is
// set by the input (two tiles of LT * 128 bytes per warp), not by a stage of kept outputs.
// Which streams: every full-rate one, and decimate-by-2 from three sections on (sweeps 22 and 25: at D = 2 the
// staged kernel still wins for K <= 2, at D >= 4 everywhere).
#ifndef CBL_RR_ENABLE
#define CBL_RR_ENABLE 1          // 0: staged kernel everywhere (A/B builds)
#endif
#ifndef CBL_RR_LT
#define CBL_RR_LT 256          // samples per run of the rerun kernel from three sections on (the input is fetched
#endif                         // twice, not kept in shared memory: iir_rerun.cuh)
#ifndef CBL_RR_LT12
#define CBL_RR_LT12 128        // ... for one or two sections: these are close to the HBM bound, and the second fetch
#endif                         // of a tile hits L2 more often when less data is in flight (sweep: profiles/r2_sweeps.txt)
#ifndef CBL_RR_WARPS
#define CBL_RR_WARPS 12
#endif
#ifndef CBL_RR_WARPS3
#define CBL_RR_WARPS3 12
#endif
#ifndef CBL_RR_WARPS4
#define CBL_RR_WARPS4 12
#endif
#ifndef CBL_RR_P3U
#define CBL_RR_P3U 1             // chunks of four outputs per iteration of the rerun kernel's phase-3 loop
#endif
#ifndef CBL_RR_RESIDENT
#define CBL_RR_RESIDENT 1        // the boxes of a tile still in the ring after phase 1 are not fetched a second time
#endif
#ifndef CBL_RR_RING
#define CBL_RR_RING 4            // 4 KiB boxes in a warp's ring
#endif
#ifndef CBL_RR_AP_WARPS
#define CBL_RR_AP_WARPS 16       // warps per SM of the AP form up to three sections (its serial section chain wants more
#endif                           // warps to hide behind; 128 registers each), with a ring of three boxes so that they fit
__host__ __device__ constexpr bool rerun_wide(int K, bool ap) { return ap && K <= 3 && CBL_RR_AP_WARPS > 12; }
__host__ __device__ constexpr int rerun_ring(int K, bool ap) { return rerun_wide(K, ap) ? 3 : CBL_RR_RING; }
// boxes of a tile of `boxes` boxes that stay in a ring of `ring` slots between the two passes (iir_rerun.cuh)
__host__ __device__ constexpr int rerun_keep(int boxes, int ring, bool ap)
{
    return CBL_RR_RESIDENT == 0 ? 0 : (boxes <= ring ? boxes : (ap ? 0 : ring));
}
constexpr int kRrPowRow = 16;             // IirTables::R rows kRrPowRow.. hold A^(samples of phase 1's first part), S x S
__host__ __device__ constexpr int iir_rerun_len(int K) { return K <= 2 ? CBL_RR_LT12 : CBL_RR_LT; }
constexpr int kRerunSmallLT = 64;       // run length of the second plan a context keeps for submits too small to fill the GPU
constexpr int kRerunP3Unroll = CBL_RR_P3U;
__host__ __device__ constexpr bool iir_is_rerun(int K, int D, bool par)
{
    return CBL_RR_ENABLE != 0 && par && (D == 1 || (D == 2 && K >= 3));
}
#ifndef CBL_IIR_STAGES
#define CBL_IIR_STAGES 2
#endif
#ifndef CBL_IIR_STAGES32
#define CBL_IIR_STAGES32 2
#endif
constexpr int kIirRingStages   = CBL_IIR_STAGES;      // TMA ring depth per warp, double-staged kernels
constexpr int kIirRingStages32 = CBL_IIR_STAGES32;    // float-staged kernels (sweep: profiles/r1_sweeps.txt)
// look-back descriptors: a slot holds this bit pattern (a NaN no computation produces) until its owner
// has written it; a value that happens to equal it is nudged to the neighbouring NaN pattern
constexpr unsigned long long kDescEmpty = 0xFFFFFFFFFFFFFFFFull;
// offsets the integer->double mantissa injection adds to the sample value
constexpr double kBiasWords = 1052672.0;   // 2^20 + 2^12
constexpr double kBiasCs16  = 2129920.0;   // 2^21 + 2^15

// returns cudaError_t as int; y32 selects the float-staged kernel (iir_has_y32 says whether one exists)
// float-staged kernels are built for the parallel realisation only
inline bool iir_has_y32(int K, bool par) { (void)K; return par; }
int launch_iir(int K, int D, int in_kind, bool par, bool y32, const CUtensorMap& tmap, const IirParams& p,
               const IirTables& t, int num_sms, cudaStream_t st);
void launch_iir_tail(const IirTailParams& p, cudaStream_t st);
// state <- state_next (n_streams x per_stream doubles), except for streams with gate[s * gate_stride] != 0
void launch_commit_state(double* state, const double* state_next, size_t n_streams, size_t per_stream, const int32_t* gate,
                         uint32_t gate_stride, cudaStream_t st);

}  // namespace cbl
