// Kernel (b): cascaded-biquad IIR as a block-parallel linear-recurrence scan with the
// decimation fused into the store - shared declarations.
//
// Math (per rail; the two rails I and Q share coefficients, never state):
//   reference stage k (LC/iir/iir/State.h:116-126):  w = in - a1 v1 - a2 v2 ; out = b0 w + b1 v1 + b2 v2
//   kernel coordinates: every stage's state is divided by g_k = prod_{j<k} b0_j, so that
//       w' = in' - a1 v1' - a2 v2' ;  in'_{k+1} = w' + c1 v1' + c2 v2'   (c = b/b0)
//   and y = G * in'_K with G = prod b0.  One multiply less per inner stage, same transfer function.
//   The 2K-vector s = (v1'_0, v2'_0, v1'_1, ...) obeys s[n+1] = A s[n] + B x[n]; in'_K[n] = Cu s[n] + x-term.
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

enum { IN_WORDS = 0, IN_CS16 = 1, IN_CF32 = 2 };
// epilogue modes
enum { OUT_F32 = 0,          // (float)y                              (parity F32)
       OUT_Q16_CS16 = 1,     // (int16)(float)y                       (parity Q16, CS16)
       OUT_Q16_CF32 = 2,     // (float)q16 / 4096                     (parity Q16, CF32)
       OUT_Q16_CF64 = 3,     // (double)q16 / 4096
       OUT_Q16_CS8 = 4,      // (int8)((q16 >> 5) & 0xFF)
       OUT_F64 = 5 };        // y as double (parity F32, CF64)

// Filter constants that sit in the kernel-parameter constant bank (operands of DFMA).
struct IirTables {
    double a1[kMaxK], a2[kMaxK], c1[kMaxK], c2[kMaxK];
    double gain;                          // G * output scale
    double Mp[5][kMaxS * kMaxS];          // M^(2^j), row-major S x S (S = 2K)
    double R[kMaxNK][kMaxS];              // R[m] = Cu A^(m D)
};

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
    int      layout;            // IN_WORDS: 0 = S1G (I = hi13), 1 = HiF (I = lo13)
    int      out_mode;
    const double* mtile_pow;    // [window+1][S*S]  M_tile^j (row-major)
    const double* plane;        // [S*S][33]        plane[e*33 + t] = (M^t)[e]
    double*   desc_agg;         // [total_tiles][2S]  zero-state tile end states
    double*   desc_incl;        // [total_tiles][2S]  true tile end states
    uint32_t* desc_flag;        // [total_tiles]      0 none, 1 aggregate, 2 inclusive (zeroed per launch)
    uint32_t* ticket;           // zeroed per launch
    const double* state_in;     // [n_streams][2S] kernel coordinates
    double*   state_out;        // [n_streams][2S]
    void*     out;              // kept outputs, NK per run, streams out_stride elements apart
    uint64_t  out_stride;
    const int32_t* gate;        // optional: the launch is a no-op when *gate != 0 (misaligned capture)
};

struct IirTailParams {
    uint32_t n_streams;
    const void* in; uint64_t in_stride;     // elements
    uint64_t first;  uint64_t n;            // samples [first, n) of every stream
    void* out; uint64_t out_stride;
    int in_kind, layout, out_mode, K, D;
    double a1[kMaxK], a2[kMaxK], c1[kMaxK], c2[kMaxK];
    double gain;
    double* state;                          // [n_streams][2S] kernel coordinates, updated in place
    const int32_t* gate;                    // optional: no-op when *gate != 0
};

// run length for a decimation factor
__host__ __device__ constexpr int iir_run_len(int D) { return D == 1 ? 32 : (D == 2 ? 64 : 128); }
constexpr int kIirWarpsPerCta = 4;
constexpr int kIirRingStages  = 2;

size_t iir_smem_bytes(int K, int D, int in_kind);
// returns cudaError_t as int
int launch_iir(int K, int D, int in_kind, const CUtensorMap& tmap, const IirParams& p, const IirTables& t,
               int grid_ctas, cudaStream_t st);
int iir_max_ctas_per_sm(int K, int D, int in_kind);
void launch_iir_tail(const IirTailParams& p, cudaStream_t st);
// state <- state_next (n doubles) unless *gate != 0
void launch_commit_state(double* state, const double* state_next, size_t n, const int32_t* gate, cudaStream_t st);

}  // namespace cbl
