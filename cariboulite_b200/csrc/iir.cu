// Kernel (b) dispatch + the sequential tail kernel (samples that do not fill a whole run).
#include "iir_kernel.cuh"

namespace cbl {

iir_launch_fn iir_launch_k1(int, int, bool, bool); iir_launch_fn iir_launch_k2(int, int, bool, bool);
iir_launch_fn iir_launch_k3(int, int, bool, bool); iir_launch_fn iir_launch_k4(int, int, bool, bool);
iir_launch_fn iir_rerun_k1(int, int); iir_launch_fn iir_rerun_k2(int, int);      // iir_r{1,2,3,4}.cu
iir_launch_fn iir_rerun_k3(int, int); iir_launch_fn iir_rerun_k4(int, int);
iir_launch_fn iir_rerun_ap_k1(int, int); iir_launch_fn iir_rerun_ap_k2(int, int);      // iir_ra{1,2,3,4}.cu
iir_launch_fn iir_rerun_ap_k3(int, int); iir_launch_fn iir_rerun_ap_k4(int, int);


int launch_iir(int K, int D, int in_kind, bool par, bool y32, const CUtensorMap& tmap, const IirParams& p,
               const IirTables& t, int num_sms, cudaStream_t st)
{
    iir_launch_fn fn = nullptr;
    if (iir_is_rerun(K, D, par) && p.ap) {
        switch (K) {
            case 1: fn = iir_rerun_ap_k1(D, in_kind); break;
            case 2: fn = iir_rerun_ap_k2(D, in_kind); break;
            case 3: fn = iir_rerun_ap_k3(D, in_kind); break;
            case 4: fn = iir_rerun_ap_k4(D, in_kind); break;
        }
    } else if (iir_is_rerun(K, D, par)) {
        switch (K) {
            case 1: fn = iir_rerun_k1(D, in_kind); break;
            case 2: fn = iir_rerun_k2(D, in_kind); break;
            case 3: fn = iir_rerun_k3(D, in_kind); break;
            case 4: fn = iir_rerun_k4(D, in_kind); break;
        }
    } else switch (K) {
        case 1: fn = iir_launch_k1(D, in_kind, par, y32); break;
        case 2: fn = iir_launch_k2(D, in_kind, par, y32); break;
        case 3: fn = iir_launch_k3(D, in_kind, par, y32); break;
        case 4: fn = iir_launch_k4(D, in_kind, par, y32); break;
    }
    if (!fn) return (int)cudaErrorInvalidValue;
    return fn(tmap, p, t, num_sms, st);
}

// ------------------------------------------------------------------ sequential tail (and state commit)
// One warp per stream walks samples [first, n) - the < one-run remainder of a submit - with the same (normalised)
// arithmetic, starting from and updating the stream state; with state_src it starts from the state the scan kernel
// left in its scratch array, so that the commit of a submit is this kernel too (first == n: commit only).
// Pure latency, so everything that is not the recurrence is taken off the walkers: the whole warp brings up to
// kTailChunk samples in with one round of coalesced loads and converts them (both rails) into shared memory; lanes 0
// and 1 (rails I and Q: same coefficients, never shared state) run the recurrence over those doubles - K and the
// realisation are compile-time, the state lives in registers, the decimation phase is carried instead of divided - and
// leave the kept outputs in shared memory; the whole warp converts and stores them.
constexpr int kTailWarps = 4;        // streams per block
constexpr int kTailChunk = 256;      // samples staged per round
__device__ __forceinline__ void tail_store(const IirTailParams& P, uint64_t o, double yi, double yq)
{
    if (P.out_mode == OUT_F32) {
        reinterpret_cast<float2*>(P.out)[o] = make_float2((float)yi, (float)yq);
    } else if (P.out_mode == OUT_F64) {
        reinterpret_cast<double2*>(P.out)[o] = make_double2(yi, yq);
    } else {
        const int qi = (int)(int16_t)__float2int_rz((float)yi), qq = (int)(int16_t)__float2int_rz((float)yq);
        if (P.out_mode == OUT_Q16_CS16) reinterpret_cast<uint32_t*>(P.out)[o] = ((uint32_t)qi & 0xFFFFu) | ((uint32_t)qq << 16);
        else if (P.out_mode == OUT_Q16_CF32) reinterpret_cast<float2*>(P.out)[o] = make_float2((float)qi * (1.0f / 4096.0f), (float)qq * (1.0f / 4096.0f));
        else if (P.out_mode == OUT_Q16_CF64) reinterpret_cast<double2*>(P.out)[o] = make_double2((double)qi * (1.0 / 4096.0), (double)qq * (1.0 / 4096.0));
        else reinterpret_cast<uint16_t*>(P.out)[o] = (uint16_t)(((uint32_t)(qi >> 5) & 0xFFu) | (((uint32_t)(qq >> 5) & 0xFFu) << 8));
    }
}
template <int K, bool PAR>
__global__ void __launch_bounds__(kTailWarps * 32) iir_tail_kernel(const __grid_constant__ IirTailParams P)
{
    __shared__ double s_x[kTailWarps][2][kTailChunk];      // converted samples per rail
    __shared__ double s_y[kTailWarps][2][kTailChunk];      // kept outputs per rail
    __shared__ uint32_t s_kept[kTailWarps];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t stream = blockIdx.x * kTailWarps + warp, rail = lane & 1u;          // rail 0 = I, 1 = Q
    if (blockIdx.x == 0 && P.mirror_dst != nullptr)        // (the verdicts of the submit's alignment search are final by now)
        for (uint32_t i = threadIdx.x; i < P.mirror_words; i += blockDim.x) P.mirror_dst[i] = P.mirror_src[i];
    if (stream >= P.n_streams) return;
    if (P.gate != nullptr && P.gate[(size_t)stream * P.gate_stride] != 0) return;
    const bool walker = lane < 2;
    constexpr int S = 2 * K;
    double v[K][2];
    double* sp = P.state + (size_t)stream * (2 * S) + rail * S;
    const double* sp_in = (P.state_src != nullptr ? P.state_src : P.state) + (size_t)stream * (2 * S) + rail * S;
#pragma unroll
    for (int k = 0; k < K; k++) { v[k][0] = sp_in[2 * k]; v[k][1] = sp_in[2 * k + 1]; }
    uint32_t ph = (uint32_t)(P.first % (uint64_t)P.D);
    uint64_t o = (uint64_t)stream * P.out_stride + (P.first + (uint64_t)P.D - 1) / (uint64_t)P.D;      // next kept sample
    const uint64_t base = (uint64_t)stream * P.in_stride;
    const bool a_upper = P.in_kind == IN_WORDS && P.layout == 0;     // words: rail A (I) = hi13 unless HiF
    const double* __restrict__ const xs = s_x[warp][rail];
    double* __restrict__ const ys = s_y[warp][rail];
    for (uint64_t n0 = P.first; n0 < P.n; n0 += kTailChunk) {
        const uint32_t cnt = (uint32_t)((P.n - n0) < (uint64_t)kTailChunk ? (P.n - n0) : (uint64_t)kTailChunk);
        for (uint32_t i = lane; i < cnt; i += 32) {
            double xa, xb;
            if (P.in_kind == IN_CF32) {
                const float2 f = __ldg(reinterpret_cast<const float2*>(P.in) + base + n0 + i);
                xa = (double)f.x; xb = (double)f.y;
            } else {
                const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(P.in) + base + n0 + i);
                if (P.in_kind == IN_WORDS) {
                    const double h = (double)hi13(w), l = (double)lo13(w);
                    xa = a_upper ? h : l; xb = a_upper ? l : h;
                } else {
                    xa = (double)(int)(int16_t)(w & 0xFFFFu); xb = (double)((int32_t)w >> 16);
                }
            }
            s_x[warp][0][i] = xa; s_x[warp][1][i] = xb;
        }
        __syncwarp();
        if (walker) {
            uint32_t nk = 0;
#pragma unroll 4
            for (uint32_t i = 0; i < cnt; i++) {
                const double x = xs[i];
                const bool kept = ph == 0;
                if (++ph == (uint32_t)P.D) ph = 0;
                double in = x;
                if (PAR) {
                    // all sections see x; the output taps the state before the update and is only worked out where
                    // it is kept (both walkers of a warp are at the same phase)
                    if (kept) {
#pragma unroll
                        for (int k = 0; k < K; k++) in = fma(P.c2[k], v[k][1], fma(P.c1[k], v[k][0], in));
                    }
#pragma unroll
                    for (int k = 0; k < K; k++) {
                        const double w = fma(-P.a1[k], v[k][0], fma(-P.a2[k], v[k][1], x));
                        v[k][1] = v[k][0]; v[k][0] = w;
                    }
                } else {
#pragma unroll
                    for (int k = 0; k < K; k++) {
                        const double tt = fma(-P.a2[k], v[k][1], in), w = fma(-P.a1[k], v[k][0], tt);
                        in = fma(P.c2[k], v[k][1], fma(P.c1[k], v[k][0], w));
                        v[k][1] = v[k][0]; v[k][0] = w;
                    }
                }
                if (kept) ys[nk++] = in * P.gain;
            }
            if (lane == 0) s_kept[warp] = nk;
        }
        __syncwarp();
        const uint32_t nk = s_kept[warp];
        for (uint32_t j = lane; j < nk; j += 32) tail_store(P, o + j, s_y[warp][0][j], s_y[warp][1][j]);
        o += nk;
        __syncwarp();
    }
    if (walker) {
#pragma unroll
        for (int k = 0; k < K; k++) { sp[2 * k] = v[k][0]; sp[2 * k + 1] = v[k][1]; }
    }
}

void launch_iir_tail(const IirTailParams& p, cudaStream_t st)
{
    if (p.n_streams == 0 || (p.first >= p.n && p.state_src == nullptr)) return;
    const unsigned grid = (p.n_streams + kTailWarps - 1) / kTailWarps;      // one warp per stream
    switch (p.K * 2 + (p.par ? 1 : 0)) {
        case 2: iir_tail_kernel<1, false><<<grid, kTailWarps * 32, 0, st>>>(p); break;
        case 3: iir_tail_kernel<1, true><<<grid, kTailWarps * 32, 0, st>>>(p); break;
        case 4: iir_tail_kernel<2, false><<<grid, kTailWarps * 32, 0, st>>>(p); break;
        case 5: iir_tail_kernel<2, true><<<grid, kTailWarps * 32, 0, st>>>(p); break;
        case 6: iir_tail_kernel<3, false><<<grid, kTailWarps * 32, 0, st>>>(p); break;
        case 7: iir_tail_kernel<3, true><<<grid, kTailWarps * 32, 0, st>>>(p); break;
        case 8: iir_tail_kernel<4, false><<<grid, kTailWarps * 32, 0, st>>>(p); break;
        case 9: iir_tail_kernel<4, true><<<grid, kTailWarps * 32, 0, st>>>(p); break;
        default: break;
    }
}

__global__ void commit_state_kernel(double* state, const double* next, size_t n, size_t per_stream, const int32_t* gate,
                                    uint32_t gate_stride)
{
    for (size_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        if (gate != nullptr && gate[(i / per_stream) * gate_stride] != 0) continue;
        state[i] = next[i];
    }
}

void launch_commit_state(double* state, const double* state_next, size_t n_streams, size_t per_stream, const int32_t* gate,
                         uint32_t gate_stride, cudaStream_t st)
{
    const size_t n = n_streams * per_stream;
    if (n == 0) return;
    commit_state_kernel<<<(unsigned)((n + 255) / 256 > 64 ? 64 : (n + 255) / 256), 256, 0, st>>>(state, state_next, n, per_stream,
                                                                                                 gate, gate_stride);
}

}  // namespace cbl
