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

// ------------------------------------------------------------------ sequential tail
// One warp per stream; lanes 0 and 1 (rails I and Q: same coefficients, never shared state) walk samples [first, n)
// with the same (normalised) arithmetic, starting from and updating the stream state.  Used for the < one-run
// remainder of a submit.  Pure latency: the whole warp first brings up to kTailChunk samples into shared memory with
// one round of coalesced loads (a thread fetching its own samples 16 at a time paid a global round trip per 16
// samples: 16 us for C4's 96-sample remainders), K and the realisation are compile-time so the state lives in
// registers, and the decimation phase is carried instead of divided.
constexpr int kTailWarps = 4;        // streams per block
constexpr int kTailChunk = 256;      // samples staged per round
template <int K, bool PAR>
__global__ void __launch_bounds__(kTailWarps * 32) iir_tail_kernel(const __grid_constant__ IirTailParams P)
{
    __shared__ uint32_t s_buf[kTailWarps][kTailChunk * 2];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t stream = blockIdx.x * kTailWarps + warp, rail = lane & 1u;          // rail 0 = I, 1 = Q
    if (stream >= P.n_streams) return;
    if (P.gate != nullptr && P.gate[(size_t)stream * P.gate_stride] != 0) return;
    const bool walker = lane < 2;
    uint32_t* const buf = s_buf[warp];
    constexpr int S = 2 * K;
    double v[K][2];
    double* sp = P.state + (size_t)stream * (2 * S) + rail * S;
#pragma unroll
    for (int k = 0; k < K; k++) { v[k][0] = sp[2 * k]; v[k][1] = sp[2 * k + 1]; }
    uint32_t ph = (uint32_t)(P.first % (uint64_t)P.D);
    uint64_t o = ((uint64_t)stream * P.out_stride + (P.first + (uint64_t)P.D - 1) / (uint64_t)P.D) * 2 + rail;   // next kept sample's component
    // one filter step + the kept-output store
    auto step = [&](double x) {
        double in = x;
#pragma unroll
        for (int k = 0; k < K; k++) {
            if (PAR) {      // all sections see x; the output taps the state before the update
                in = fma(P.c2[k], v[k][1], fma(P.c1[k], v[k][0], in));
                const double w = fma(-P.a1[k], v[k][0], fma(-P.a2[k], v[k][1], x));
                v[k][1] = v[k][0]; v[k][0] = w;
            } else {
                const double tt = fma(-P.a2[k], v[k][1], in), w = fma(-P.a1[k], v[k][0], tt);
                in = fma(P.c2[k], v[k][1], fma(P.c1[k], v[k][0], w));
                v[k][1] = v[k][0]; v[k][0] = w;
            }
        }
        const bool kept = ph == 0;
        if (++ph == (uint32_t)P.D) ph = 0;
        if (kept) {
            const double y = in * P.gain;
            if (P.out_mode == OUT_F32) {
                reinterpret_cast<float*>(P.out)[o] = (float)y;
            } else if (P.out_mode == OUT_F64) {
                reinterpret_cast<double*>(P.out)[o] = y;
            } else {
                const int q = (int)(int16_t)__float2int_rz((float)y);
                if (P.out_mode == OUT_Q16_CS16) reinterpret_cast<int16_t*>(P.out)[o] = (int16_t)q;
                else if (P.out_mode == OUT_Q16_CF32) reinterpret_cast<float*>(P.out)[o] = (float)q * (1.0f / 4096.0f);
                else if (P.out_mode == OUT_Q16_CF64) reinterpret_cast<double*>(P.out)[o] = (double)q * (1.0 / 4096.0);
                else reinterpret_cast<uint8_t*>(P.out)[o] = (uint8_t)((uint32_t)(q >> 5) & 0xFFu);
            }
            o += 2;
        }
    };
    constexpr int CH = 8;
    const uint64_t base = (uint64_t)stream * P.in_stride;
    const uint32_t wps = P.in_kind == IN_CF32 ? 2u : 1u;         // 32-bit words per sample
    // which half of a word / element this rail reads
    const bool upper = P.in_kind == IN_WORDS ? (rail == 0) != (P.layout != 0) : rail == 1;    // words: I = hi13 unless HiF
    for (uint64_t n0 = P.first; n0 < P.n; n0 += kTailChunk) {
        const uint32_t cnt = (uint32_t)((P.n - n0) < (uint64_t)kTailChunk ? (P.n - n0) : (uint64_t)kTailChunk);
        const uint32_t* src = reinterpret_cast<const uint32_t*>(P.in) + (base + n0) * wps;
        for (uint32_t i = lane; i < cnt * wps; i += 32) buf[i] = __ldg(src + i);
        __syncwarp();
        if (walker) {
            for (uint32_t k0 = 0; k0 < cnt; k0 += CH) {
                uint32_t w[CH];
#pragma unroll
                for (int k = 0; k < CH; k++) w[k] = k0 + k < cnt ? buf[(k0 + k) * wps + (wps == 2 ? rail : 0u)] : 0u;
#pragma unroll
                for (int k = 0; k < CH; k++) {
                    if (k0 + k >= cnt) break;
                    if (P.in_kind == IN_CF32) step((double)__uint_as_float(w[k]));
                    else if (P.in_kind == IN_WORDS) step((double)(upper ? hi13(w[k]) : lo13(w[k])));
                    else step(upper ? (double)((int32_t)w[k] >> 16) : (double)(int)(int16_t)(w[k] & 0xFFFFu));
                }
            }
        }
        __syncwarp();
    }
    if (walker) {
#pragma unroll
        for (int k = 0; k < K; k++) { sp[2 * k] = v[k][0]; sp[2 * k + 1] = v[k][1]; }
    }
}

void launch_iir_tail(const IirTailParams& p, cudaStream_t st)
{
    if (p.n_streams == 0 || p.first >= p.n) return;
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
