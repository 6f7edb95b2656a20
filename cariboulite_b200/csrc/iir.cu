// Kernel (b) dispatch + the sequential tail kernel (samples that do not fill a whole run).
#include "iir_kernel.cuh"

namespace cbl {

iir_launch_fn iir_launch_k1(int, int, bool, bool); iir_launch_fn iir_launch_k2(int, int, bool, bool);
iir_launch_fn iir_launch_k3(int, int, bool, bool); iir_launch_fn iir_launch_k4(int, int, bool, bool);

bool iir_has_y32(int K, bool par) { return K == 1 || par; }

int launch_iir(int K, int D, int in_kind, bool par, bool y32, const CUtensorMap& tmap, const IirParams& p,
               const IirTables& t, int num_sms, cudaStream_t st)
{
    iir_launch_fn fn = nullptr;
    switch (K) {
        case 1: fn = iir_launch_k1(D, in_kind, par, y32); break;
        case 2: fn = iir_launch_k2(D, in_kind, par, y32); break;
        case 3: fn = iir_launch_k3(D, in_kind, par, y32); break;
        case 4: fn = iir_launch_k4(D, in_kind, par, y32); break;
    }
    if (!fn) return (int)cudaErrorInvalidValue;
    return fn(tmap, p, t, num_sms, st);
}

// ------------------------------------------------------------------ sequential tail
// One thread per stream walks samples [first, n) with the same (normalised) arithmetic, starting
// from and updating the stream state.  Used for the < one-run remainder of a submit.
template <int K, bool PAR>
__global__ void iir_tail_kernel(const __grid_constant__ IirTailParams P)
{
    const uint32_t stream = blockIdx.x * blockDim.x + threadIdx.x;
    if (stream >= P.n_streams) return;
    if (P.gate != nullptr && *P.gate != 0) return;
    constexpr int S = 2 * K;
    double vi[K][2], vq[K][2];                  // registers: K and the realisation are compile-time here
    double* sp = P.state + (size_t)stream * (2 * S);
#pragma unroll
    for (int k = 0; k < K; k++) {
        vi[k][0] = sp[2 * k]; vi[k][1] = sp[2 * k + 1];
        vq[k][0] = sp[S + 2 * k]; vq[k][1] = sp[S + 2 * k + 1];
    }
    // decimation phase and output index are carried along: a 64-bit divide per sample would dominate this kernel
    uint32_t ph = (uint32_t)(P.first % (uint64_t)P.D);
    uint64_t o = (uint64_t)stream * P.out_stride + (P.first + (uint64_t)P.D - 1) / (uint64_t)P.D;   // next kept sample's slot
    // one filter step + the kept-output store
    auto step = [&](double xi, double xq) {
        double ii = xi, iq = xq;
#pragma unroll
        for (int k = 0; k < K; k++) {
            if (PAR) {    // all sections see x; the output taps the state before the update
                ii = fma(P.c2[k], vi[k][1], fma(P.c1[k], vi[k][0], ii));
                iq = fma(P.c2[k], vq[k][1], fma(P.c1[k], vq[k][0], iq));
                const double wi = fma(-P.a1[k], vi[k][0], fma(-P.a2[k], vi[k][1], xi));
                const double wq = fma(-P.a1[k], vq[k][0], fma(-P.a2[k], vq[k][1], xq));
                vi[k][1] = vi[k][0]; vi[k][0] = wi;
                vq[k][1] = vq[k][0]; vq[k][0] = wq;
            } else {
                const double ti = fma(-P.a2[k], vi[k][1], ii), wi = fma(-P.a1[k], vi[k][0], ti);
                const double tq = fma(-P.a2[k], vq[k][1], iq), wq = fma(-P.a1[k], vq[k][0], tq);
                ii = fma(P.c2[k], vi[k][1], fma(P.c1[k], vi[k][0], wi));
                iq = fma(P.c2[k], vq[k][1], fma(P.c1[k], vq[k][0], wq));
                vi[k][1] = vi[k][0]; vi[k][0] = wi;
                vq[k][1] = vq[k][0]; vq[k][0] = wq;
            }
        }
        const bool kept = ph == 0;
        if (++ph == (uint32_t)P.D) ph = 0;
        if (kept) {
            const double yi = ii * P.gain, yq = iq * P.gain;
            if (P.out_mode == OUT_F32) {
                reinterpret_cast<float2*>(P.out)[o] = make_float2((float)yi, (float)yq);
            } else if (P.out_mode == OUT_F64) {
                reinterpret_cast<double2*>(P.out)[o] = make_double2(yi, yq);
            } else {
                const int qi = (int)(int16_t)__float2int_rz((float)yi);
                const int qq = (int)(int16_t)__float2int_rz((float)yq);
                if (P.out_mode == OUT_Q16_CS16)
                    reinterpret_cast<uint32_t*>(P.out)[o] = ((uint32_t)qi & 0xFFFFu) | ((uint32_t)qq << 16);
                else if (P.out_mode == OUT_Q16_CF32)
                    reinterpret_cast<float2*>(P.out)[o] = make_float2((float)qi * (1.0f / 4096.0f), (float)qq * (1.0f / 4096.0f));
                else if (P.out_mode == OUT_Q16_CF64)
                    reinterpret_cast<double2*>(P.out)[o] = make_double2((double)qi * (1.0 / 4096.0), (double)qq * (1.0 / 4096.0));
                else
                    reinterpret_cast<uint16_t*>(P.out)[o] =
                        (uint16_t)(((uint32_t)(qi >> 5) & 0xFFu) | (((uint32_t)(qq >> 5) & 0xFFu) << 8));
            }
            o++;
        }
    };
    // The loads do not depend on the recurrence: fetch 16 samples at a time, then walk them (one global round
    // trip per 16 samples instead of one per sample - the tail is pure latency).
    constexpr int CH = 16;
    const uint64_t base = (uint64_t)stream * P.in_stride;
    for (uint64_t n0 = P.first; n0 < P.n; n0 += CH) {
        const int cnt = (int)((P.n - n0) < (uint64_t)CH ? (P.n - n0) : (uint64_t)CH);
        if (P.in_kind == IN_CF32) {
            float2 v[CH];
#pragma unroll
            for (int k = 0; k < CH; k++) v[k] = k < cnt ? __ldg(reinterpret_cast<const float2*>(P.in) + base + n0 + k) : make_float2(0.f, 0.f);
#pragma unroll
            for (int k = 0; k < CH; k++) if (k < cnt) step((double)v[k].x, (double)v[k].y);
        } else {
            uint32_t w[CH];
#pragma unroll
            for (int k = 0; k < CH; k++) w[k] = k < cnt ? __ldg(reinterpret_cast<const uint32_t*>(P.in) + base + n0 + k) : 0u;
#pragma unroll
            for (int k = 0; k < CH; k++) {
                if (k >= cnt) break;
                if (P.in_kind == IN_WORDS) {
                    const int h = hi13(w[k]), l = lo13(w[k]);
                    step((double)(P.layout ? l : h), (double)(P.layout ? h : l));
                } else {
                    step((double)(int)(int16_t)(w[k] & 0xFFFFu), (double)((int32_t)w[k] >> 16));
                }
            }
        }
    }
#pragma unroll
    for (int k = 0; k < K; k++) {
        sp[2 * k] = vi[k][0]; sp[2 * k + 1] = vi[k][1];
        sp[S + 2 * k] = vq[k][0]; sp[S + 2 * k + 1] = vq[k][1];
    }
}

void launch_iir_tail(const IirTailParams& p, cudaStream_t st)
{
    if (p.n_streams == 0 || p.first >= p.n) return;
    const unsigned grid = (p.n_streams + 31) / 32;
    switch (p.K * 2 + (p.par ? 1 : 0)) {
        case 2: iir_tail_kernel<1, false><<<grid, 32, 0, st>>>(p); break;
        case 3: iir_tail_kernel<1, true><<<grid, 32, 0, st>>>(p); break;
        case 4: iir_tail_kernel<2, false><<<grid, 32, 0, st>>>(p); break;
        case 5: iir_tail_kernel<2, true><<<grid, 32, 0, st>>>(p); break;
        case 6: iir_tail_kernel<3, false><<<grid, 32, 0, st>>>(p); break;
        case 7: iir_tail_kernel<3, true><<<grid, 32, 0, st>>>(p); break;
        case 8: iir_tail_kernel<4, false><<<grid, 32, 0, st>>>(p); break;
        case 9: iir_tail_kernel<4, true><<<grid, 32, 0, st>>>(p); break;
        default: break;
    }
}

__global__ void commit_state_kernel(double* state, const double* next, size_t n, const int32_t* gate)
{
    if (gate != nullptr && *gate != 0) return;
    for (size_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) state[i] = next[i];
}

void launch_commit_state(double* state, const double* state_next, size_t n, const int32_t* gate, cudaStream_t st)
{
    if (n == 0) return;
    commit_state_kernel<<<(unsigned)((n + 255) / 256 > 64 ? 64 : (n + 255) / 256), 256, 0, st>>>(state, state_next, n, gate);
}

}  // namespace cbl
