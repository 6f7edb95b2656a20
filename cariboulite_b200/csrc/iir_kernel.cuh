// Kernel (b) template.  See iir.cuh for the math.  Included by iir_k{1,2,3,4}.cu.
#pragma once
#include "iir.cuh"

namespace cbl {

// ------------------------------------------------------------------ PTX helpers: mbarrier + TMA
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
// 3-D tiled TMA load: box (32 uint32 = 128 bytes of one run, 32 runs, 1 stream) -> 4 KiB of shared
// memory, 128B-swizzled.  x counts uint32 units whatever the sample type.
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* tmap, int x, int y, int z, uint32_t bar)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(dst), "l"(tmap), "r"(x), "r"(y), "r"(z), "r"(bar) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t* p, uint32_t v)
{
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// explicit shared-space accesses (the carve-up below goes through integers, which would
// otherwise demote these to generic LD/ST)
__device__ __forceinline__ uint4 lds128(uint32_t a)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ double lds64(uint32_t a)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts64(uint32_t a, double v)
{
    asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory");
}
// int32 -> double.  Default: I2F.F64 (conversion pipe).  -DCBL_CONV_MAGIC: exponent trick + one DADD.
__device__ __forceinline__ double i2d(int v)
{
#ifdef CBL_CONV_MAGIC
    return __hiloint2double(0x43300000, (int)((uint32_t)v ^ 0x80000000u)) - 4503601774854144.0;   // 2^52 + 2^31
#else
    return (double)v;
#endif
}
__device__ __forceinline__ double shfl_up_f64(double v, int d)
{
    return __shfl_up_sync(0xffffffffu, v, d);
}

template <int K, int D, int IN>
struct IirCfg {
    static constexpr int S     = 2 * K;                       // state dimension per rail
    static constexpr int LT    = iir_run_len(D);              // samples per run
    static constexpr int NK    = LT / D;                      // kept outputs per run
    static constexpr int SPS   = (IN == IN_CF32) ? 16 : 32;   // samples per 128-byte row step
    static constexpr int NSTEP = LT / SPS;
    static constexpr int STAGES = (NSTEP < kIirRingStages) ? NSTEP : kIirRingStages;
    static constexpr int WPC   = kIirWarpsPerCta;
    static constexpr int RING_BYTES    = STAGES * 4096;
    static constexpr int YSTAGE_BYTES  = NK * 2 * 32 * 8;
    static constexpr int SCRATCH_BYTES = 2 * S * 8 * 2;        // S_in broadcast + spare
    static constexpr int WARP_BYTES = RING_BYTES + YSTAGE_BYTES + SCRATCH_BYTES;
    static constexpr int SMEM_BYTES = WPC * WARP_BYTES + WPC * STAGES * 8 + 1024;   // + mbarriers + alignment slack
};

// One cascade step on one rail.  v = (v1,v2) per stage.  Returns in'_K if LAST else undefined.
template <int K, bool NEED_OUT>
__device__ __forceinline__ double cascade_step(const IirTables& T, double (&v)[K][2], double x)
{
    double in = x;
#pragma unroll
    for (int k = 0; k < K; k++) {
        const double t = fma(-T.a2[k], v[k][1], in);
        const double w = fma(-T.a1[k], v[k][0], t);
        if (k < K - 1 || NEED_OUT) in = fma(T.c2[k], v[k][1], fma(T.c1[k], v[k][0], w));
        v[k][1] = v[k][0];
        v[k][0] = w;
    }
    return in;
}

template <int K, int D, int IN>
__global__ void __launch_bounds__(kIirWarpsPerCta * 32)
iir_scan_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ IirParams P,
                const __grid_constant__ IirTables T)
{
    using C = IirCfg<K, D, IN>;
    constexpr int S = C::S;
    extern __shared__ uint8_t smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // shared-memory carve-up, as 32-bit shared-window addresses (1024-byte aligned for the 128B swizzle)
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t ring_u32 = base + warp * C::RING_BYTES;
    const uint32_t ystage = base + C::WPC * C::RING_BYTES + warp * C::YSTAGE_BYTES;
    const uint32_t scratch = base + C::WPC * (C::RING_BYTES + C::YSTAGE_BYTES) + warp * C::SCRATCH_BYTES;
    const uint32_t bar_u32 = base + C::WPC * C::WARP_BYTES + warp * C::STAGES * 8;

    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < C::STAGES; s++) mbar_init(bar_u32 + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();

    if (P.gate != nullptr && __ldg(P.gate) != 0) return;

    uint32_t gstep = 0;    // steps consumed by this warp so far (ring position / parity)

    for (;;) {
        uint32_t tile = 0;
        if (lane == 0) tile = atomicAdd(P.ticket, 1u);
        tile = __shfl_sync(0xffffffffu, tile, 0);
        if (tile >= P.total_tiles) break;
        const uint32_t stream = tile / P.tiles_per_stream;
        const uint32_t tis = tile - stream * P.tiles_per_stream;       // tile index within the stream
        const int run0 = (int)tis * 32;

        // ---- phase 1: zero-state pass over the run, TMA-fed through a per-warp ring
        if (lane == 0) {
#pragma unroll
            for (int s = 0; s < C::STAGES; s++) {
                const uint32_t st = (gstep + s) % C::STAGES;
                mbar_expect_tx(bar_u32 + 8 * st, 4096);
                tma_load_3d(ring_u32 + st * 4096, &tmap, s * 32, run0, (int)stream, bar_u32 + 8 * st);
            }
        }
        double vi[K][2], vq[K][2];
#pragma unroll
        for (int k = 0; k < K; k++) { vi[k][0] = vi[k][1] = vq[k][0] = vq[k][1] = 0.0; }

#pragma unroll 1
        for (int step = 0; step < C::NSTEP; step++) {
            const uint32_t st = gstep % C::STAGES;
            mbar_wait(bar_u32 + 8 * st, (gstep / C::STAGES) & 1u);
            const uint32_t row = ring_u32 + st * 4096 + lane * 128;
#pragma unroll
            for (int c = 0; c < 8; c++) {
                const uint4 q = lds128(row + ((c ^ (lane & 7)) << 4));
                const uint32_t wv[4] = {q.x, q.y, q.z, q.w};
                constexpr int PER = (IN == IN_CF32) ? 2 : 4;     // samples per 16 bytes
#pragma unroll
                for (int e = 0; e < PER; e++) {
                    const int i = c * PER + e;                   // sample index within the step
                    double xi, xq;
                    if (IN == IN_WORDS) {
                        const int h = hi13(wv[e]), l = lo13(wv[e]);
                        xi = i2d(P.layout ? l : h);
                        xq = i2d(P.layout ? h : l);
                    } else if (IN == IN_CS16) {
                        xi = i2d((int)(int16_t)(wv[e] & 0xFFFFu));
                        xq = i2d((int32_t)wv[e] >> 16);
                    } else {
                        xi = (double)__uint_as_float(wv[2 * e]);
                        xq = (double)__uint_as_float(wv[2 * e + 1]);
                    }
                    // kept iff (step*SPS + i) % D == 0
                    const bool kept_static = (D <= C::SPS) ? (i % D == 0) : (i == 0);
                    if (kept_static) {
                        const bool kept = (D <= C::SPS) ? true : (((step * C::SPS) % D) == 0);
                        const double yi = cascade_step<K, true>(T, vi, xi);
                        const double yq = cascade_step<K, true>(T, vq, xq);
                        if (kept) {
                            const int m = (step * C::SPS + i) / D;
                            sts64(ystage + ((m * 2 + 0) * 32 + lane) * 8, yi);
                            sts64(ystage + ((m * 2 + 1) * 32 + lane) * 8, yq);
                        }
                    } else {
                        cascade_step<K, false>(T, vi, xi);
                        cascade_step<K, false>(T, vq, xq);
                    }
                }
            }
            __syncwarp();
            if (lane == 0 && step + C::STAGES < C::NSTEP) {
                mbar_expect_tx(bar_u32 + 8 * st, 4096);
                tma_load_3d(ring_u32 + st * 4096, &tmap, (step + C::STAGES) * 32, run0, (int)stream, bar_u32 + 8 * st);
            }
            gstep++;
        }

        // ---- phase 2: warp scan of run end states: E_t = sum_{j<=t} M^(t-j) f_j
        double f[2][S];
#pragma unroll
        for (int k = 0; k < K; k++) {
            f[0][2 * k] = vi[k][0]; f[0][2 * k + 1] = vi[k][1];
            f[1][2 * k] = vq[k][0]; f[1][2 * k + 1] = vq[k][1];
        }
#pragma unroll
        for (int j = 0; j < 5; j++) {
            const int d = 1 << j;
            double g[2][S];
#pragma unroll
            for (int r = 0; r < 2; r++)
#pragma unroll
                for (int c = 0; c < S; c++) g[r][c] = shfl_up_f64(f[r][c], d);
            if (lane >= d) {
#pragma unroll
                for (int r = 0; r < 2; r++)
#pragma unroll
                    for (int row = 0; row < S; row++) {
                        double acc = f[r][row];
                        // A (hence every power of it) is block lower-triangular: stage k only sees stages <= k
#pragma unroll
                        for (int col = 0; col < 2 * (row / 2 + 1); col++) acc = fma(T.Mp[j][row * S + col], g[r][col], acc);
                        f[r][row] = acc;
                    }
            }
        }
        // f = inclusive scan value E_lane.  Lane 31 holds the tile aggregate.
        double* agg = P.desc_agg + (size_t)tile * (2 * S);
        if (lane == 31) {
#pragma unroll
            for (int r = 0; r < 2; r++)
#pragma unroll
                for (int c = 0; c < S; c++) agg[r * S + c] = f[r][c];
            if (tis != 0 || !P.chained) { __threadfence(); st_release_u32(P.desc_flag + tile, 1u); }
        }

        // ---- decoupled look-back: state entering this tile, component `lane` (< 2S) per lane
        const int comp = lane < 2 * S ? lane : 0;
        const int crail = comp / S, crow = comp % S;
        double sin_c = 0.0;
        if (tis == 0) {
            sin_c = P.state_in[(size_t)stream * (2 * S) + comp];
        } else if (P.chained) {
            while (ld_acquire_u32(P.desc_flag + tile - 1) != 2u) { }
            sin_c = __ldcg(P.desc_incl + (size_t)(tile - 1) * (2 * S) + comp);
        } else {
            const int depth = (int)tis < P.window ? (int)tis : P.window;
            for (int j = 1; j <= depth; j++) {
                while (ld_acquire_u32(P.desc_flag + tile - j) == 0u) { }
                const double* a = P.desc_agg + (size_t)(tile - j) * (2 * S) + crail * S;
                const double* mp = P.mtile_pow + (size_t)(j - 1) * (S * S) + crow * S;
#pragma unroll
                for (int col = 0; col < S; col++) sin_c = fma(__ldg(mp + col), __ldcg(a + col), sin_c);
            }
            if ((int)tis <= P.window) {     // the window reaches the start of the stream
                const double* a = P.state_in + (size_t)stream * (2 * S) + crail * S;
                const double* mp = P.mtile_pow + (size_t)tis * (S * S) + crow * S;
#pragma unroll
                for (int col = 0; col < S; col++) sin_c = fma(__ldg(mp + col), __ldg(a + col), sin_c);
            }
        }
        __syncwarp();
        if (lane < 2 * S) sts64(scratch + lane * 8, sin_c);
        __syncwarp();
        double sin[2][S];
#pragma unroll
        for (int r = 0; r < 2; r++)
#pragma unroll
            for (int c = 0; c < S; c++) sin[r][c] = lds64(scratch + (r * S + c) * 8);

        // true state entering this lane's run: s0 = M^lane S_in + E_{lane-1}
        double s0[2][S];
#pragma unroll
        for (int r = 0; r < 2; r++)
#pragma unroll
            for (int c = 0; c < S; c++) {
                const double e = shfl_up_f64(f[r][c], 1);
                s0[r][c] = lane == 0 ? 0.0 : e;
            }
#pragma unroll
        for (int row = 0; row < S; row++)
#pragma unroll
            for (int col = 0; col < 2 * (row / 2 + 1); col++) {
                const double m = __ldg(P.plane + (row * S + col) * 33 + lane);
                s0[0][row] = fma(m, sin[0][col], s0[0][row]);
                s0[1][row] = fma(m, sin[1][col], s0[1][row]);
            }

        // true end state of the tile (chained mode) / of the stream (last tile)
        const bool last_tile = (tis == P.tiles_per_stream - 1);
        if (P.chained || (last_tile && P.write_state)) {
            const int lastl = last_tile && P.write_state ? (int)P.last_tile_runs - 1 : 31;
            // state after run `lastl`: M^(lastl+1) S_in + E_lastl, evaluated by lane lastl
            if (lane == lastl) {
                double* incl = P.desc_incl + (size_t)tile * (2 * S);
#pragma unroll
                for (int row = 0; row < S; row++) {
                    double ai = f[0][row], aq = f[1][row];
#pragma unroll
                    for (int col = 0; col < 2 * (row / 2 + 1); col++) {
                        const double m = __ldg(P.plane + (row * S + col) * 33 + lastl + 1);
                        ai = fma(m, sin[0][col], ai);
                        aq = fma(m, sin[1][col], aq);
                    }
                    incl[row] = ai; incl[S + row] = aq;
                    if (last_tile && P.write_state) {
                        P.state_out[(size_t)stream * (2 * S) + row] = ai;
                        P.state_out[(size_t)stream * (2 * S) + S + row] = aq;
                    }
                }
                if (P.chained) { __threadfence(); st_release_u32(P.desc_flag + tile, 2u); }
            }
        }

        // ---- phase 3: correct the staged outputs and store (decimation already applied)
        const uint32_t run = (uint32_t)run0 + lane;
        if (run < P.n_runs) {
            const uint64_t o0 = (uint64_t)stream * P.out_stride + (uint64_t)run * C::NK;
#pragma unroll 4
            for (int m = 0; m < C::NK; m++) {
                double yi = lds64(ystage + ((m * 2 + 0) * 32 + lane) * 8);
                double yq = lds64(ystage + ((m * 2 + 1) * 32 + lane) * 8);
#pragma unroll
                for (int c = 0; c < S; c++) {
                    yi = fma(T.R[m][c], s0[0][c], yi);
                    yq = fma(T.R[m][c], s0[1][c], yq);
                }
                yi *= T.gain; yq *= T.gain;
                const uint64_t o = o0 + m;
                if (P.out_mode == OUT_F32) {
                    __stcs(reinterpret_cast<float2*>(P.out) + o, make_float2((float)yi, (float)yq));
                } else if (P.out_mode == OUT_F64) {
                    __stcs(reinterpret_cast<double2*>(P.out) + o, make_double2(yi, yq));
                } else {
                    // (int16_t)(float)y, CaribouliteStream.cpp:295-296: round to float, truncate toward zero
                    const int qi = (int)(int16_t)__float2int_rz((float)yi);
                    const int qq = (int)(int16_t)__float2int_rz((float)yq);
                    if (P.out_mode == OUT_Q16_CS16) {
                        reinterpret_cast<uint32_t*>(P.out)[o] = ((uint32_t)qi & 0xFFFFu) | ((uint32_t)qq << 16);
                    } else if (P.out_mode == OUT_Q16_CF32) {
                        __stcs(reinterpret_cast<float2*>(P.out) + o,
                               make_float2((float)qi * (1.0f / 4096.0f), (float)qq * (1.0f / 4096.0f)));
                    } else if (P.out_mode == OUT_Q16_CF64) {
                        __stcs(reinterpret_cast<double2*>(P.out) + o,
                               make_double2((double)qi * (1.0 / 4096.0), (double)qq * (1.0 / 4096.0)));
                    } else {   // OUT_Q16_CS8, CaribouliteStream.cpp:362-363
                        reinterpret_cast<uint16_t*>(P.out)[o] =
                            (uint16_t)(((uint32_t)(qi >> 5) & 0xFFu) | (((uint32_t)(qq >> 5) & 0xFFu) << 8));
                    }
                }
            }
        }
        __syncwarp();   // ystage / scratch are reused by the next tile
    }
}

template <int K, int D, int IN>
int launch_iir_t(const CUtensorMap& tmap, const IirParams& p, const IirTables& t, int grid, cudaStream_t st)
{
    using C = IirCfg<K, D, IN>;
    auto kern = iir_scan_kernel<K, D, IN>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    kern<<<grid, C::WPC * 32, C::SMEM_BYTES, st>>>(tmap, p, t);
    return (int)cudaGetLastError();
}

template <int K, int D, int IN>
int occupancy_iir_t()
{
    using C = IirCfg<K, D, IN>;
    auto kern = iir_scan_kernel<K, D, IN>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES);
    int n = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, C::WPC * 32, C::SMEM_BYTES) != cudaSuccess) return 0;
    return n;
}

#define CBL_IIR_DISPATCH_D(KK, IN_, CALL)                    \
    switch (D) {                                            \
        case 1:  return CALL<KK, 1, IN_>;                    \
        case 2:  return CALL<KK, 2, IN_>;                    \
        case 4:  return CALL<KK, 4, IN_>;                    \
        case 8:  return CALL<KK, 8, IN_>;                    \
        case 16: return CALL<KK, 16, IN_>;                   \
        case 32: return CALL<KK, 32, IN_>;                   \
        default: return nullptr;                            \
    }

typedef int (*iir_launch_fn)(const CUtensorMap&, const IirParams&, const IirTables&, int, cudaStream_t);
typedef int (*iir_occ_fn)();

template <int KK>
iir_launch_fn iir_pick_launch(int D, int in_kind)
{
    if (in_kind == IN_WORDS) { CBL_IIR_DISPATCH_D(KK, IN_WORDS, &launch_iir_t) }
    if (in_kind == IN_CS16)  { CBL_IIR_DISPATCH_D(KK, IN_CS16, &launch_iir_t) }
    if (in_kind == IN_CF32)  { CBL_IIR_DISPATCH_D(KK, IN_CF32, &launch_iir_t) }
    return nullptr;
}
template <int KK>
iir_occ_fn iir_pick_occ(int D, int in_kind)
{
    if (in_kind == IN_WORDS) { CBL_IIR_DISPATCH_D(KK, IN_WORDS, &occupancy_iir_t) }
    if (in_kind == IN_CS16)  { CBL_IIR_DISPATCH_D(KK, IN_CS16, &occupancy_iir_t) }
    if (in_kind == IN_CF32)  { CBL_IIR_DISPATCH_D(KK, IN_CF32, &occupancy_iir_t) }
    return nullptr;
}

}  // namespace cbl
