// Kernel (b), full-rate and lightly decimated form (iir_is_rerun(): D = 1, and D = 2 from three sections on; parallel
// realisation): the "rerun" kernel.
//
// The staged kernel (iir_kernel.cuh) keeps the zero-state outputs of a run in shared memory and corrects each
// kept one with R[m] . s0 later.  That pays when few outputs are kept; at D = 1 every sample is kept, the stage
// limits a run to 32 samples and every output costs a shared-memory round trip plus 2S correction FMAs.  Here a
// run is filtered TWICE instead: phase 1 from the steady state for its end state only (2K FMAs per rail and
// sample, no outputs), phase 2 and the look-back are those of the staged kernel, and phase 3 filters the same
// run again from its true entry state s0 - now in exactly the reference's sample order - producing the outputs
// straight into 32-byte stores.  FP64 work per complex sample at D = 1: 4K + 8K + 2.
//
// The input is FETCHED twice as well - except for what is still in the ring: a tile goes through phase 1, the scan and
// its look-back, and its boxes are then requested a second time for phase 3, a few microseconds after the first request
// (about half of those are L2 hits).  Shared memory therefore holds only a small ring of 4 KiB TMA boxes per warp, the
// run length is not tied to it (256 samples; 64 fitted when two whole tiles had to stay resident), and the per-tile
// work - warp scan, look-back, entry states, tickets - which was a third of the time at 64-sample runs is spread over
// four times as many samples; the look-back window shrinks with it (2 tiles instead of 6 for the narrowest reference
// filter) and CF32 input is no longer held to 7 warps by its 8-byte samples.  Phase 1 takes the boxes in the order
// keep .. N-1, 0 .. keep-1 so that the `keep` boxes the ring holds when it ends are the ones phase 3 starts with: those
// are not requested again (rerun_keep(); a tile that fits the ring is not fetched twice at all).
//
// Phase 3 follows the look-back at once (the first version ran it one tile late, like the staged kernel, to keep
// a warp from idling on a predecessor that is a little behind; the second fetch then came a whole tile later and
// missed L2 every time - ncu: 8.6 GB read for 4.3 GB of input).  A tile is ~35k cycles long now and its predecessor
// was claimed a ticket earlier: the aggregate it waits for is practically always there.
//
// Integer inputs of filters whose zeros all sit at z = -1 take the AP form of phase 3 (template parameter AP, see the
// phase-3 block): the numerator on the integer samples, an all-pole cascade on the FP64 pipe - 4K + 2 FMAs per
// complex sample instead of 8K + 2.
//
//   one tile of a warp:   phase 1 out of the ring (a slot whose box is done with goes to the next request of the
//                         tile's sequence: the rest of phase 1's order, then the second requests of boxes keep .. N-1)
//                         phase 2, aggregate out, look-back, entry states
//                         ticket for the next tile (its round trip hides behind phase 3)
//                         phase 3 out of the ring (after the tile's last second request: the next tile's first boxes)
#pragma once
#include "iir_kernel.cuh"

namespace cbl {

#ifndef CBL_RR_L2HINT
#define CBL_RR_L2HINT 0          // L2 eviction-priority hints on the two requests of a box (sweeps 29 and 40: no gain)
#endif
template <int K, int D, int IN, bool AP = false>
struct RerunCfg {
    static constexpr int S     = 2 * K;
    static constexpr int SPS   = (IN == IN_CF32) ? 16 : 32;   // samples per 128-byte row step = per box
    static constexpr bool WIDE = rerun_wide(K, AP);
    static constexpr int R     = rerun_ring(K, AP);           // ring slots (a tile of fewer boxes uses as many as it has)
    // the run length is a launch parameter (IirParams::rr_boxes boxes of SPS samples): the plan fixes it
    static constexpr int RING_BYTES    = R * 4096;
    static constexpr int SCRATCH_BYTES = 2 * S * 8 * 2;
    static constexpr int WARP_BYTES = RING_BYTES + SCRATCH_BYTES;
    static constexpr int SMEM_WARPS = (232448 - 1024) / (WARP_BYTES + R * 8);
    // registers (cuobjdump -res-usage): 12 warps allow 168 per thread
    static constexpr int REG_WARPS  = WIDE ? CBL_RR_AP_WARPS : ((K <= 2) ? CBL_RR_WARPS : (K == 3 ? CBL_RR_WARPS3 : CBL_RR_WARPS4));
    static constexpr int WPC = SMEM_WARPS < REG_WARPS ? SMEM_WARPS : REG_WARPS;
    static_assert(WPC >= 1, "a warp's buffers do not fit shared memory");
    static_assert(D == 1 || D == 2, "phase 3 works in groups of 8 or 16 samples: whole chunks of 4 kept outputs");
    static constexpr int SMEM_BYTES = WPC * WARP_BYTES + WPC * R * 8 + 1024;
};

// four consecutive outputs of one lane, already scaled (y = gain * y/G + out_bias), in the stream's output format
__device__ __forceinline__ void store_out4(const IirParams& P, uint64_t o, const double (&yi)[4], const double (&yq)[4])
{
    if (P.out_mode == OUT_F32) {
        float2* op = reinterpret_cast<float2*>(P.out) + o;
        if (P.out_vec) {
            stg_cf32x4(op, make_float2((float)yi[0], (float)yq[0]), make_float2((float)yi[1], (float)yq[1]),
                       make_float2((float)yi[2], (float)yq[2]), make_float2((float)yi[3], (float)yq[3]));
        } else {
#pragma unroll
            for (int i = 0; i < 4; i++) __stcs(op + i, make_float2((float)yi[i], (float)yq[i]));
        }
    } else if (P.out_mode == OUT_F64) {
        double2* op = reinterpret_cast<double2*>(P.out) + o;
        if (P.out_vec) {
            stg_f64x4(&op[0].x, yi[0], yq[0], yi[1], yq[1]);
            stg_f64x4(&op[2].x, yi[2], yq[2], yi[3], yq[3]);
        } else {
#pragma unroll
            for (int i = 0; i < 4; i++) __stcs(op + i, make_double2(yi[i], yq[i]));
        }
    } else {
        // (int16_t)(float)y, CaribouliteStream.cpp:295-296: round to float, truncate toward zero
        int qi[4], qq[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            qi[i] = (int)(int16_t)__float2int_rz((float)yi[i]);
            qq[i] = (int)(int16_t)__float2int_rz((float)yq[i]);
        }
        if (P.out_mode == OUT_Q16_CS16) {
            uint32_t w[4];
#pragma unroll
            for (int i = 0; i < 4; i++) w[i] = ((uint32_t)qi[i] & 0xFFFFu) | ((uint32_t)qq[i] << 16);
            uint32_t* op = reinterpret_cast<uint32_t*>(P.out) + o;
            if (P.out_vec) {
                __stcs(reinterpret_cast<uint4*>(op), make_uint4(w[0], w[1], w[2], w[3]));
            } else {
#pragma unroll
                for (int i = 0; i < 4; i++) op[i] = w[i];
            }
        } else if (P.out_mode == OUT_Q16_CF32) {
            float2* op = reinterpret_cast<float2*>(P.out) + o;
            const float k = 1.0f / 4096.0f;
            if (P.out_vec) {
                stg_cf32x4(op, make_float2((float)qi[0] * k, (float)qq[0] * k), make_float2((float)qi[1] * k, (float)qq[1] * k),
                           make_float2((float)qi[2] * k, (float)qq[2] * k), make_float2((float)qi[3] * k, (float)qq[3] * k));
            } else {
#pragma unroll
                for (int i = 0; i < 4; i++) __stcs(op + i, make_float2((float)qi[i] * k, (float)qq[i] * k));
            }
        } else if (P.out_mode == OUT_Q16_CF64) {
            double2* op = reinterpret_cast<double2*>(P.out) + o;
            const double k = 1.0 / 4096.0;
            if (P.out_vec) {
                stg_f64x4(&op[0].x, (double)qi[0] * k, (double)qq[0] * k, (double)qi[1] * k, (double)qq[1] * k);
                stg_f64x4(&op[2].x, (double)qi[2] * k, (double)qq[2] * k, (double)qi[3] * k, (double)qq[3] * k);
            } else {
#pragma unroll
                for (int i = 0; i < 4; i++) __stcs(op + i, make_double2((double)qi[i] * k, (double)qq[i] * k));
            }
        } else {   // OUT_Q16_CS8, CaribouliteStream.cpp:362-363
            uint32_t h[4];
#pragma unroll
            for (int i = 0; i < 4; i++) h[i] = ((uint32_t)(qi[i] >> 5) & 0xFFu) | (((uint32_t)(qq[i] >> 5) & 0xFFu) << 8);
            uint16_t* op = reinterpret_cast<uint16_t*>(P.out) + o;
            if (P.out_vec) {
                __stcs(reinterpret_cast<uint2*>(op), make_uint2(h[0] | (h[1] << 16), h[2] | (h[3] << 16)));
            } else {
#pragma unroll
                for (int i = 0; i < 4; i++) op[i] = (uint16_t)h[i];
            }
        }
    }
}

// AP: phase 3 in the "integer FIR + all-pole cascade" form (see the phase-3 block); integer inputs of filters whose
// zeros all sit at z = -1 (every Butterworth low-pass).
template <int K, int D, int IN, bool AP = false>
__global__ void __launch_bounds__((RerunCfg<K, D, IN, AP>::WPC) * 32, 1)
iir_rerun_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ IirParams P,
                 const __grid_constant__ IirTables T)
{
    using C = RerunCfg<K, D, IN, AP>;
    constexpr int S = C::S;
    constexpr bool PAR = true;
    constexpr bool INJECT = IN != IN_CF32;
    const int N = (int)P.rr_boxes;                              // boxes per tile
    const uint32_t R = N < C::R ? (uint32_t)N : (uint32_t)C::R;  // ring slots in use
    // The last `keep` boxes phase 1 looks at stay in the ring for phase 3, which starts with them: phase 1 takes the
    // boxes in the order keep .. N-1, 0 .. keep-1 (two parts, each from the steady state; their end vectors combine as
    // f = A^len(first part) f_second + f_first) and only boxes keep .. N-1 are fetched a second time.  keep = the ring
    // size; a tile that fits the ring (keep = N) is not fetched twice at all.
    // (the AP kernels only keep whole tiles: with their 16 warps and 128 registers the split costs what the saved
    // fetches bring - profiles/r2_sweeps.txt, sweep 42)
    const int keep = rerun_keep(N, (int)C::R, AP);
    constexpr bool SPLIT = !AP;                                  // whether phase 1 can come in two parts at all
    const int LEN = 2 * N - keep;                                // box requests of a tile
    // request idx of a tile's sequence -> box: phase 1's order, then the second requests
    auto seq_box = [&](int idx) { return idx < N ? (idx + keep < N ? idx + keep : idx + keep - N) : keep + (idx - N); };
    extern __shared__ uint8_t smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t ring_u32 = base + warp * C::RING_BYTES;
    const uint32_t scratch = base + C::WPC * C::RING_BYTES + warp * C::SCRATCH_BYTES;
    const uint32_t bar_u32 = base + C::WPC * C::WARP_BYTES + warp * C::R * 8;

    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < C::R; s++) mbar_init(bar_u32 + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();

    const uint32_t tps = P.tiles_per_stream;

    const uint32_t inj_xor = P.inj_xor, inj_mul = P.inj_mul;

    // box `b` of tile (stream, tis) into ring slot `slot` (lane 0 only)
    // `again`: the second request of a box (phase 3).  CBL_RR_L2HINT: the first request asks L2 to keep the lines
    // (evict_last), the second one releases them (evict_first)
    auto issue_box = [&](uint32_t stream, uint32_t tis, int b, uint32_t slot, bool again = false) {
        const uint32_t bar = bar_u32 + 8 * slot;
        mbar_expect_tx(bar, 4096);
#if CBL_RR_L2HINT
        tma_load_3d_hint(ring_u32 + slot * 4096, &tmap, b * 32, (int)tis * 32, (int)stream, bar, again ? kL2EvictFirst : kL2EvictLast);
#else
        (void)again;
        tma_load_3d(ring_u32 + slot * 4096, &tmap, b * 32, (int)tis * 32, (int)stream, bar);
#endif
    };

    // 16-byte piece c of this lane's row in box `box_u32`
    auto row_piece = [&](uint32_t box_u32, int c) {
        return lds128((box_u32 + lane * 128 + ((lane & 7) << 4)) ^ (uint32_t)(c << 4));
    };

    uint32_t slot = 0, pm = 0;     // ring slot of the next box to consume; bit s of pm = parity the pending request of slot s
                                   // completes its barrier with (flipped whenever a slot is handed to a new request)
    uint32_t tile = 0xFFFFFFFFu;
    uint32_t* const ticket = P.ticket + __ldg(P.ticket + 32 + lane);      // words 32..63 of the ticket block stay zero
    if (lane == 0) {
        tile = take_ticket(ticket);
    }
    tile = __shfl_sync(0xffffffffu, tile, 0);
    uint32_t stream = tile / tps, tis = tile - stream * tps;
    if (lane == 0 && tile < P.total_tiles) {
        for (uint32_t s = 0; s < R; s++) issue_box(stream, tis, seq_box((int)s), s);
    }

    while (tile < P.total_tiles) {
        // gated stream (its capture lost word alignment, api.cu): the tile drains the boxes already requested for
        // it and is otherwise skipped
        const bool live = P.gate == nullptr || __ldg(P.gate + (size_t)stream * P.gate_stride) == 0;
        uint32_t next = 0xFFFFFFFFu;       // meaningful in lane 0
        double f[2][S];

        // ---- phase 1: the run from the steady state, end state only
        {
            double va[K][2], vb[K][2];
#pragma unroll
            for (int k = 0; k < K; k++) {
                va[k][0] = vb[k][0] = T.sbias[2 * k];
                va[k][1] = vb[k][1] = T.sbias[2 * k + 1];
            }
            int li = (int)R;                 // next request of this tile's sequence (R are out when the tile starts)
            double fa[2][S];                 // end vector of the first part (boxes keep .. N-1)
#pragma unroll 1
            for (int j = 0; j < N; j++) {
                if (SPLIT && j == N - keep && j != 0) {
                    // second part (boxes 0 .. keep-1, the ones that stay): from the steady state again
#pragma unroll
                    for (int k = 0; k < K; k++) {
                        fa[0][2 * k] = va[k][0] - T.sbias[2 * k]; fa[0][2 * k + 1] = va[k][1] - T.sbias[2 * k + 1];
                        fa[1][2 * k] = vb[k][0] - T.sbias[2 * k]; fa[1][2 * k + 1] = vb[k][1] - T.sbias[2 * k + 1];
                        va[k][0] = vb[k][0] = T.sbias[2 * k];
                        va[k][1] = vb[k][1] = T.sbias[2 * k + 1];
                    }
                }
                const uint32_t st = slot;
                mbar_wait(bar_u32 + 8 * st, (pm >> st) & 1u);
                const uint32_t box = ring_u32 + st * 4096;
                // the 128-byte row in two groups of four pieces: the loops of a tile have to share the instruction
                // cache of the SM with eleven other warps in other phases
#pragma unroll 1
                for (int co = 0; co < (live ? 8 : 0); co += 4) {
                    uint4 q[4];
#pragma unroll
                    for (int ci = 0; ci < 4; ci++) q[ci] = row_piece(box, co + ci);
#pragma unroll
                    for (int ci = 0; ci < 4; ci++) {
                        const uint32_t wv[4] = {q[ci].x, q[ci].y, q[ci].z, q[ci].w};
#pragma unroll
                        for (int e = 0; e < (IN == IN_CF32 ? 2 : 4); e++) {
                            double xa, xb;
                            if (INJECT) inject_f64<IN>(wv[e], inj_xor, inj_mul, xa, xb);
                            else sample_to_f64<IN, false>(wv[2 * e], wv[2 * e + 1], xa, xb);
                            filter_step<K, PAR, false>(T, va, xa);
                            filter_step<K, PAR, false>(T, vb, xb);
                        }
                    }
                }
                __syncwarp();
                // a box of the first part is done with: its slot goes to the next request of the tile's sequence (the
                // rest of phase 1, then the second requests).  A gated tile has no phase 3: every box is done with and
                // only first requests go out.
                const bool frees = !live || j < N - keep;
                if (frees && li < (live ? LEN : N)) {
                    if (lane == 0) issue_box(stream, tis, seq_box(li), st, li >= N);
                    li++;
                    pm ^= 1u << st;
                }
                if (++slot == R) slot = 0;
            }
            if (SPLIT && keep > 0 && keep < N) {
                // f = A^(samples of the first part) f_second + f_first (the first part is the later one in time)
#pragma unroll
                for (int row = 0; row < S; row++) {
                    double ea = fa[0][row], eb = fa[1][row];
#pragma unroll
                    for (int col = col_lo<PAR>(row); col < col_hi(row); col++) {
                        const double m = T.R[kRrPowRow + row][col];
                        ea = fma(m, (col & 1 ? va[col / 2][1] : va[col / 2][0]) - T.sbias[col], ea);
                        eb = fma(m, (col & 1 ? vb[col / 2][1] : vb[col / 2][0]) - T.sbias[col], eb);
                    }
                    f[0][row] = ea; f[1][row] = eb;
                }
            } else {
#pragma unroll
                for (int k = 0; k < K; k++) {
                    f[0][2 * k] = va[k][0] - T.sbias[2 * k]; f[0][2 * k + 1] = va[k][1] - T.sbias[2 * k + 1];
                    f[1][2 * k] = vb[k][0] - T.sbias[2 * k]; f[1][2 * k + 1] = vb[k][1] - T.sbias[2 * k + 1];
                }
            }
        }

        if (!live) {
            // nothing of this tile's is in flight any more and every slot is free; the next tile's first requests go out
            // after its ticket
            if (lane == 0) next = take_ticket(ticket);
            const uint32_t ns_ = next / tps, nt_ = next - ns_ * tps;
            if (lane == 0 && next < P.total_tiles) {
                for (uint32_t s = 0; s < R; s++) issue_box(ns_, nt_, seq_box((int)s), (slot + s) % R);
            }
            pm ^= (1u << R) - 1u;
            tile = __shfl_sync(0xffffffffu, next, 0);
            stream = __shfl_sync(0xffffffffu, ns_, 0);
            tis = __shfl_sync(0xffffffffu, nt_, 0);
            continue;
        }

        // the first look-back descriptor: its L2 round trip hides behind the scan
        const int comp = lane < 2 * S ? lane : 0;
        unsigned long long pre = kDescEmpty;
        if (tis != 0)
            pre = ld_desc((P.chained ? P.desc_incl : P.desc_agg) + (size_t)(tile - 1) * (2 * S) + comp);

        // ---- phase 2: warp scan of run end states: E_t = sum_{j<=t} M^(t-j) f_j
#pragma unroll
        for (int j = 0; j < 5; j++) {
            if (j < P.scan_rounds) {
                const int d = 1 << j;
                double gq[2][S];
#pragma unroll
                for (int r = 0; r < 2; r++)
#pragma unroll
                    for (int c = 0; c < S; c++) gq[r][c] = shfl_up_f64(f[r][c], d);
                if (lane >= d) {
#pragma unroll
                    for (int r = 0; r < 2; r++)
#pragma unroll
                        for (int row = 0; row < S; row++) {
                            double acc = f[r][row];
#pragma unroll
                            for (int col = col_lo<PAR>(row); col < col_hi(row); col++) acc = fma(T.Mp[j][row * S + col], gq[r][col], acc);
                            f[r][row] = acc;
                        }
                }
            }
        }
        if (lane == 31 && !P.chained) {
            double* agg = P.desc_agg + (size_t)tile * (2 * S);
#pragma unroll
            for (int r = 0; r < 2; r++)
#pragma unroll
                for (int c = 0; c < S; c++) st_desc(agg + r * S + c, f[r][c]);
        }

        // ---- decoupled look-back: the state entering the tile
        {
            const int crail = comp / S, crow = comp % S;
            // Every lane walks the look-back (lanes >= 2S repeat component 0: same addresses, no extra traffic) and
            // only the store is predicated: with the walk under `if (lane < 2S)` and more than one trip through its
            // loop, the two groups of lanes stayed apart all the way through phase 3 - every instruction issued
            // twice, at 16 threads (ncu, order 6 at fc = 10 kHz: profiles/r1_sweeps.txt, sweep 26).
            double sin_c = 0.0;
            if (tis == 0) {
                sin_c = P.state_in[(size_t)stream * (2 * S) + crail * S + crow];
            } else {
                const double* first = (P.chained ? P.desc_incl : P.desc_agg) + (size_t)(tile - 1) * (2 * S) + comp;
                while (pre == kDescEmpty) { __nanosleep(32); pre = ld_desc(first); }
                sin_c = __longlong_as_double((long long)pre);
                if (!P.chained) {
                    const int depth = (int)tis < P.window ? (int)tis : P.window;
                    // two tiles' aggregates are requested together (one L2 round trip for both)
                    for (int j = 2; j <= depth; j += 2) {
                        const bool two = j + 1 <= depth;
                        const double* a = P.desc_agg + (size_t)(tile - j) * (2 * S) + crail * S;
                        const double* a2 = a - (two ? 2 * S : 0);
                        const double* mp = P.mtile_pow + (size_t)(j - 1) * (S * S) + crow * S;
                        unsigned long long dv[S], dw[S];
#pragma unroll
                        for (int col = 0; col < S; col++) { dv[col] = ld_desc(a + col); dw[col] = ld_desc(a2 + col); }
#pragma unroll
                        for (int col = 0; col < S; col++) {
                            while (dv[col] == kDescEmpty) { __nanosleep(32); dv[col] = ld_desc(a + col); }
                            sin_c = fma(__ldg(mp + col), __longlong_as_double((long long)dv[col]), sin_c);
                        }
                        if (two) {
#pragma unroll
                            for (int col = 0; col < S; col++) {
                                while (dw[col] == kDescEmpty) { __nanosleep(32); dw[col] = ld_desc(a2 + col); }
                                sin_c = fma(__ldg(mp + S * S + col), __longlong_as_double((long long)dw[col]), sin_c);
                            }
                        }
                    }
                    if ((int)tis <= P.window) {     // the window reaches the start of the stream
                        const double* a = P.state_in + (size_t)stream * (2 * S) + crail * S;
                        const double* mp = P.mtile_pow + (size_t)tis * (S * S) + crow * S;
#pragma unroll
                        for (int col = 0; col < S; col++) sin_c = fma(__ldg(mp + col), __ldg(a + col), sin_c);
                    }
                }
            }
            if (lane < 2 * S) sts64(scratch + lane * 8, sin_c);
        }
        __syncwarp();
        if (lane == 0) next = take_ticket(ticket);          // looked at towards the end of phase 3
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
            for (int col = col_lo<PAR>(row); col < col_hi(row); col++) {
                const double m = __ldg(P.plane + (row * S + col) * 33 + lane);      // (M^lane)[row][col]
                s0[0][row] = fma(m, sin[0][col], s0[0][row]);
                s0[1][row] = fma(m, sin[1][col], s0[1][row]);
            }

        // true end state of the tile (chained mode) / of the stream (last tile)
        const bool last_tile = (tis == tps - 1);
        if (P.chained || (last_tile && P.write_state)) {
            const int lastl = last_tile && P.write_state ? (int)P.last_tile_runs - 1 : 31;
            if (lane == lastl) {
                double* incl = P.desc_incl + (size_t)tile * (2 * S);
#pragma unroll
                for (int row = 0; row < S; row++) {
                    double ea = f[0][row], eb = f[1][row];
#pragma unroll
                    for (int col = col_lo<PAR>(row); col < col_hi(row); col++) {
                        const double m = __ldg(P.plane + (row * S + col) * 33 + lastl + 1);
                        ea = fma(m, sin[0][col], ea);
                        eb = fma(m, sin[1][col], eb);
                    }
                    if (P.chained) { st_desc(incl + row, ea); st_desc(incl + S + row, eb); }
                    if (last_tile && P.write_state) {
                        P.state_out[(size_t)stream * (2 * S) + row] = ea;
                        P.state_out[(size_t)stream * (2 * S) + S + row] = eb;
                    }
                }
            }
        }

        // ---- phase 3: the run again, box by box, from its true entry state (offset by the steady state the
        // injected inputs imply), outputs at the kept positions
        uint32_t nstream = 0, ntis = 0;
        {
            const uint32_t run = tis * 32u + lane;
            const bool mine = run < P.n_runs;
            double va[K][2], vb[K][2];
            // AP form.  H(q) = G (1 + q)^2K / prod_k (1 + a1_k q + a2_k q^2): the numerator is applied to the INTEGER
            // samples - a cascade of 2K two-tap sums, exact; for SMI words the first three stages act on both 13-bit
            // fields at once in the 16-bit halves of one register - the result is dropped into a double's mantissa
            // (the last sum also adds the exponent bits, one IMAD.WIDE spreads it over the register pair), and the
            // all-pole cascade that is left costs 2 FMAs per section, rail and sample with the output for free:
            // 4K + 2 FMAs per complex sample instead of the parallel form's 8K + 2.  The run starts from "all earlier
            // samples were zero" (histories preset to the offset-binary zero) and the all-pole state
            // sigma0 = bias + Tap s0, Tap = the map between the two realisations' zero-input responses (tables.cpp).
            constexpr bool APW = IN != IN_CS16;                         // SMI words: 13-bit fields, packed stages
            constexpr int  SG = 2 * K;                                  // two-tap stages
            constexpr int  NPK = AP ? (APW ? (SG < 3 ? SG : 3) : 0) : 0;  // stages on the packed halves
            constexpr int  NUP = AP ? SG - NPK : 0;                     // stages per rail after unpacking
            constexpr uint32_t BETA = APW ? 4096u : 32768u;             // offset-binary zero
            uint32_t hp[NPK > 0 ? NPK : 1], ga[NUP > 0 ? NUP : 1], gb[NUP > 0 ? NUP : 1];
            const uint32_t ap_exp = P.ap_exp, ap_mul = P.ap_mul, ap_xor = P.ap_xor;
            if constexpr (AP) {
#pragma unroll
                for (int i = 0; i < NPK; i++) hp[i] = (BETA << i) * 0x00010001u;
#pragma unroll
                for (int i = 0; i < NUP; i++) ga[i] = gb[i] = BETA << (NPK + i);
#pragma unroll
                for (int row = 0; row < S; row++) {
                    double ea = T.R[kApBiasRow][row], eb = ea;
#pragma unroll
                    for (int col = 0; col < S; col++) {
                        ea = fma(T.R[row][col], s0[0][col], ea);
                        eb = fma(T.R[row][col], s0[1][col], eb);
                    }
                    va[row / 2][row % 2] = ea;
                    vb[row / 2][row % 2] = eb;
                }
            } else {
#pragma unroll
                for (int k = 0; k < K; k++) {
                    va[k][0] = T.sbias[2 * k] + s0[0][2 * k];     va[k][1] = T.sbias[2 * k + 1] + s0[0][2 * k + 1];
                    vb[k][0] = T.sbias[2 * k] + s0[1][2 * k];     vb[k][1] = T.sbias[2 * k + 1] + s0[1][2 * k + 1];
                }
            }
            (void)hp; (void)ga; (void)gb; (void)ap_exp; (void)ap_mul; (void)ap_xor;
            uint64_t o = (uint64_t)stream * P.out_stride + (uint64_t)run * (uint32_t)(N * C::SPS / D);
            constexpr int PER = (IN == IN_CF32) ? 2 : 4;      // samples per 16-byte piece
            constexpr int PP  = 4 * D / PER;                  // pieces that hold four kept outputs = one store
            constexpr int CPB = 8 / PP;                       // chunks of four kept outputs per box
            static_assert(PP >= 1 && 8 % PP == 0, "a box holds whole chunks");
            const bool fast_f32 = P.out_mode == OUT_F32 && P.out_vec != 0;
            int li = (int)R + N - keep;       // where phase 1 left the tile's request sequence
#pragma unroll 1
            for (int b = 0; b < N; b++) {
                const uint32_t st = slot;
                mbar_wait(bar_u32 + 8 * st, (pm >> st) & 1u);      // (a box that stayed: completed long ago)
                const uint32_t box = ring_u32 + st * 4096;
                if (mine) {
                    // chunks of four kept outputs, kRerunP3Unroll per iteration (one copy of the format switch in
                    // store_out4 each); the next chunk's pieces are read while this one is filtered
                    uint4 qn[PP];
#pragma unroll
                    for (int i = 0; i < PP; i++) qn[i] = row_piece(box, i);
#pragma unroll (kRerunP3Unroll)
                    for (int ch = 0; ch < CPB; ch++) {
                        uint4 q[PP];
#pragma unroll
                        for (int i = 0; i < PP; i++) q[i] = qn[i];
                        if (ch + 1 < CPB) {
#pragma unroll
                            for (int i = 0; i < PP; i++) qn[i] = row_piece(box, (ch + 1) * PP + i);
                        }
                        double yi[4], yq[4];
#pragma unroll
                        for (int ci = 0; ci < PP; ci++) {
                            const uint32_t wv[4] = {q[ci].x, q[ci].y, q[ci].z, q[ci].w};
#pragma unroll
                            for (int e = 0; e < PER; e++) {
                                const int j = ci * PER + e;          // sample within the chunk (chunks start on multiples of D)
                                [[maybe_unused]] double xa, xb;
                                if constexpr (AP) {
                                    uint32_t ta, tb;
                                    if constexpr (APW) {
                                        uint32_t pk = and_xor<0x1FFF1FFFu>(wv[e] >> 1, ap_xor);   // both fields, offset binary
#pragma unroll
                                        for (int i = 0; i < NPK; i++) { const uint32_t t = pk + hp[i]; hp[i] = pk; pk = t; }
                                        const uint32_t lo = pk & 0xFFFFu, hi = pk >> 16;
                                        ta = IN == IN_WORDS ? hi : lo;       // rail A = I: bits 29:17 (S1G) / 13:1 (HiF)
                                        tb = IN == IN_WORDS ? lo : hi;
                                    } else {
                                        const uint32_t u = wv[e] ^ ap_xor;   // 0x80008000: both int16 halves, offset binary
                                        ta = u & 0xFFFFu;
                                        tb = u >> 16;
                                    }
#pragma unroll
                                    for (int i = 0; i < NUP; i++) {
                                        const uint32_t xe = i == NUP - 1 ? ap_exp : 0u;
                                        const uint32_t t = ta + ga[i] + xe, u = tb + gb[i] + xe;
                                        ga[i] = ta; gb[i] = tb;
                                        ta = t; tb = u;
                                    }
                                    if (NUP == 0) { ta += ap_exp; tb += ap_exp; }
                                    double ia = mul_wide_f64(ta, ap_mul), ib = mul_wide_f64(tb, ap_mul);
#pragma unroll
                                    for (int k = 0; k < K; k++) {
                                        const double wa = fma(-T.a1[k], va[k][0], fma(-T.a2[k], va[k][1], ia));
                                        const double wb = fma(-T.a1[k], vb[k][0], fma(-T.a2[k], vb[k][1], ib));
                                        va[k][1] = va[k][0]; va[k][0] = wa; ia = wa;
                                        vb[k][1] = vb[k][0]; vb[k][0] = wb; ib = wb;
                                    }
                                    if (j % D == 0) {
                                        yi[j / D] = fma(ia, T.gain, T.out_bias);
                                        yq[j / D] = fma(ib, T.gain, T.out_bias);
                                    }
                                } else {
                                    if (INJECT) inject_f64<IN>(wv[e], inj_xor, inj_mul, xa, xb);
                                    else sample_to_f64<IN, false>(wv[2 * e], wv[2 * e + 1], xa, xb);
                                    if (j % D == 0) {
                                        yi[j / D] = fma(filter_step<K, PAR, true>(T, va, xa), T.gain, T.out_bias);
                                        yq[j / D] = fma(filter_step<K, PAR, true>(T, vb, xb), T.gain, T.out_bias);
                                    } else {
                                        filter_step<K, PAR, false>(T, va, xa);
                                        filter_step<K, PAR, false>(T, vb, xb);
                                    }
                                }
                            }
                        }
                        if (fast_f32) {      // plain CF32 into an aligned buffer: one test instead of the format switch
                            stg_cf32x4(reinterpret_cast<float2*>(P.out) + o, make_float2((float)yi[0], (float)yq[0]),
                                       make_float2((float)yi[1], (float)yq[1]), make_float2((float)yi[2], (float)yq[2]),
                                       make_float2((float)yi[3], (float)yq[3]));
                        } else {
                            store_out4(P, o, yi, yq);
                        }
                        o += 4;
                    }
                }
                __syncwarp();
                // the slot goes to the next request: the rest of this tile's second requests, then the first requests
                // of the next tile
                if (lane == 0) {
                    if (li < LEN) {
                        issue_box(stream, tis, seq_box(li), st, true);
                    } else {
                        nstream = next / tps;
                        ntis = next - nstream * tps;
                        if (next < P.total_tiles) issue_box(nstream, ntis, seq_box(li - LEN), st);
                    }
                }
                li++;
                pm ^= 1u << st;
                if (++slot == R) slot = 0;
            }
        }
        tile = __shfl_sync(0xffffffffu, next, 0);
        stream = __shfl_sync(0xffffffffu, nstream, 0);
        tis = __shfl_sync(0xffffffffu, ntis, 0);
    }
}

template <int K, int D, int IN, bool AP = false>
int launch_rerun_t(const CUtensorMap& tmap, const IirParams& p, const IirTables& t, int num_sms, cudaStream_t st)
{
    using C = RerunCfg<K, D, IN, AP>;
    auto kern = iir_rerun_kernel<K, D, IN, AP>;
    static bool ready_dev[kMaxDevices] = {false};      // per device: the shared-memory opt-in is a per-device attribute
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= kMaxDevices) return (int)cudaErrorInvalidDevice;
    if (!ready_dev[dev]) {
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES);
        if (e != cudaSuccess) return (int)e;
        ready_dev[dev] = true;
    }
    const uint64_t want = ((uint64_t)p.total_tiles + C::WPC - 1) / C::WPC;
    const int grid = (int)(want < (uint64_t)num_sms ? want : (uint64_t)num_sms);
    kern<<<grid, C::WPC * 32, C::SMEM_BYTES, st>>>(tmap, p, t);
    return (int)cudaGetLastError();
}

template <int KK, int D>
iir_launch_fn rerun_pick_kind(int in_kind)
{
    if constexpr (iir_is_rerun(KK, D, true)) {
        if (in_kind == IN_WORDS) return &launch_rerun_t<KK, D, IN_WORDS>;
        if (in_kind == IN_WORDS_HIF) return &launch_rerun_t<KK, D, IN_WORDS_HIF>;
        if (in_kind == IN_CS16) return &launch_rerun_t<KK, D, IN_CS16>;
        if (in_kind == IN_CF32) return &launch_rerun_t<KK, D, IN_CF32>;
    }
    return nullptr;
}

template <int KK>
iir_launch_fn rerun_pick_launch(int D, int in_kind)
{
    return D == 1 ? rerun_pick_kind<KK, 1>(in_kind) : (D == 2 ? rerun_pick_kind<KK, 2>(in_kind) : nullptr);
}

// the AP form of phase 3 (integer inputs only)
template <int KK, int D>
iir_launch_fn rerun_ap_pick_kind(int in_kind)
{
    if constexpr (iir_is_rerun(KK, D, true)) {
        if (in_kind == IN_WORDS) return &launch_rerun_t<KK, D, IN_WORDS, true>;
        if (in_kind == IN_WORDS_HIF) return &launch_rerun_t<KK, D, IN_WORDS_HIF, true>;
        if (in_kind == IN_CS16) return &launch_rerun_t<KK, D, IN_CS16, true>;
    }
    return nullptr;
}
template <int KK>
iir_launch_fn rerun_ap_pick_launch(int D, int in_kind)
{
    return D == 1 ? rerun_ap_pick_kind<KK, 1>(in_kind) : (D == 2 ? rerun_ap_pick_kind<KK, 2>(in_kind) : nullptr);
}

}  // namespace cbl
