// Kernel (b) template.  See iir.cuh for the math.  Included by iir_k{1,2,3,4}.cu.
//
// Shape of one warp's life (persistent; 32 lanes = 32 consecutive runs of LT samples = one tile):
//
//   claim a tile (ticket) ---------------------------------------------.
//   phase 1  every lane filters its run from the steady state s*,       |  the NEXT tile is claimed at the top
//            fed by a per-warp ring of 4 KiB TMA boxes (128B swizzle:   |  of phase 1 and its first boxes are
//            lane = row, bank-conflict-free 16-byte reads); zero-state  |  requested while phase 1 drains, so
//            outputs at the kept (decimated) positions go to smem       |  the ring never runs dry between tiles
//   phase 2  warp scan of the run end states with the constant powers   |
//            M^(2^j) (shuffles); lane 31 publishes the tile aggregate --'
//   look-back  state entering the tile from the predecessors' aggregates (window bounded by the filter's
//            contraction) or, without contraction, from the predecessor's inclusive state.  Descriptor
//            doubles validate themselves (kDescEmpty until written): no flag, no fence.
//   phase 3  y[m] = gain * (y_zs[m] + R[m] . s0) + out_bias, converted and stored.
//
// Look-back and phase 3 of a tile run one tile LATE (after phases 1-2 of the warp's next tile, from the
// other half of a double-buffered output stage): a warp never idles on a predecessor that is a little
// behind, which otherwise synchronises all warps of the GPU into one convoy.
#pragma once
#include "iir.cuh"
#ifndef CBL_P1_UNROLL
#define CBL_P1_UNROLL 8
#endif
#ifndef CBL_P3_UNROLL
#define CBL_P3_UNROLL 5          // float-staged phase 3: up to this many chunks of four kept outputs are unrolled completely
#endif                           // (R[m][c] become constant-bank operands of the DFMAs instead of LDC.64 loads: C4 1.054 -> 1.027 ms)
namespace cbl { constexpr int kP1Unroll = CBL_P1_UNROLL; constexpr int kP3Unroll = CBL_P3_UNROLL; }

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
// the same with an L2 eviction-priority hint (the rerun kernel fetches every tile twice: the first fetch asks L2 to keep
// the lines, the second one releases them).  The policy words are the fixed encodings createpolicy.fractional
// produces for a fraction of 1.0 (the values CUTLASS's CacheHintSm90 carries).
constexpr unsigned long long kL2EvictFirst = 0x12F0000000000000ull;
constexpr unsigned long long kL2EvictLast  = 0x14F0000000000000ull;
__device__ __forceinline__ void tma_load_3d_hint(uint32_t dst, const CUtensorMap* tmap, int x, int y, int z, uint32_t bar,
                                                 unsigned long long policy)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%2, %3, %4}], [%5], %6;"
        ::"r"(dst), "l"(tmap), "r"(x), "r"(y), "r"(z), "r"(bar), "l"(policy) : "memory");
}
// look-back descriptor slots: relaxed, L2-coherent 8-byte accesses; the value is its own "ready" flag
__device__ __forceinline__ unsigned long long ld_desc(const double* p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_desc(double* p, double x)
{
    unsigned long long v = (unsigned long long)__double_as_longlong(x);
    if (v == kDescEmpty) v = kDescEmpty - 1;      // still a NaN, no longer "empty"
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// ticket: plain PTX atomic.  atomicAdd() - and a PTX atom on an address ptxas can prove uniform - is compiled into
// a warp-aggregated sequence whose shuffle waits for the atomic's round trip on the spot; the kernel therefore
// adds a zero it loaded from memory to the ticket address, and the value is not looked at until one tile later.
__device__ __forceinline__ uint32_t take_ticket(uint32_t* p)
{
    uint32_t v;
    asm volatile("atom.relaxed.gpu.global.add.u32 %0, [%1], 1;" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ double wait_desc(const double* p)
{
    unsigned long long v = ld_desc(p);
    while (v == kDescEmpty) { __nanosleep(32); v = ld_desc(p); }
    return __longlong_as_double((long long)v);
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
__device__ __forceinline__ void sts_f32x2(uint32_t a, float x, float y)
{
    asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(a), "f"(x), "f"(y) : "memory");
}
__device__ __forceinline__ void sts_u32x2(uint32_t a, uint32_t x, uint32_t y)
{
    asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(a), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ uint2 lds_u32x2(uint32_t a)
{
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ float2 lds_f32x2(uint32_t a)
{
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a) : "memory");
    return v;
}
// 32-byte streaming store: four consecutive CF32 outputs of one lane = one full sector (STG.256).  A warp's
// outputs are NK * 8 bytes apart from lane to lane, so 8-byte stores touch each sector four times.
__device__ __forceinline__ void stg_cf32x4(float2* p, float2 a, float2 b, float2 c, float2 d)
{
    asm volatile("st.global.cs.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 ::"l"(p), "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y), "f"(d.x), "f"(d.y) : "memory");
}
__device__ __forceinline__ void stg_f64x4(double* p, double a, double b, double c, double d)
{
    asm volatile("st.global.cs.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}
__device__ __forceinline__ double shfl_up_f64(double v, int d)
{
    return __shfl_up_sync(0xffffffffu, v, d);
}

// ------------------------------------------------------------------ sample -> (rail A, rail B) doubles
// rail A = hi13 / int16 low half / float re, rail B = lo13 / int16 high half / float im.
// Injection: integer fields become doubles offset by kBiasWords / kBiasCs16 with two integer ops each.
// One LOP3 isolates the field, flips its sign bit (offset binary u = v + 2^(bits-1)) and ORs in the
// exponent bits shifted left by one; one IMAD.WIDE by 2^31 then spreads that word over a register pair
// whose high half is (exponent | u) and whose low half is (the last field bit, zeros): the double
// 2^20 + u (SMI words) or 2^21 + u (CS16).  The second field of the word is shifted into place first.
template <uint32_t MASK>
__device__ __forceinline__ uint32_t and_xor(uint32_t a, uint32_t c)     // (a & MASK) ^ c, c in a register
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0x6A;" : "=r"(d) : "r"(a), "n"(MASK), "r"(c));
    return d;
}
// the multiplier comes in a register (kernel parameter): an immediate power of two is strength-reduced
// by ptxas into shifts plus moves
__device__ __forceinline__ double mul_wide_f64(uint32_t a, uint32_t mul)
{
    unsigned long long r;
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(r) : "r"(a), "r"(mul));
    return __longlong_as_double((long long)r);
}
template <int IN>
__device__ __forceinline__ void inject_f64(uint32_t w, uint32_t xc, uint32_t m31, double& xa, double& xb)
{
    if (IN == IN_WORDS) {      // S1G: I = bits 29:17, Q = bits 13:1
        xa = mul_wide_f64(and_xor<0x00003FFEu>(w >> 16, xc), m31);
        xb = mul_wide_f64(and_xor<0x00003FFEu>(w, xc), m31);
    } else if (IN == IN_WORDS_HIF) {   // HiF: I = bits 13:1, Q = bits 29:17 (caribou_smi.c:371-373)
        xa = mul_wide_f64(and_xor<0x00003FFEu>(w, xc), m31);
        xb = mul_wide_f64(and_xor<0x00003FFEu>(w >> 16, xc), m31);
    } else {                   // rail A: bits 15:0, rail B: bits 31:16
        xa = mul_wide_f64(and_xor<0x0000FFFFu>(w, xc), m31);
        xb = mul_wide_f64(and_xor<0x0000FFFFu>(w >> 16, xc), m31);
    }
}

template <int IN, bool INJECT>
__device__ __forceinline__ void sample_to_f64(uint32_t w0, uint32_t w1, double& xa, double& xb)
{
    if (IN == IN_WORDS) {
        xa = (double)hi13(w0);
        xb = (double)lo13(w0);
    } else if (IN == IN_WORDS_HIF) {
        xa = (double)lo13(w0);
        xb = (double)hi13(w0);
    } else if (IN == IN_CS16) {
        xa = (double)(int)(int16_t)(w0 & 0xFFFFu);
        xb = (double)((int32_t)w0 >> 16);
    } else {
        xa = (double)__uint_as_float(w0);
        xb = (double)__uint_as_float(w1);
    }
}

// Y32: the zero-state outputs are staged as (already scaled) floats instead of doubles - half the shared memory
// per warp, which buys occupancy; used when the output itself is CF32 without the int16 truncation.
// One CTA per SM holding as many warps as shared memory and the register file allow: the same number of warps
// in six 2-warp CTAs is 7 % slower (profiles/r1_sweeps.txt, sweep 18).  The register file is four banks of 16K,
// one per SMSP, and warp w of a CTA lives on SMSP w % 4: up to 8 warps (2 per SMSP) a thread may use 255
// registers, from 9 on (3 on some SMSP) 168.  Registers per thread (cuobjdump -res-usage): parallel form
// 120-168 for K <= 2 with float staging, up to 190 with double staging, ~205 (K = 3), ~250 (K = 4); the cascade
// form of K > 1 and the D = 1 kernels (32 kept outputs per run) 200-250.
__host__ __device__ constexpr int iir_reg_warps(int K, int D, bool par, bool y32)
{
    return (K == 3 && par && y32 && D != 1) ? CBL_IIR_K3_WARPS
         : (((!par && K > 1) || K >= 3 || D == 1) ? 8 : ((K == 2 && !y32) ? 8 : 12));
}
#ifndef CBL_IIR_WPC_CAP
#define CBL_IIR_WPC_CAP 16
#endif
template <int K, int D, int IN, bool PAR, bool Y32>
struct IirCfg {
    static constexpr int S     = 2 * K;                       // state dimension per rail
    static constexpr int LT    = iir_run_len(K, D, Y32);         // samples per run
    static constexpr int NK    = LT / D;                      // kept outputs per run
    static constexpr int SPS   = (IN == IN_CF32) ? 16 : 32;   // samples per 128-byte row step
    static constexpr int NSTEP = LT / SPS;
    static constexpr int RSTG  = Y32 ? kIirRingStages32 : kIirRingStages;
    static constexpr int STAGES = (NSTEP < RSTG) ? NSTEP : RSTG;
    static constexpr int RING_BYTES    = STAGES * 4096;
    static constexpr int YSLOT_BYTES   = Y32 ? 256 : 512;      // one kept output of all 32 lanes: float2 / 2 doubles each
    static constexpr int YSTAGE_BYTES  = NK * YSLOT_BYTES;     // one tile's zero-state outputs; kIirStageBufs of them per warp
    static constexpr int SCRATCH_BYTES = 2 * S * 8 * 2;        // S_in broadcast + spare
    static constexpr int P3U = (NK / 4 <= kP3Unroll) ? (NK / 4 > 0 ? NK / 4 : 1) : 1;   // phase-3 chunks unrolled (all or none)
    static constexpr int WARP_BYTES = RING_BYTES + kIirStageBufs * YSTAGE_BYTES + SCRATCH_BYTES;
    // 227 KiB per CTA minus the alignment slack of the carve-up
    static constexpr int SMEM_WARPS = (232448 - 1024) / (WARP_BYTES + STAGES * 8);
    static constexpr int REG_WARPS  = iir_reg_warps(K, D, PAR, Y32);
    static constexpr int WPC_ = SMEM_WARPS < REG_WARPS ? SMEM_WARPS : REG_WARPS;
    static constexpr int WPC  = WPC_ < CBL_IIR_WPC_CAP ? WPC_ : CBL_IIR_WPC_CAP;     // warps per CTA = per SM
    static_assert(WPC >= 1, "a warp's buffers do not fit shared memory");
    static constexpr int SMEM_BYTES = WPC * WARP_BYTES + WPC * STAGES * 8 + 1024;   // + mbarriers + alignment slack
};

// One filter step on one rail.  v = (v1,v2) per section.  Returns y/G at this sample if NEED_OUT.
template <int K, bool PAR, bool NEED_OUT>
__device__ __forceinline__ double filter_step(const IirTables& T, double (&v)[K][2], double x)
{
    if (PAR) {
        double out = x;
#pragma unroll
        for (int k = 0; k < K; k++) {
            if (NEED_OUT) out = fma(T.c2[k], v[k][1], fma(T.c1[k], v[k][0], out));
            const double w = fma(-T.a1[k], v[k][0], fma(-T.a2[k], v[k][1], x));
            v[k][1] = v[k][0];
            v[k][0] = w;
        }
        return out;
    } else {
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
}

// non-zero columns of row `row` of A^n: the section's own 2x2 block (parallel) / everything up to and
// including it (cascade: block lower-triangular)
template <bool PAR> __device__ __forceinline__ constexpr int col_lo(int row) { return PAR ? 2 * (row / 2) : 0; }
__device__ __forceinline__ constexpr int col_hi(int row) { return 2 * (row / 2 + 1); }

template <int K, int D, int IN, bool PAR, bool Y32>
__global__ void __launch_bounds__((IirCfg<K, D, IN, PAR, Y32>::WPC) * 32, 1)
iir_scan_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ IirParams P,
                const __grid_constant__ IirTables T)
{
    using C = IirCfg<K, D, IN, PAR, Y32>;
    constexpr int S = C::S;
    constexpr bool INJECT = PAR && IN != IN_CF32;
    constexpr bool FIXED = IN != IN_CF32;      // float-staged kernels fed with integers stage fixed-point words
    extern __shared__ uint8_t smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // shared-memory carve-up, as 32-bit shared-window addresses (1024-byte aligned for the 128B swizzle)
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t ring_u32 = base + warp * C::RING_BYTES;
    const uint32_t ystage0 = base + C::WPC * C::RING_BYTES + warp * (kIirStageBufs * C::YSTAGE_BYTES);
    const uint32_t scratch = base + C::WPC * (C::RING_BYTES + kIirStageBufs * C::YSTAGE_BYTES) + warp * C::SCRATCH_BYTES;
    const uint32_t bar_u32 = base + C::WPC * C::WARP_BYTES + warp * C::STAGES * 8;

    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < C::STAGES; s++) mbar_init(bar_u32 + 8 * s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();

    constexpr int swap = 0;     // rail A is I, rail B is Q for every input kind (the HiF layout is its own instantiation)
    const uint32_t tps = P.tiles_per_stream;

    // this lane's M^lane, non-zero pattern only (s0 = M^lane S_in + E_{lane-1})
    constexpr int NPL = PAR ? 2 : S;
    double pl[S][NPL];
#pragma unroll
    for (int row = 0; row < S; row++)
#pragma unroll
        for (int col = col_lo<PAR>(row); col < col_hi(row); col++)
            pl[row][col - col_lo<PAR>(row)] = __ldg(P.plane + (row * S + col) * 33 + lane);

    // mantissa-injection constants.  They come from the kernel parameters because ptxas folds anything it can
    // see through, and then splits the LOP3 and the IMAD.WIDE into longer immediate-friendly sequences.
    const uint32_t inj_xor = P.inj_xor, inj_mul = P.inj_mul;

    // request box `lstep` of tile (stream, tis) into ring slot `slot` (lane 0 only)
    auto issue = [&](uint32_t stream, uint32_t tis, int lstep, uint32_t slot) {
        mbar_expect_tx(bar_u32 + 8 * slot, 4096);
        tma_load_3d(ring_u32 + slot * 4096, &tmap, lstep * 32, (int)tis * 32, (int)stream, bar_u32 + 8 * slot);
    };

    // Tickets run two tiles ahead: `tile` is being filtered, `next` is already known (its first boxes are
    // requested while `tile` drains), and the atomic for the tile after that is in flight meanwhile, so
    // its round trip never sits on the critical path.  next / next2 are meaningful in lane 0 only.
    uint32_t gstep = 0;    // steps consumed by this warp so far (ring position / parity)
    uint32_t tile = 0xFFFFFFFFu, next = 0xFFFFFFFFu;
    uint32_t* const ticket = P.ticket + __ldg(P.ticket + 32 + lane);      // words 32..63 of the ticket block stay zero
    if (lane == 0) {
        tile = take_ticket(ticket);
        if (kIirDefer) next = take_ticket(ticket);      // (non-deferred form: one ticket at a time, see the loop)
    }
    tile = __shfl_sync(0xffffffffu, tile, 0);
    uint32_t stream = tile / tps, tis = tile - stream * tps;
    if (lane == 0 && tile < P.total_tiles) {
#pragma unroll
        for (int s = 0; s < C::STAGES; s++) issue(stream, tis, s, (uint32_t)s);
    }

    // the tile whose look-back + phase 3 are still owed (one behind the tile in phases 1-2)
    bool owe = false;
    uint32_t ptile = 0, pstream = 0, ptis = 0;
    double pf[2][S];       // its inclusive scan values E_lane
    uint32_t buf = 0;      // half of the output stage the CURRENT tile writes
#pragma unroll
    for (int r = 0; r < 2; r++)
#pragma unroll
        for (int c = 0; c < S; c++) pf[r][c] = 0.0;

    for (;;) {
        const bool have = tile < P.total_tiles;
        // a stream whose capture lost word alignment is left to the slow path (api.cu): its tiles still drain the
        // ring (the boxes were requested before anybody could know) but are neither filtered nor stored
        const bool live = have && (P.gate == nullptr || __ldg(P.gate + (size_t)stream * P.gate_stride) == 0);
        uint32_t next2 = 0xFFFFFFFFu, nstream = 0, ntis = 0;
        double f[2][S];
        double va[K][2], vb[K][2];
        if (have) {
            // deferred form: the tile after `next`, consumed one iteration later.  Non-deferred form: `next` itself,
            // first looked at when phase 1 requests the next tile's boxes - a warp that held two consecutive tiles
            // would process them a whole iteration apart, and the neighbour that needs the second one's aggregate
            // for its look-back would sit waiting for it
            if (lane == 0) { if (kIirDefer) next2 = take_ticket(ticket); else next = take_ticket(ticket); }
            if (kIirDefer) {
                nstream = next / tps;
                ntis = next - nstream * tps;
            }
            const uint32_t ystage = ystage0 + buf * C::YSTAGE_BYTES;

            // ---- phase 1: zero-state pass over the run (started at s*, see iir.cuh)
#pragma unroll
            for (int k = 0; k < K; k++) {
                va[k][0] = vb[k][0] = T.sbias[2 * k];
                va[k][1] = vb[k][1] = T.sbias[2 * k + 1];
            }
#pragma unroll 1
            for (int step = 0; step < C::NSTEP; step++) {
                const uint32_t st = gstep % C::STAGES;
                mbar_wait(bar_u32 + 8 * st, (gstep / C::STAGES) & 1u);
                const uint32_t row = ring_u32 + st * 4096 + lane * 128;
                const uint32_t rowx = row + ((lane & 7) << 4);       // bits 6:4 of row are zero (128-byte rows)
                // the 128-byte row in groups of U 16-byte pieces: the group body is unrolled, the group loop is
                // not (the whole row unrolled is ~9 KiB of code per kernel, which the instruction cache feels)
                constexpr int U = kP1Unroll;                     // 16-byte pieces per group
                constexpr int PER = (IN == IN_CF32) ? 2 : 4;     // samples per 16 bytes
                constexpr int GS = U * PER;                      // samples per group
                const uint32_t ylane = ystage + lane * 8;
#pragma unroll 1
                for (int co = 0; co < (live ? 8 / U : 0); co++) {
                    const int gbase = step * C::SPS + co * GS;   // first sample of the group within the run
                    uint4 q[U];
#pragma unroll
                    for (int ci = 0; ci < U; ci++) q[ci] = lds128(rowx ^ (uint32_t)((co * U + ci) << 4));   // = row + ((c ^ (lane & 7)) << 4): one XOR per piece
#pragma unroll
                    for (int ci = 0; ci < U; ci++) {
                        const uint32_t wv[4] = {q[ci].x, q[ci].y, q[ci].z, q[ci].w};
#pragma unroll
                        for (int e = 0; e < PER; e++) {
                            const int j = ci * PER + e;              // sample index within the group
                            double xa, xb;
                            if (INJECT) {
                                inject_f64<IN>(wv[e], inj_xor, inj_mul, xa, xb);
                            } else if (IN == IN_CF32) {
                                sample_to_f64<IN, false>(wv[2 * e], wv[2 * e + 1], xa, xb);
                            } else {
                                sample_to_f64<IN, false>(wv[e], 0u, xa, xb);
                            }
                            // kept iff (gbase + j) % D == 0; gbase is a multiple of GS
                            const bool kept_static = (D <= GS) ? (j % D == 0) : (j == 0);
                            if (kept_static) {
                                const bool kept = (D <= GS) ? true : ((gbase % D) == 0);
                                const double ya = filter_step<K, PAR, true>(T, va, xa);
                                const double yb = filter_step<K, PAR, true>(T, vb, xb);
                                if (kept) {
                                    // slot m = (gbase + j) / D: 2 rails x 32 lanes x 8 bytes, or 32 lanes x float2
                                    const uint32_t ya_addr = ylane + (uint32_t)(gbase / D) * C::YSLOT_BYTES + (uint32_t)((D <= GS) ? (j / D) : 0) * C::YSLOT_BYTES;
                                    if (Y32 && FIXED) {
                                        // T.out_bias carries stage_mo here: the low words are the staged fixed-point values
                                        sts_u32x2(ya_addr, (uint32_t)__double2loint(fma(ya, T.gain, T.out_bias)),
                                                  (uint32_t)__double2loint(fma(yb, T.gain, T.out_bias)));
                                    } else if (Y32) {
                                        sts_f32x2(ya_addr, (float)fma(ya, T.gain, T.out_bias), (float)fma(yb, T.gain, T.out_bias));
                                    } else {
                                        sts64(ya_addr, ya);
                                        sts64(ya_addr + 256, yb);
                                    }
                                }
                            } else {
                                filter_step<K, PAR, false>(T, va, xa);
                                filter_step<K, PAR, false>(T, vb, xb);
                            }
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) {
                    const int ls = step + C::STAGES;                 // the box that follows the ones in flight
                    if (ls < C::NSTEP) {
                        issue(stream, tis, ls, st);
                    } else {
                        if (!kIirDefer) {        // the ticket taken at the top of the tile is looked at here for the first time
                            nstream = next / tps;
                            ntis = next - nstream * tps;
                        }
                        if (next < P.total_tiles) issue(nstream, ntis, ls - C::NSTEP, st);
                    }
                }
                gstep++;
            }
        }

        // The owed tile's first look-back descriptor is requested now and looked at after phase 2: its
        // L2 round trip hides behind the scan.
        const int comp = lane < 2 * S ? lane : 0;
        unsigned long long pre = kDescEmpty;
        if (!kIirDefer) {          // the tile in hand is the one whose look-back follows
            owe = live;
            ptile = tile; pstream = stream; ptis = tis;
        }
        if (owe && ptis != 0)
            pre = ld_desc((P.chained ? P.desc_incl : P.desc_agg) + (size_t)(ptile - 1) * (2 * S) + comp);

        if (live) {
            // ---- phase 2: warp scan of run end states: E_t = sum_{j<=t} M^(t-j) f_j
#pragma unroll
            for (int k = 0; k < K; k++) {
                f[0][2 * k] = va[k][0] - T.sbias[2 * k]; f[0][2 * k + 1] = va[k][1] - T.sbias[2 * k + 1];
                f[1][2 * k] = vb[k][0] - T.sbias[2 * k]; f[1][2 * k + 1] = vb[k][1] - T.sbias[2 * k + 1];
            }
#pragma unroll
            for (int j = 0; j < 5; j++) {
                if (j < P.scan_rounds) {      // later rounds multiply by a matrix that is numerically zero
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
#pragma unroll
                                for (int col = col_lo<PAR>(row); col < col_hi(row); col++) acc = fma(T.Mp[j][row * S + col], g[r][col], acc);
                                f[r][row] = acc;
                            }
                    }
                }
            }
            // f = inclusive scan value E_lane.  Lane 31 holds the tile aggregate.
            if (lane == 31 && !P.chained) {
                double* agg = P.desc_agg + (size_t)tile * (2 * S);
#pragma unroll
                for (int r = 0; r < 2; r++)
#pragma unroll
                    for (int c = 0; c < S; c++) st_desc(agg + r * S + c, f[r][c]);
            }
        }

        if (!kIirDefer && live) {
#pragma unroll
            for (int r = 0; r < 2; r++)
#pragma unroll
                for (int c = 0; c < S; c++) pf[r][c] = f[r][c];
        }
        // ---- the owed tile: decoupled look-back, then phase 3 out of the other half of the output stage (deferred
        // form) / out of the stage phase 1 has just filled
        if (owe) {
            const uint32_t ystage = ystage0 + (kIirDefer ? (buf ^ 1u) : 0u) * C::YSTAGE_BYTES;
            // state entering the tile, component `lane` (< 2S) per lane
            const int crail = comp / S, crow = comp % S;
            // Every lane walks the look-back (lanes >= 2S repeat component 0: same addresses, no extra traffic) and
            // only the store is predicated: with the walk under `if (lane < 2S)` and more than one trip through its
            // loop, the two groups of lanes stayed apart all the way through phase 3 - every instruction issued
            // twice, at 16 threads (ncu, order 6 at fc = 10 kHz: profiles/r1_sweeps.txt, sweep 26).
            {
                double sin_c = 0.0;
                if (ptis == 0) {
                    sin_c = P.state_in[(size_t)pstream * (2 * S) + (crail ^ swap) * S + crow];
                } else {
                    const double* first = (P.chained ? P.desc_incl : P.desc_agg) + (size_t)(ptile - 1) * (2 * S) + comp;
                    while (pre == kDescEmpty) { __nanosleep(32); pre = ld_desc(first); }
                    sin_c = __longlong_as_double((long long)pre);                               // M_tile^0 = I
                    if (!P.chained) {
                        const int depth = (int)ptis < P.window ? (int)ptis : P.window;
                        for (int j = 2; j <= depth; j++) {
                            const double* a = P.desc_agg + (size_t)(ptile - j) * (2 * S) + crail * S;
                            const double* mp = P.mtile_pow + (size_t)(j - 1) * (S * S) + crow * S;
                            // all S loads go out together; tiles this far back are practically always published
                            unsigned long long dv[S];
#pragma unroll
                            for (int col = 0; col < S; col++) dv[col] = ld_desc(a + col);
#pragma unroll
                            for (int col = 0; col < S; col++) {
                                while (dv[col] == kDescEmpty) { __nanosleep(32); dv[col] = ld_desc(a + col); }
                                sin_c = fma(__ldg(mp + col), __longlong_as_double((long long)dv[col]), sin_c);
                            }
                        }
                        if ((int)ptis <= P.window) {     // the window reaches the start of the stream
                            const double* a = P.state_in + (size_t)pstream * (2 * S) + (crail ^ swap) * S;
                            const double* mp = P.mtile_pow + (size_t)ptis * (S * S) + crow * S;
#pragma unroll
                            for (int col = 0; col < S; col++) sin_c = fma(__ldg(mp + col), __ldg(a + col), sin_c);
                        }
                    }
                }
                if (lane < 2 * S) sts64(scratch + lane * 8, sin_c);
            }
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
                    const double e = shfl_up_f64(pf[r][c], 1);
                    s0[r][c] = lane == 0 ? 0.0 : e;
                }
#pragma unroll
            for (int row = 0; row < S; row++)
#pragma unroll
                for (int col = col_lo<PAR>(row); col < col_hi(row); col++) {
                    const double m = pl[row][col - col_lo<PAR>(row)];
                    s0[0][row] = fma(m, sin[0][col], s0[0][row]);
                    s0[1][row] = fma(m, sin[1][col], s0[1][row]);
                }

            // true end state of the tile (chained mode) / of the stream (last tile)
            const bool last_tile = (ptis == tps - 1);
            if (P.chained || (last_tile && P.write_state)) {
                const int lastl = last_tile && P.write_state ? (int)P.last_tile_runs - 1 : 31;
                // state after run `lastl`: M^(lastl+1) S_in + E_lastl, evaluated by lane lastl
                if (lane == lastl) {
                    double* incl = P.desc_incl + (size_t)ptile * (2 * S);
#pragma unroll
                    for (int row = 0; row < S; row++) {
                        double ea = pf[0][row], eb = pf[1][row];
#pragma unroll
                        for (int col = col_lo<PAR>(row); col < col_hi(row); col++) {
                            const double m = __ldg(P.plane + (row * S + col) * 33 + lastl + 1);
                            ea = fma(m, sin[0][col], ea);
                            eb = fma(m, sin[1][col], eb);
                        }
                        if (P.chained) { st_desc(incl + row, ea); st_desc(incl + S + row, eb); }
                        if (last_tile && P.write_state) {
                            P.state_out[(size_t)pstream * (2 * S) + (0 ^ swap) * S + row] = ea;
                            P.state_out[(size_t)pstream * (2 * S) + (1 ^ swap) * S + row] = eb;
                        }
                    }
                }
            }

            // ---- phase 3: correct the staged outputs and store (decimation already applied).
            // The I/Q <-> rail A/B assignment is resolved once per tile, not once per output.
            const uint32_t run = ptis * 32u + lane;
            if (run < P.n_runs) {
                double si[S], sq[S];
#pragma unroll
                for (int c = 0; c < S; c++) { si[c] = swap ? s0[1][c] : s0[0][c]; sq[c] = swap ? s0[0][c] : s0[1][c]; }
                const uint32_t yi0 = ystage + ((swap ? 32 : 0) + lane) * 8, yq0 = ystage + ((swap ? 0 : 32) + lane) * 8;
                const uint64_t o0 = (uint64_t)pstream * P.out_stride + (uint64_t)run * C::NK;
                if (Y32) {
                    // staged values are gain * y_zs already, T.R is gain * R (run_iir scales it for these launches)
                    float2* o = reinterpret_cast<float2*>(P.out) + o0;
                    static_assert(C::NK % 4 == 0, "kept outputs per run come in chunks of 4");
#pragma unroll (C::P3U)
                    for (int m0 = 0; m0 < C::NK; m0 += 4) {
                        float2 r[4];
#pragma unroll
                        for (int i = 0; i < 4; i++) {
                            const int m = m0 + i;
                            double yi, yq;
                            if (FIXED) {
                                const uint2 z = lds_u32x2(ystage + m * 256 + lane * 8);
                                yi = __hiloint2double((int)T.stage_hi, (int)z.x) - T.stage_mo;      // exact
                                yq = __hiloint2double((int)T.stage_hi, (int)z.y) - T.stage_mo;
                            } else {
                                const float2 z = lds_f32x2(ystage + m * 256 + lane * 8);
                                yi = (double)z.x; yq = (double)z.y;
                            }
#pragma unroll
                            for (int c = 0; c < S; c++) {
                                yi = fma(T.R[m][c], si[c], yi);
                                yq = fma(T.R[m][c], sq[c], yq);
                            }
                            r[i] = make_float2((float)yi, (float)yq);
                        }
                        if (P.out_vec) {
                            stg_cf32x4(o + m0, r[0], r[1], r[2], r[3]);
                        } else {
#pragma unroll
                            for (int i = 0; i < 4; i++) __stcs(o + m0 + i, r[i]);
                        }
                    }
                } else if (P.out_mode == OUT_F32) {
                    float2* o = reinterpret_cast<float2*>(P.out) + o0;
#pragma unroll 1
                    for (int m0 = 0; m0 < C::NK; m0 += 4) {
                        float2 r[4];
#pragma unroll
                        for (int i = 0; i < 4; i++) {
                            const int m = m0 + i;
                            double yi = lds64(yi0 + m * 512), yq = lds64(yq0 + m * 512);
#pragma unroll
                            for (int c = 0; c < S; c++) {
                                yi = fma(T.R[m][c], si[c], yi);
                                yq = fma(T.R[m][c], sq[c], yq);
                            }
                            yi = fma(yi, T.gain, T.out_bias);
                            yq = fma(yq, T.gain, T.out_bias);
                            r[i] = make_float2((float)yi, (float)yq);
                        }
                        if (P.out_vec) {
                            stg_cf32x4(o + m0, r[0], r[1], r[2], r[3]);
                        } else {
#pragma unroll
                            for (int i = 0; i < 4; i++) __stcs(o + m0 + i, r[i]);
                        }
                    }
                } else if (!Y32) {
                    // the Soapy formats and CF64: chunks of 4 outputs per lane, one vector store per chunk where
                    // the buffer allows it (out_vec)
#pragma unroll 1
                    for (int m0 = 0; m0 < C::NK; m0 += 4) {
                        double yi[4], yq[4];
#pragma unroll
                        for (int i = 0; i < 4; i++) {
                            const int m = m0 + i;
                            double a = lds64(yi0 + m * 512), b = lds64(yq0 + m * 512);
#pragma unroll
                            for (int c = 0; c < S; c++) {
                                a = fma(T.R[m][c], si[c], a);
                                b = fma(T.R[m][c], sq[c], b);
                            }
                            yi[i] = fma(a, T.gain, T.out_bias);
                            yq[i] = fma(b, T.gain, T.out_bias);
                        }
                        const uint64_t o = o0 + m0;
                        if (P.out_mode == OUT_F64) {
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
                                float2 r[4];
#pragma unroll
                                for (int i = 0; i < 4; i++) r[i] = make_float2((float)qi[i] * (1.0f / 4096.0f), (float)qq[i] * (1.0f / 4096.0f));
                                float2* op = reinterpret_cast<float2*>(P.out) + o;
                                if (P.out_vec) {
                                    stg_cf32x4(op, r[0], r[1], r[2], r[3]);
                                } else {
#pragma unroll
                                    for (int i = 0; i < 4; i++) __stcs(op + i, r[i]);
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
                }
            }
            __syncwarp();   // this half of the output stage and the scratch are free again
        }
        if (!have) break;
        if (kIirDefer) {
            owe = live;
            ptile = tile; pstream = stream; ptis = tis;
            if (live) {
#pragma unroll
                for (int r = 0; r < 2; r++)
#pragma unroll
                    for (int c = 0; c < S; c++) pf[r][c] = f[r][c];
                buf ^= 1u;
            }
        }
        // lane 0's view of the ticket pipeline becomes everybody's current tile
        tile = __shfl_sync(0xffffffffu, next, 0);
        stream = __shfl_sync(0xffffffffu, nstream, 0);
        tis = __shfl_sync(0xffffffffu, ntis, 0);
        if (kIirDefer) next = next2;
    }
}

// Launches one instantiation on a persistent grid: as many CTAs as fit on the device (occupancy is queried
// once per instantiation), never more than the tiles need.
template <int K, int D, int IN, bool PAR, bool Y32>
int launch_iir_t(const CUtensorMap& tmap, const IirParams& p, const IirTables& t, int num_sms, cudaStream_t st)
{
    using C = IirCfg<K, D, IN, PAR, Y32>;
    auto kern = iir_scan_kernel<K, D, IN, PAR, Y32>;
    // the shared-memory opt-in is a per-device attribute of the function: one slot per device ordinal (the
    // writes are idempotent, so two host threads racing on a slot only repeat the query)
    static int occ_dev[kMaxDevices] = {0};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= kMaxDevices) return (int)cudaErrorInvalidDevice;
    if (occ_dev[dev] == 0) {
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES);
        if (e != cudaSuccess) return (int)e;
        int n = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, C::WPC * 32, C::SMEM_BYTES);
        if (e != cudaSuccess) return (int)e;
        occ_dev[dev] = n > 0 ? n : 1;
    }
    const int occ = occ_dev[dev];
    const uint64_t want = ((uint64_t)p.total_tiles + C::WPC - 1) / C::WPC;
    const uint64_t cap = (uint64_t)occ * (uint64_t)num_sms;
    const int grid = (int)(want < cap ? want : cap);
    kern<<<grid, C::WPC * 32, C::SMEM_BYTES, st>>>(tmap, p, t);
    return (int)cudaGetLastError();
}

typedef int (*iir_launch_fn)(const CUtensorMap&, const IirParams&, const IirTables&, int, cudaStream_t);

// the staged kernel for (K, D, ...), unless that combination belongs to the rerun kernel (iir_rerun.cuh)
template <int K, int D, int IN, bool PAR, bool Y32>
iir_launch_fn staged_launch()
{
    if constexpr (iir_is_rerun(K, D, PAR)) return nullptr;
    else return &launch_iir_t<K, D, IN, PAR, Y32>;
}

#define CBL_IIR_DISPATCH_D(KK, IN_, PAR_, Y32_)                       \
    switch (D) {                                                     \
        case 1:  return staged_launch<KK, 1, IN_, PAR_, Y32_>();      \
        case 2:  return staged_launch<KK, 2, IN_, PAR_, Y32_>();      \
        case 4:  return staged_launch<KK, 4, IN_, PAR_, Y32_>();      \
        case 8:  return staged_launch<KK, 8, IN_, PAR_, Y32_>();      \
        case 16: return staged_launch<KK, 16, IN_, PAR_, Y32_>();     \
        case 32: return staged_launch<KK, 32, IN_, PAR_, Y32_>();     \
        default: return nullptr;                                     \
    }

// The float-staged (Y32) kernels exist for the preferred realisation only: the parallel form, also for a single
// section (x + alpha v1 + beta v2 is what lets the integer inputs enter by mantissa injection).
template <int KK, bool Y32>
iir_launch_fn iir_pick_launch(int D, int in_kind, bool par)
{
    constexpr bool PREF = true;          // the realisation AUTO picks when the poles allow it
    if (par == PREF) {
        if (in_kind == IN_WORDS) { CBL_IIR_DISPATCH_D(KK, IN_WORDS, PREF, Y32) }
        if (in_kind == IN_WORDS_HIF) { CBL_IIR_DISPATCH_D(KK, IN_WORDS_HIF, PREF, Y32) }
        if (in_kind == IN_CS16)  { CBL_IIR_DISPATCH_D(KK, IN_CS16, PREF, Y32) }
        if (in_kind == IN_CF32)  { CBL_IIR_DISPATCH_D(KK, IN_CF32, PREF, Y32) }
        return nullptr;
    }
    if (Y32) return nullptr;
    if (in_kind == IN_WORDS) { CBL_IIR_DISPATCH_D(KK, IN_WORDS, false, false) }
    if (in_kind == IN_WORDS_HIF) { CBL_IIR_DISPATCH_D(KK, IN_WORDS_HIF, false, false) }
    if (in_kind == IN_CS16)  { CBL_IIR_DISPATCH_D(KK, IN_CS16, false, false) }
    if (in_kind == IN_CF32)  { CBL_IIR_DISPATCH_D(KK, IN_CF32, false, false) }
    return nullptr;
}

}  // namespace cbl
