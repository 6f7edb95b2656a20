// Kernel (a) of libcaribou_cuda: per-chunk alignment search and the fused
// unpack / demux / sync-check / sign-extend / CS16 / CF32 kernel, plus the ordered
// compaction of flagged words.
//
// What it computes is fixed by the reference (LC/caribou_smi/caribou_smi.c:235-393); how it
// is computed is not.  A tile is 4096 consecutive raw words; a plain tile (whole words, every
// chunk it touches aligned) is decoded with four coalesced 128-bit loads in flight per thread,
// one 128-bit CS16 store, one 32-bit meta store and two 128-bit CF32 stores per 4 samples;
// everything else takes the per-sample path with the reference's integer arithmetic.
// Flagged words are counted per tile by the decode kernel (warp ballots, one atomic per warp
// that saw any); a one-CTA scan turns the counts into list offsets and a work list, and only
// the flagged tiles are read again to emit their indices in ascending order.
#include "unpack.cuh"
#include <limits.h>

namespace cbl {

// ------------------------------------------------------------------ alignment search (A2)
// A CTA takes kFindChunks consecutive chunks of one stream.  The aligned common case is four loads: warp w tests
// candidate 0 of chunk w with its first four lanes (one CTA per chunk was 8192 CTAs of which one thread each did that:
// 7.5 us for C4's 8192 chunks, launch- and latency-bound).  A chunk that fails is swept by the whole CTA,
// 4 * blockDim candidates at a time, stopping at the first hit.
constexpr int kFindChunks = 8;
constexpr int kFindThreads = kFindChunks * 32;
__global__ void __launch_bounds__(kFindThreads) find_offsets_kernel(OffsetsParams p)
{
    __shared__ int s_res[kFindChunks];
    __shared__ int s_found;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t c0 = blockIdx.x * kFindChunks;
    p.sv.words = reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(p.sv.words) +
                                                   (uint64_t)blockIdx.y * p.stream_stride_bytes);
    auto chunk_len = [&](uint32_t c) {
        const uint64_t rem = p.sv.nbytes - (uint64_t)c * p.chunk_bytes;
        return rem < p.chunk_bytes ? (uint32_t)rem : p.chunk_bytes;
    };
    {
        const uint32_t c = c0 + warp;
        int res = 0;
        if (c < p.n_chunks && chunk_len(c) > 16) {             // len <= 16: caribou_smi.c:240-243
            const bool ok = lane < 4 ? hdr_ok(sv_u32_at(p.sv, (uint64_t)c * p.chunk_bytes + 4 * lane)) : true;
            res = __all_sync(0xffffffffu, ok) ? 0 : INT_MAX;
        }
        if (lane == 0) s_res[warp] = res;
    }
    __syncthreads();
    for (int w = 0; w < kFindChunks; w++) {
        if (s_res[w] != INT_MAX) continue;                     // the same for every thread of the CTA
        const uint32_t c = c0 + (uint32_t)w;
        const uint64_t start = (uint64_t)c * p.chunk_bytes;
        const uint32_t ncand = chunk_len(c) - 16;              // candidates offs in [0, len-16), caribou_smi.c:248
        if (threadIdx.x == 0) s_found = INT_MAX;
        __syncthreads();
        int result = INT_MAX;
        for (uint32_t base = 0; base < ncand; base += 4 * kFindThreads) {
            // this thread: candidates base + 4*tid + {0,1,2,3}; they share aligned words w[0..4]
            const uint32_t o0 = base + 4 * threadIdx.x;
            if (o0 < ncand) {
                uint32_t wd[5];
                const uint64_t wi = (start + o0) >> 2;   // start is a multiple of 4 (chunk_bytes % 16 == 0)
#pragma unroll
                for (int k = 0; k < 5; k++) wd[k] = sv_word(p.sv, wi + k);
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    if (o0 + b < ncand) {
                        bool ok = true;
#pragma unroll
                        for (int k = 0; k < 4; k++) {
                            const uint32_t u = b ? __funnelshift_r(wd[k], wd[k + 1], 8 * b) : wd[k];
                            ok = ok && hdr_ok(u);
                        }
                        if (ok) { atomicMin(&s_found, (int)(o0 + b)); break; }
                    }
                }
            }
            __syncthreads();
            result = s_found;
            __syncthreads();
            if (result != INT_MAX) break;
        }
        if (threadIdx.x == 0) s_res[w] = result == INT_MAX ? -1 : result;    // caribou_smi.c:285-289
        __syncthreads();
    }
    if (threadIdx.x < kFindChunks && c0 + threadIdx.x < p.n_chunks) {
        const uint32_t c = c0 + threadIdx.x;
        const int result = s_res[threadIdx.x];
        p.offs[(size_t)blockIdx.y * p.offs_stride + c] = result;
        if (result < 0) {
            atomicMin(p.first_bad + (size_t)blockIdx.y * p.first_bad_stride, (int)c);
            if (p.lost) *p.lost = 1;
        }
        if (result > 0) {
            atomicAdd(p.n_misaligned, 1);
            if (p.nmis) atomicAdd(p.nmis + blockIdx.y, 1);
        }
        if (result != 0) {
            *p.any_misaligned = 1;
            if (p.gate) {
                p.gate[blockIdx.y] = 1;
                atomicMin(p.gate_seq + blockIdx.y, p.seq);
            }
        }
    }
}

void launch_find_offsets(const OffsetsParams& p, cudaStream_t st)
{
    if (p.n_chunks == 0 || p.n_streams == 0) return;
    find_offsets_kernel<<<dim3((p.n_chunks + kFindChunks - 1) / kFindChunks, p.n_streams), kFindThreads, 0, st>>>(p);
}

// ------------------------------------------------------------------ decode helpers (A3)
__device__ __forceinline__ void decode_iq(uint32_t w, int layout, int& i, int& q)
{
    const int h = hi13(w), l = lo13(w);
    i = layout ? l : h;      // S1G: i = bits 29:17 (caribou_smi.c:351-354); HiF: i = bits 13:1 (:371-373)
    q = layout ? h : l;
}
__device__ __forceinline__ uint32_t pack_cs16(int i, int q)
{
    return ((uint32_t)i & 0xFFFFu) | ((uint32_t)q << 16);
}
__device__ __forceinline__ float2 to_cf32(int i, int q)
{
    // (float)v / 4096.0f, exact for |v| <= 4096 (LC/CaribouLiteRadioCpp.cpp:91)
    return make_float2((float)i * (1.0f / 4096.0f), (float)q * (1.0f / 4096.0f));
}

// Four consecutive output slots j0..j0+3 of sub-stream `s` whose aligned source words are w[0..3].
__device__ __forceinline__ void emit_group(const UnpackParams& p, int s, uint64_t j0, const uint32_t (&w)[4])
{
    const SubView& sv = p.sv[s];
    if (j0 * 4 >= sv.nbytes) return;
    const uint32_t c = (uint32_t)(j0 / p.spc);
    if ((int)c >= *p.first_bad[s]) return;                 // caribou_smi.c:665-668: nothing after a lost sync
    const uint32_t l0 = (uint32_t)(j0 - (uint64_t)c * p.spc);
    const int offs = p.offs[s][c];
    const uint64_t cstart = (uint64_t)c * p.chunk_bytes;
    const uint64_t rem = sv.nbytes - cstart;
    const uint32_t len = rem < p.chunk_bytes ? (uint32_t)rem : p.chunk_bytes;
    const uint32_t nfull = len >> 2;
    const int layout = p.layout[s];
    if (offs == 0 && l0 + 4 <= nfull && p.vec_ok) {
        int i[4], q[4];
#pragma unroll
        for (int e = 0; e < 4; e++) decode_iq(w[e], layout, i[e], q[e]);
        if (p.cs16[s]) {
            uint4 v = make_uint4(pack_cs16(i[0], q[0]), pack_cs16(i[1], q[1]), pack_cs16(i[2], q[2]), pack_cs16(i[3], q[3]));
            __stcs(reinterpret_cast<uint4*>(p.cs16[s] + j0), v);
        }
        if (p.meta[s]) {
            const uint32_t m = (w[0] & 1u) | ((w[1] & 1u) << 8) | ((w[2] & 1u) << 16) | ((w[3] & 1u) << 24);
            __stcs(reinterpret_cast<uint32_t*>(p.meta[s] + j0), m);
        }
        if (p.cf32[s]) {
            const float2 a = to_cf32(i[0], q[0]), b = to_cf32(i[1], q[1]);
            const float2 cc = to_cf32(i[2], q[2]), d = to_cf32(i[3], q[3]);
            float4* o = reinterpret_cast<float4*>(p.cf32[s] + j0);
            __stcs(o, make_float4(a.x, a.y, b.x, b.y));
            __stcs(o + 1, make_float4(cc.x, cc.y, d.x, d.y));
        }
        return;
    }
    // general path: misaligned chunk, chunk tail, or unaligned output pointers
    const uint32_t shorten = offs > 0 ? (uint32_t)offs / 4 + 1 : 0;        // caribou_smi.c:323
    const uint32_t nvalid = nfull - shorten;
#pragma unroll
    for (int e = 0; e < 4; e++) {
        const uint32_t l = l0 + e;
        const uint64_t j = j0 + e;
        if (l < nvalid) {
            const uint32_t u = offs == 0 ? w[e] : sv_u32_at(sv, cstart + offs + 4ull * l);
            int i, q;
            decode_iq(u, layout, i, q);
            if (p.cs16[s]) p.cs16[s][j] = pack_cs16(i, q);
            if (p.meta[s]) p.meta[s][j] = (uint8_t)(u & 1u);
            if (p.cf32[s]) p.cf32[s][j] = to_cf32(i, q);
        } else if (l == nvalid && shorten > 0 && nvalid >= 2 && (p.cs16[s] || p.cf32[s])) {
            // caribou_smi.c:382-389: one extrapolated sample, C int arithmetic, narrowed to int16;
            // its meta slot is not written
            int i1, q1, i2, q2;
            decode_iq(sv_u32_at(sv, cstart + offs + 4ull * (l - 1)), layout, i1, q1);
            decode_iq(sv_u32_at(sv, cstart + offs + 4ull * (l - 2)), layout, i2, q2);
            const int i = (int)(int16_t)(110 * i1 / 100 - i2 / 10);
            const int q = (int)(int16_t)(110 * q1 / 100 - q2 / 10);
            if (p.cs16[s]) p.cs16[s][j] = pack_cs16(i, q);
            if (p.cf32[s]) p.cf32[s][j] = to_cf32(i, q);
        }
        // l > nvalid: slot left untouched, as the reference leaves it
    }
}

// ------------------------------------------------------------------ flag compaction (A10)
// Flagged words are rare (one sync word per second of signal, invalid headers only on a damaged link), so the
// decode kernel only COUNTS them per tile - warps that saw a flag add their packed count to the tile's slot,
// nobody waits for anybody.  A one-CTA scan turns the per-tile counts into list offsets and a work list of the
// tiles that have anything to emit; those tiles alone are read a second time to write their indices in order.
__device__ __forceinline__ uint64_t pack_counts(uint32_t inv, uint32_t syn) { return ((uint64_t)inv << 32) | syn; }

// Is every output slot of tile `tile` a plain aligned decode?  True when the tile holds whole raw words only,
// all chunks it touches (in every sub-stream) have offs == 0, lie before the first un-alignable chunk and the
// tile does not reach into a sub-stream's partial last word.  Uniform over the CTA.
__device__ bool tile_is_plain(const UnpackParams& p, uint64_t w0, uint64_t w1)
{
    if (!p.vec_ok || w1 > (p.nbytes >> 2)) return false;
    for (int s = 0; s < p.n_subs; s++) {
        uint64_t j_lo, j_hi;           // (superset of the) output slots of sub-stream s this tile writes
        if (p.B == 0)      { j_lo = w0; j_hi = w1; }
        else if (p.B == 1) { j_lo = w0 >> 1; j_hi = (w1 + 1) >> 1; }
        else               { j_lo = (w0 / (2ull * p.B)) * p.B; j_hi = ((w1 - 1) / (2ull * p.B)) * p.B + p.B; }
        if (j_hi > (p.sv[s].nbytes >> 2)) return false;
        const uint64_t c_lo = j_lo / p.spc, c_hi = (j_hi - 1) / p.spc;
        if (c_hi - c_lo > 3) return false;
        if ((int64_t)c_hi >= (int64_t)*p.first_bad[s]) return false;
        for (uint64_t c = c_lo; c <= c_hi; c++)
            if (p.offs[s][c] != 0) return false;
    }
    return true;
}

// plain decode of 4 consecutive samples of sub-stream s from their 4 source words: 128-bit CS16 store, 32-bit
// meta store, two 128-bit CF32 stores
__device__ __forceinline__ void emit_plain4(const UnpackParams& p, int s, uint64_t j0, const uint32_t (&w)[4])
{
    const int layout = p.layout[s];
    int i[4], q[4];
#pragma unroll
    for (int e = 0; e < 4; e++) decode_iq(w[e], layout, i[e], q[e]);
    if (p.cs16[s])
        __stcs(reinterpret_cast<uint4*>(p.cs16[s] + j0),
               make_uint4(pack_cs16(i[0], q[0]), pack_cs16(i[1], q[1]), pack_cs16(i[2], q[2]), pack_cs16(i[3], q[3])));
    if (p.meta[s])
        __stcs(reinterpret_cast<uint32_t*>(p.meta[s] + j0),
               (w[0] & 1u) | ((w[1] & 1u) << 8) | ((w[2] & 1u) << 16) | ((w[3] & 1u) << 24));
    if (p.cf32[s]) {
        const float2 a = to_cf32(i[0], q[0]), b = to_cf32(i[1], q[1]), cc = to_cf32(i[2], q[2]), d = to_cf32(i[3], q[3]);
        float4* o = reinterpret_cast<float4*>(p.cf32[s] + j0);
        __stcs(o, make_float4(a.x, a.y, b.x, b.y));
        __stcs(o + 1, make_float4(cc.x, cc.y, d.x, d.y));
    }
}

// A tile is kUnpackTileWords consecutive raw words.  Single-channel and block-interleaved captures are walked
// in groups of 4 words per thread (one fully coalesced 128-bit load per warp instruction); a word-interleaved
// capture (B == 1) in groups of 8, which give 4 samples to each channel.
__global__ void __launch_bounds__(kUnpackThreads) unpack_kernel(const __grid_constant__ UnpackParams p)
{
    __shared__ int s_plain;
    const int lane = threadIdx.x & 31;
    const uint32_t tile = blockIdx.x;
    if (threadIdx.x == 0)
        s_plain = tile_is_plain(p, (uint64_t)tile * kUnpackTileWords, (uint64_t)(tile + 1) * kUnpackTileWords) ? 1 : 0;
    __syncthreads();
    const bool plain = s_plain != 0;
    const uint64_t tile_w0 = (uint64_t)tile * kUnpackTileWords;
    const uint64_t n_raw = p.nbytes >> 2;            // whole raw words
    const int W = (p.B == 1) ? 8 : 4;                // words per group
    const int iters = kUnpackTileWords / (kUnpackThreads * W);

    uint32_t invb[kUnpackMaxIters], synb[kUnpackMaxIters];
#pragma unroll
    for (int it = 0; it < kUnpackMaxIters; it++) invb[it] = synb[it] = 0;

    if (plain && W == 4) {
        // ---- aligned single-channel / block-interleaved tile: nothing to decide per sample
        uint4 v[kUnpackMaxIters];
#pragma unroll
        for (int it = 0; it < kUnpackMaxIters; it++)
            v[it] = __ldcs(reinterpret_cast<const uint4*>(p.words + tile_w0 + ((uint64_t)it * kUnpackThreads + threadIdx.x) * 4));
#pragma unroll
        for (int it = 0; it < kUnpackMaxIters; it++) {
            const uint64_t k0 = tile_w0 + ((uint64_t)it * kUnpackThreads + threadIdx.x) * 4;
            const uint32_t w[4] = {v[it].x, v[it].y, v[it].z, v[it].w};
#pragma unroll
            for (int e = 0; e < 4; e++) {
                invb[it] |= (hdr_ok(w[e]) ? 0u : 1u) << e;
                synb[it] |= (w[e] & 1u) << e;
            }
            if (p.B == 0) {
                emit_plain4(p, 0, k0, w);
            } else {
                const uint32_t kk = (uint32_t)k0, blk = kk / p.B;         // at most 2^31 words per launch
                emit_plain4(p, (int)(blk & 1u), (uint64_t)(blk >> 1) * p.B + (kk - blk * p.B), w);
            }
        }
    } else if (plain) {
        // ---- aligned word-interleaved tile: even words to S1G, odd words to HiF
#pragma unroll
        for (int it = 0; it < kUnpackMaxIters / 2; it++) {
            const uint64_t k0 = tile_w0 + ((uint64_t)it * kUnpackThreads + threadIdx.x) * 8;
            const uint4 a = __ldcs(reinterpret_cast<const uint4*>(p.words + k0));
            const uint4 b = __ldcs(reinterpret_cast<const uint4*>(p.words + k0 + 4));
            const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
            for (int e = 0; e < 8; e++) {
                invb[it] |= (hdr_ok(w[e]) ? 0u : 1u) << e;
                synb[it] |= (w[e] & 1u) << e;
            }
            const uint32_t ev[4] = {w[0], w[2], w[4], w[6]}, od[4] = {w[1], w[3], w[5], w[7]};
            emit_plain4(p, 0, k0 >> 1, ev);
            emit_plain4(p, 1, k0 >> 1, od);
        }
    } else {
        // ---- general tile: misaligned chunks, lost sync, stream tail, unaligned output pointers
        for (int it = 0; it < iters; it++) {
            const uint64_t k0 = tile_w0 + ((uint64_t)it * kUnpackThreads + threadIdx.x) * W;
            if (k0 * 4 >= p.nbytes) continue;
            uint32_t w[8];
            uint32_t ib = 0, sb = 0;
            for (int e = 0; e < W; e++) {
                w[e] = (k0 + e < n_raw) ? p.words[k0 + e] : 0u;
                if (k0 + e < n_raw) {
                    ib |= (hdr_ok(w[e]) ? 0u : 1u) << e;
                    sb |= (w[e] & 1u) << e;
                }
            }
#pragma unroll
            for (int q = 0; q < kUnpackMaxIters; q++) if (q == it) { invb[q] = ib; synb[q] = sb; }
            if (p.B == 0) {
                // the (possibly partial) tail word still has to reach sv_word's byte-wise path
                if (k0 + 4 > n_raw)
                    for (int e = 0; e < 4; e++) w[e] = sv_word(p.sv[0], k0 + e);
                const uint32_t lo[4] = {w[0], w[1], w[2], w[3]};
                emit_group(p, 0, k0, lo);
            } else if (p.B == 1) {
                const uint32_t ev[4] = {w[0], w[2], w[4], w[6]}, od[4] = {w[1], w[3], w[5], w[7]};
                emit_group(p, 0, k0 >> 1, ev);
                emit_group(p, 1, k0 >> 1, od);
            } else {   // B % 8 == 0: the whole group belongs to one sub-stream
                const int s = (int)((k0 / p.B) & 1u);
                const uint64_t j0 = (k0 / (2ull * p.B)) * p.B + k0 % p.B;
                const uint32_t lo[4] = {w[0], w[1], w[2], w[3]};
                emit_group(p, s, j0, lo);
            }
        }
    }
    if (!p.want_flags) return;
    // ---- per-tile counts of flagged words
    uint32_t ci = 0, cs = 0;
#pragma unroll
    for (int it = 0; it < kUnpackMaxIters; it++) { ci += __popc(invb[it]); cs += __popc(synb[it]); }
    if (__ballot_sync(0xffffffffu, (ci | cs) != 0)) {
        uint32_t v = ci | (cs << 16);                     // at most 8 * 4 * 32 = 1024 per warp and kind
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) atomicAdd(reinterpret_cast<unsigned long long*>(p.tile_counts + tile), pack_counts(v & 0xFFFFu, v >> 16));
    }
}

// One CTA: exclusive scan of the per-tile counts, running totals, work list of the tiles that have flags.
__global__ void __launch_bounds__(1024) flag_scan_kernel(const __grid_constant__ UnpackParams p)
{
    __shared__ uint64_t s_warp[32];
    __shared__ uint64_t s_carry;
    __shared__ uint32_t s_nwork;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { s_carry = 0; s_nwork = 0; }
    __syncthreads();
    for (uint32_t base = 0; base < p.n_tiles; base += 1024) {
        const uint32_t t = base + threadIdx.x;
        const uint64_t mine = t < p.n_tiles ? p.tile_counts[t] : 0;
        uint64_t v = mine;                                 // both 32-bit halves scan independently (no carry: < 2^31 words)
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint64_t u = __shfl_up_sync(0xffffffffu, v, d);
            if (lane >= d) v += u;
        }
        if (lane == 31) s_warp[warp] = v;
        __syncthreads();
        if (warp == 0) {
            uint64_t w = s_warp[lane];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint64_t u = __shfl_up_sync(0xffffffffu, w, d);
                if (lane >= d) w += u;
            }
            s_warp[lane] = w;
        }
        __syncthreads();
        const uint64_t excl = s_carry + (warp ? s_warp[warp - 1] : 0) + v - mine;
        if (t < p.n_tiles) {
            p.tile_offsets[t] = excl;
            if (mine != 0 && (p.inv_idx || p.syn_idx)) p.work[atomicAdd(&s_nwork, 1u)] = t;
        }
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = excl + mine;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        p.totals_out[0] = p.totals_in[0] + (s_carry >> 32);
        p.totals_out[1] = p.totals_in[1] + (s_carry & 0xFFFFFFFFull);
        *p.n_work = s_nwork;
    }
}

// Flagged tiles only: read the tile again and write the indices of its flagged words, ascending, at the offsets
// the scan assigned.  Word order within a tile = (iteration, warp, lane, bit) order.
__global__ void __launch_bounds__(kUnpackThreads) flag_emit_kernel(const __grid_constant__ UnpackParams p)
{
    __shared__ uint32_t s_cnt[kUnpackMaxIters][kUnpackThreads / 32][2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t n_raw = p.nbytes >> 2;
    const int W = (p.B == 1) ? 8 : 4;
    const int iters = kUnpackTileWords / (kUnpackThreads * W);
    const uint32_t n_work = *p.n_work;
    for (uint32_t wi = blockIdx.x; wi < n_work; wi += gridDim.x) {
        const uint32_t tile = p.work[wi];
        const uint64_t tile_w0 = (uint64_t)tile * kUnpackTileWords;
        uint32_t invb[kUnpackMaxIters], synb[kUnpackMaxIters], pre[kUnpackMaxIters];
#pragma unroll
        for (int it = 0; it < kUnpackMaxIters; it++) {
            invb[it] = synb[it] = 0; pre[it] = 0;
            uint32_t c2 = 0;
            if (it < iters) {
                const uint64_t k0 = tile_w0 + ((uint64_t)it * kUnpackThreads + threadIdx.x) * W;
                for (int e = 0; e < W; e++) {
                    if (k0 + e < n_raw) {
                        const uint32_t w = __ldg(p.words + k0 + e);
                        invb[it] |= (hdr_ok(w) ? 0u : 1u) << e;
                        synb[it] |= (w & 1u) << e;
                    }
                }
                // packed inclusive warp scan of (invalid, sync) counts: 16 bits each (at most 256 per warp)
                uint32_t v = __popc(invb[it]) | (__popc(synb[it]) << 16);
                const uint32_t own = v;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const uint32_t u = __shfl_up_sync(0xffffffffu, v, d);
                    if (lane >= d) v += u;
                }
                pre[it] = v - own;                               // exclusive, this thread
                c2 = __shfl_sync(0xffffffffu, v, 31);
            }
            if (lane == 0) { s_cnt[it][warp][0] = c2 & 0xFFFFu; s_cnt[it][warp][1] = c2 >> 16; }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t ti = 0, ts = 0;
            for (int it = 0; it < kUnpackMaxIters; it++)
                for (int wv = 0; wv < kUnpackThreads / 32; wv++) {
                    const uint32_t a = s_cnt[it][wv][0], b = s_cnt[it][wv][1];
                    s_cnt[it][wv][0] = ti; s_cnt[it][wv][1] = ts;       // exclusive base within the tile
                    ti += a; ts += b;
                }
        }
        __syncthreads();
        const uint64_t off = p.tile_offsets[tile];
        const uint64_t bi = p.totals_in[0] + (off >> 32), bs = p.totals_in[1] + (off & 0xFFFFFFFFull);
#pragma unroll
        for (int it = 0; it < kUnpackMaxIters; it++) {
            if ((invb[it] | synb[it]) == 0) continue;
            const uint64_t k0 = p.index_base + tile_w0 + ((uint64_t)it * kUnpackThreads + threadIdx.x) * W;
            uint64_t pi = bi + s_cnt[it][warp][0] + (pre[it] & 0xFFFFu);
            uint64_t ps = bs + s_cnt[it][warp][1] + (pre[it] >> 16);
            for (int e = 0; e < 8; e++) {
                if ((invb[it] >> e) & 1u) { if (p.inv_idx && pi < p.inv_cap) p.inv_idx[pi] = (uint32_t)(k0 + e); pi++; }
                if ((synb[it] >> e) & 1u) { if (p.syn_idx && ps < p.syn_cap) p.syn_idx[ps] = (uint32_t)(k0 + e); ps++; }
            }
        }
        __syncthreads();
    }
}

int launch_unpack(const UnpackParams& p, int num_sms, cudaStream_t st)
{
    if (p.n_tiles == 0) return 0;
    unpack_kernel<<<p.n_tiles, kUnpackThreads, 0, st>>>(p);
    if (!p.want_flags) return 1;
    flag_scan_kernel<<<1, 1024, 0, st>>>(p);
    if (!p.inv_idx && !p.syn_idx) return 2;
    flag_emit_kernel<<<num_sms, kUnpackThreads, 0, st>>>(p);
    return 3;
}

// ------------------------------------------------------------------ synthetic words
__device__ __forceinline__ uint64_t splitmix64(uint64_t x)
{
    uint64_t z = x + 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__global__ void synth_kernel(uint32_t* out, uint64_t n0, uint64_t n, int hif_layout, uint64_t seed,
                             const int16_t* __restrict__ tone_i, const int16_t* __restrict__ tone_q)
{
    for (uint64_t k = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t g = n0 + k;
        const int ph = (int)(g % 800u);
        const int ui = (int)(splitmix64(seed ^ (g * 2 + 0)) & 127u) - 64;
        const int uq = (int)(splitmix64(seed ^ (g * 2 + 1)) & 127u) - 64;
        int I = tone_i[ph] + ui, Q = tone_q[ph] + uq;
        I = max(-4096, min(4095, I));
        Q = max(-4096, min(4095, Q));
        const uint32_t S = (g % 4000000u) == 0;
        const uint32_t hi = (uint32_t)(hif_layout ? Q : I) & 0x1FFFu;
        const uint32_t lo = (uint32_t)(hif_layout ? I : Q) & 0x1FFFu;
        out[k] = 0x80004000u | (hi << 17) | (lo << 1) | S;
    }
}

void launch_synth(uint32_t* out, uint64_t n0, uint64_t n, int hif_layout, uint64_t seed,
                  const int16_t* tone_i, const int16_t* tone_q, int num_sms, cudaStream_t st)
{
    if (n == 0) return;
    synth_kernel<<<(num_sms > 0 ? num_sms : 148) * 8, 256, 0, st>>>(out, n0, n, hif_layout, seed, tone_i, tone_q);
}

}  // namespace cbl
