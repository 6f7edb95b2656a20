// Shared device/host helpers for libcaribou_cuda (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace cbl {

// ------------------------------------------------------------------ SMI word format (A0)
// [31:30]='10' [29:17]=hi13 [16]='0' [15:14]='01' [13:1]=lo13 [0]=S
// firmware/lvds_rx.v:17-18,44-81; LC/caribou_smi/caribou_smi.c:338-340
constexpr uint32_t kHdrMask = 0xC001C000u;   // LC/caribou_smi/caribou_smi.c:256-259
constexpr uint32_t kHdrGood = 0x80004000u;

__host__ __device__ __forceinline__ bool hdr_ok(uint32_t w) { return (w & kHdrMask) == kHdrGood; }
// 13-bit two's complement fields, sign-extended (caribou_smi.c:351-357)
__host__ __device__ __forceinline__ int hi13(uint32_t w) { return ((int32_t)(w << 2)) >> 19; }
__host__ __device__ __forceinline__ int lo13(uint32_t w) { return ((int32_t)(w << 18)) >> 19; }

// One sub-stream of a capture: the whole byte stream (B == 0) or every other block of B words
// of a dual-channel capture.
struct SubView {
    const uint32_t* words;   // raw capture, 4-byte aligned
    uint64_t nbytes;         // bytes in this sub-stream
    uint64_t n_sub_words;    // ceil(nbytes / 4)
    uint32_t B;              // 0 = not interleaved
    uint32_t sub;            // 0 = S1G blocks, 1 = HiF blocks
};

#ifdef __CUDACC__
// i-th aligned 32-bit word of the sub-stream; bytes past the end read as zero.
__device__ __forceinline__ uint32_t sv_word(const SubView& v, uint64_t i)
{
    if (v.B == 0) {
        if ((i + 1) * 4 <= v.nbytes) return __ldg(v.words + i);
        uint32_t w = 0;
        const uint8_t* b = reinterpret_cast<const uint8_t*>(v.words);
        for (int k = 0; k < 4; k++)
            if (i * 4 + k < v.nbytes) w |= (uint32_t)b[i * 4 + k] << (8 * k);
        return w;
    }
    if (i >= v.n_sub_words) return 0;
    const uint64_t raw = (v.B == 1) ? (2 * i + v.sub) : (((i / v.B) * 2 + v.sub) * v.B + i % v.B);
    return __ldg(v.words + raw);
}
// little-endian u32 at an arbitrary byte offset of the sub-stream
// (the reference's *(uint32_t*)&buffer[offs], caribou_smi.c:250-253)
__device__ __forceinline__ uint32_t sv_u32_at(const SubView& v, uint64_t byte_off)
{
    const uint64_t i = byte_off >> 2;
    const uint32_t sh = (uint32_t)(byte_off & 3) * 8;
    const uint32_t lo = sv_word(v, i);
    if (sh == 0) return lo;
    const uint32_t hi = sv_word(v, i + 1);
    return __funnelshift_r(lo, hi, sh);
}

__device__ __forceinline__ uint32_t smem_u32(const void* p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}
#endif

}  // namespace cbl
