// Kernel (c), SURVEY 8f N3: the TX direction - native int16 pairs (or the Soapy formats converted to them) packed
// into 32-bit SMI TX words.  Follows caribou_smi_generate_data (LC/caribou_smi/caribou_smi.c:688-717, layout comment
// :693-696) with the sample read the reference has commented out put back, and Stream::WriteSamples
// (LC/soapy_api/CaribouliteStream.cpp:199-244) for the conversions.
//
//   byte 0 (first on the wire)  [SOF TXC CTX I12 I11 I10 I9 I8]
//   byte 1                      [ 0  I7  I6  I5  I4  I3  I2 I1]
//   byte 2                      [ 0  I0 Q12 Q11 Q10  Q9  Q8 Q7]
//   byte 3                      [ 0  Q6  Q5  Q4  Q3  Q2  Q1 Q0]
//
// Pure streaming: 4 B (CS16) .. 16 B (CF64) in, 4 B out per sample; four samples per thread and step through
// 128-bit accesses, a grid-stride loop over a grid that fills the SMs.  HBM-bound, nothing to stage.
#include "tx.cuh"

namespace cbl {

__device__ __forceinline__ uint32_t tx_word(int ii, int qq)
{
    const uint32_t i = (uint32_t)ii, q = (uint32_t)qq;
    const uint32_t s = 0xE0000000u                 // SOF | MODEM_TX_CTRL | COND_TX_CTRL, caribou_smi.c:684-686,704
                     | ((i & 0x1F00u) << 16)       // I12..I8 -> bits 28:24
                     | ((i & 0x00FEu) << 15)       // I7..I1  -> bits 22:16
                     | ((i & 0x0001u) << 14)       // I0      -> bit 14
                     | ((q & 0x1F80u) << 1)        // Q12..Q7 -> bits 13:8
                     | (q & 0x007Fu);              // Q6..Q0  -> bits 6:0
    return __byte_perm(s, 0u, 0x0123);             // __builtin_bswap32, caribou_smi.c:714
}

// (int16_t)(x * 4096): truncation toward zero to int32 and the low 16 bits of that, as the reference's cast
// compiles on x86-64; out of int32 range (undefined in C) it yields the low half of 0x80000000, i.e. 0.
__device__ __forceinline__ int tx_i16(float v)
{
    return (v > -2147483648.0f && v < 2147483648.0f) ? (int)(int16_t)__float2int_rz(v) : 0;
}
__device__ __forceinline__ int tx_i16(double v)
{
    return (v > -2147483648.0 && v < 2147483648.0) ? (int)(int16_t)__double2int_rz(v) : 0;
}

// one sample of format FMT at element index k -> (i, q) as the native int16 pair
template <int FMT>
__device__ __forceinline__ void tx_load1(const void* in, uint64_t k, int& i, int& q)
{
    if (FMT == TX_CS16) {
        const uint32_t w = reinterpret_cast<const uint32_t*>(in)[k];
        i = (int)(int16_t)(w & 0xFFFFu); q = (int)w >> 16;
    } else if (FMT == TX_CS8) {
        const uint16_t w = reinterpret_cast<const uint16_t*>(in)[k];
        i = (int)(int16_t)((int16_t)(int8_t)(w & 0xFFu) << 5);              // CaribouliteStream.cpp:236-237
        q = (int)(int16_t)((int16_t)(int8_t)(w >> 8) << 5);
    } else if (FMT == TX_CF32) {
        const float2 v = reinterpret_cast<const float2*>(in)[k];
        i = tx_i16(v.x * 4096.0f); q = tx_i16(v.y * 4096.0f);              // :202-207
    } else {
        const double2 v = reinterpret_cast<const double2*>(in)[k];
        i = tx_i16(v.x * 4096.0); q = tx_i16(v.y * 4096.0);                // :217-222
    }
}

template <int FMT>
__global__ void __launch_bounds__(kTxThreads) tx_pack_kernel(TxParams p)
{
    if (p.gate != nullptr && __ldg(p.gate) != 0) return;
    const uint64_t groups = p.n / 4;                       // 4 samples = one 128-bit store
    const uint64_t stride = (uint64_t)gridDim.x * kTxThreads;
    uint4* const out4 = reinterpret_cast<uint4*>(p.out);
    for (uint64_t g = (uint64_t)blockIdx.x * kTxThreads + threadIdx.x; g < groups; g += stride) {
        int i[4], q[4];
        if (FMT == TX_CS16) {
            const uint4 w = __ldcs(reinterpret_cast<const uint4*>(p.in) + g);
            const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
            for (int e = 0; e < 4; e++) { i[e] = (int)(int16_t)(ww[e] & 0xFFFFu); q[e] = (int)ww[e] >> 16; }
        } else if (FMT == TX_CS8) {
            const uint2 w = __ldcs(reinterpret_cast<const uint2*>(p.in) + g);
            const uint32_t ww[2] = {w.x, w.y};
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const uint32_t h = ww[e >> 1] >> ((e & 1) * 16);
                i[e] = (int)(int8_t)(h & 0xFFu) * 32;          // << 5 of a value that fits 13 bits
                q[e] = (int)(int8_t)((h >> 8) & 0xFFu) * 32;
            }
        } else if (FMT == TX_CF32) {
            const float4 a = __ldcs(reinterpret_cast<const float4*>(p.in) + 2 * g);
            const float4 b = __ldcs(reinterpret_cast<const float4*>(p.in) + 2 * g + 1);
            i[0] = tx_i16(a.x * 4096.0f); q[0] = tx_i16(a.y * 4096.0f);
            i[1] = tx_i16(a.z * 4096.0f); q[1] = tx_i16(a.w * 4096.0f);
            i[2] = tx_i16(b.x * 4096.0f); q[2] = tx_i16(b.y * 4096.0f);
            i[3] = tx_i16(b.z * 4096.0f); q[3] = tx_i16(b.w * 4096.0f);
        } else {
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const double2 v = __ldcs(reinterpret_cast<const double2*>(p.in) + 4 * g + e);
                i[e] = tx_i16(v.x * 4096.0); q[e] = tx_i16(v.y * 4096.0);
            }
        }
        __stcs(out4 + g, make_uint4(tx_word(i[0], q[0]), tx_word(i[1], q[1]), tx_word(i[2], q[2]), tx_word(i[3], q[3])));
    }
    // ragged tail (n % 4 samples), one thread each
    const uint64_t t = groups * 4 + (uint64_t)blockIdx.x * kTxThreads + threadIdx.x;
    if (t < p.n) {
        int i, q;
        tx_load1<FMT>(p.in, t, i, q);
        p.out[t] = tx_word(i, q);
    }
}

int launch_tx_pack(const TxParams& p, int num_sms, cudaStream_t st)
{
    if (p.n == 0) return 0;
    const uint64_t groups = p.n / 4 + 1;
    const uint64_t want = (groups + kTxThreads - 1) / kTxThreads;
    const uint64_t cap = (uint64_t)num_sms * 8;              // 8 x 256 threads fill an SM; the loop strides the rest
    const unsigned grid = (unsigned)(want < cap ? want : cap);
    switch (p.fmt) {
        case TX_CS16: tx_pack_kernel<TX_CS16><<<grid, kTxThreads, 0, st>>>(p); break;
        case TX_CS8:  tx_pack_kernel<TX_CS8><<<grid, kTxThreads, 0, st>>>(p); break;
        case TX_CF32: tx_pack_kernel<TX_CF32><<<grid, kTxThreads, 0, st>>>(p); break;
        case TX_CF64: tx_pack_kernel<TX_CF64><<<grid, kTxThreads, 0, st>>>(p); break;
        default: return (int)cudaErrorInvalidValue;
    }
    return (int)cudaGetLastError();
}

}  // namespace cbl
