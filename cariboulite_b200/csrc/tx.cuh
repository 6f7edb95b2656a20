// Kernel (c): CS16 / CS8 / CF32 / CF64 samples -> SMI TX words (tx.cu).
#pragma once
#include "common.cuh"

namespace cbl {

// same values as caribou_cuda_format (include/caribou_cuda.h)
enum { TX_CF32 = 0, TX_CS16 = 1, TX_CS8 = 2, TX_CF64 = 3 };

struct TxParams {
    const void* in;          // n elements of the format, device memory, 16-byte aligned
    uint32_t*   out;         // n words, 16-byte aligned
    uint64_t    n;
    int         fmt;
    const int32_t* gate;     // optional: no-op when *gate != 0
};

constexpr int kTxThreads = 256;

// returns cudaError_t as int
int launch_tx_pack(const TxParams& p, int num_sms, cudaStream_t st);

}  // namespace cbl
