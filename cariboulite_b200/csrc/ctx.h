// Internal context of libcaribou_cuda (host side).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/caribou_cuda.h"
#include "iir.cuh"
#include "tables.h"
#include "unpack.cuh"
#include "tx.cuh"

namespace cbl {

void set_error(const char* fmt, ...);
bool cuda_ok(cudaError_t e, const char* what);
#define CBL_CHECK(call)                                  \
    do {                                                 \
        if (!::cbl::cuda_ok((call), #call)) return CARIBOU_CUDA_ERR; \
    } while (0)

// device buffer that only ever grows
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    bool reserve(size_t n);
    void release();
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// small block of flags/counters mirrored to pinned host memory after every submit
struct Counters {
    int32_t  first_bad[2];        // per sub-stream, min over streams is not needed: stream 0.. handled per launch
    int32_t  n_misaligned;
    int32_t  any_misaligned;
    uint64_t totals[2];           // n_invalid, n_sync
};

struct PendingRx {                // what caribou_cuda_rx_submit has to redo if a chunk was misaligned
    bool active = false;
    const void* bytes = nullptr; size_t nbytes = 0; size_t stride = 0; int mem = 0;
    void* filt_out = nullptr; size_t filt_stride = 0;
    const void* d_bytes = nullptr; void* d_out = nullptr;
};

}  // namespace cbl

struct caribou_cuda_ctx {
    caribou_cuda_config cfg;
    int device = 0;
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    uint64_t launches = 0;
    bool profiling = false;
    std::vector<cudaEvent_t> prof_ev;   // ring of (start, stop) pairs around the dominant kernel
    uint64_t prof_n = 0;                // pairs recorded since the last read

    // filter
    bool has_filter = false;
    cbl::FilterPlan plan;
    int out_mode = 0;
    size_t out_elem = 8;
    double* d_plane = nullptr;
    double* d_mtile_pow = nullptr;
    double* d_state = nullptr;        // [n_streams][2S] kernel coordinates
    double* d_state_next = nullptr;
    uint64_t pos = 0;                 // samples consumed per stream since create/reset
    double in_scale = 1.0;            // notional filter input = raw sample value * in_scale (1/4096 for integer
                                      // input under F32 parity); the state API speaks in notional units
    cbl::DevBuf desc_agg, desc_incl, desc_flag;
    uint32_t* d_ticket = nullptr;     // [4] tickets (iir, unpack, spare)

    // unpack
    cbl::DevBuf offs;                 // [n_streams][2][n_chunks] int32
    cbl::DevBuf desc_inv, desc_syn;
    cbl::Counters* d_counters = nullptr;
    cbl::Counters* h_counters = nullptr;   // pinned
    int32_t* d_first_bad = nullptr;   // [n_streams][2]
    int32_t* h_first_bad = nullptr;   // pinned mirror

    // staging for host memory
    cbl::DevBuf stage_in, stage_cs16[2], stage_meta[2], stage_cf32[2], stage_filt, stage_inv, stage_syn, scratch_cs16;

    // window of streams the launch helpers work on (pipelined host-buffer submits walk the batch group by group)
    uint32_t win0 = 0, win_cnt = 0;
    // copy engines' streams and events of the pipelined host-buffer path (created on first use)
    cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;
    std::vector<cudaEvent_t> pipe_ev;
    double* d_state_backup = nullptr;   // filter state before a pipelined submit (restored if it has to be redone)
    bool pipelined_last = false;

    // synth
    int16_t* d_tone = nullptr;

    // last submit, for caribou_cuda_wait
    int64_t last_ret = 0;
    bool last_is_unpack = false;
    size_t last_nbytes = 0;
    int64_t last_n_filtered = 0;
    cbl::PendingRx pending;
};
