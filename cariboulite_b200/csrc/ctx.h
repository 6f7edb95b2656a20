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

// page-locked host buffer that only ever grows (contents are not kept)
struct HostBuf {
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

struct PendingRx {                // one queued caribou_cuda_rx_submit: what has to be redone for a capture that
    int32_t seq = 0;              // turned out misaligned (position in the queue since the last resolve)
    const void* bytes = nullptr; size_t nbytes = 0; size_t stride = 0; int mem = 0;
    void* filt_out = nullptr; size_t filt_stride = 0;
    const void* d_bytes = nullptr; void* d_out = nullptr;
};
constexpr size_t kMaxQueuedRx = 64;   // rx submits queued without a wait; the next one resolves the queue first

struct Shard;                     // multi-GPU time sharding (shard.cu)
void shard_release(Shard* s);
int shard_flush(caribou_cuda_ctx* c);   // enqueue the hand-off a peer-mode sharded call left for later (no-op otherwise)

// ---- launch sequences shared by api.cu, compat.cu and shard.cu
// Filters n samples per stream (device memory) of the streams in the context's window.  `state` is the buffer
// the streams start from and, with commit, end in (kernel coordinates; nullptr: the context's own d_state);
// max_tiles > 0 limits the launch to the first max_tiles tiles of every stream (fix-up pass, no commit).
// Tiles of stream s are skipped when gate[s * gate_stride] != 0.
int run_iir(caribou_cuda_ctx* c, int in_kind, const void* d_in, size_t n, size_t in_stride, void* d_out,
            size_t out_stride, const int32_t* gate, uint32_t gate_stride, bool commit, uint32_t max_tiles,
            double* state);
// alignment search of the window's streams; returns the chunk count (>= 0) or an error code.
// with_gates: also raise the sticky per-stream gates the fused launches look at
int run_find_offsets(caribou_cuda_ctx* c, const void* d_bytes, size_t nbytes, size_t stride_bytes, bool with_gates, size_t iir_n = 0);
// the plan run_iir takes for a submit of n samples per stream (the short-run plan when the long-run one would leave most
// of the GPU without a tile)
const FilterPlan& pick_plan(const caribou_cuda_ctx* c, size_t n);
int mirror_counters(caribou_cuda_ctx* c);
int ensure_scale(caribou_cuda_ctx* c, int in_kind);
// blocks until every queued rx submit has been carried out (misaligned captures redone on the slow path)
int resolve_queue(caribou_cuda_ctx* c);
// caribou_cuda_filter_submit on device memory with an optional gate (a no-op on the device when *gate != 0)
int filter_device(caribou_cuda_ctx* c, int in_kind, const void* in, size_t n, void* out, const int32_t* gate);

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
    // rerun plans only: the same filter with short runs (kRerunSmallLT) for submits too small to fill the GPU with the
    // long-run plan's tiles (an MTU of 131072 samples is 16 tiles of 32 x 256-sample runs, but 64 of 32 x 64)
    bool has_small = false;
    cbl::FilterPlan plan_small;
    double* d_plane_small = nullptr;
    double* d_mtile_pow_small = nullptr;
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
    uint8_t* d_ctrl = nullptr;        // control block: counters, first bad chunks, per-stream flags (pointers below)
    uint8_t* h_ctrl = nullptr;        // its page-locked mirror
    uint8_t* h_ctrl_dev = nullptr;    // the mirror as the device sees it (kernels may write it directly)
    bool mirror_in_tail = false;      // request: the tail launch of the next run_iir copies d_ctrl into the mirror
    struct RrPow { uint64_t len = 0; double m[cbl::kMaxS * cbl::kMaxS]; } rr_pow[2];   // A^len of the main / short-run plan (rerun kernel, run_iir)
    const void* precleared_ptr = nullptr;   // descriptors (and the ticket) an earlier kernel of the submit has emptied for
    uint64_t precleared_doubles = 0;        // the next run_iir: it skips its memsets when they are exactly the ones it needs
    size_t ctrl_bytes = 0;
    cbl::Counters* d_counters = nullptr;
    cbl::Counters* h_counters = nullptr;   // pinned
    int32_t* d_first_bad = nullptr;   // [n_streams][2]
    int32_t* h_first_bad = nullptr;   // pinned mirror
    // per-stream verdicts of the fused path: [0..ns) sticky gate, [ns..2ns) first submit that raised it,
    // [2ns..3ns) misaligned chunks of the last submit; [3ns] sticky "a chunk had no alignment" flag the shims
    // clear before a read and gate the filter with (a read that lost sync delivers nothing, CaribouliteStream.cpp:266-276)
    int32_t* d_sflags = nullptr;
    int32_t* h_sflags = nullptr;      // pinned mirror

    // staging for host memory
    cbl::DevBuf stage_in, stage_cs16[2], stage_meta[2], stage_cf32[2], stage_filt, stage_inv, stage_syn, scratch_cs16;

    cbl::HostBuf h_bounce;            // page-locked bounce buffer of the pageable-output unpack path

    // window of streams the launch helpers work on (pipelined host-buffer submits walk the batch group by group)
    uint32_t win0 = 0, win_cnt = 0;
    // copy engines' streams and events of the pipelined host-buffer path (created on first use)
    cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;
    std::vector<cudaEvent_t> pipe_ev;

    // synth
    int16_t* d_tone = nullptr;

    // last submit, for caribou_cuda_wait
    int64_t last_ret = 0;
    bool last_is_unpack = false;
    bool last_is_shard = false;       // the last submit was a sharded call: wait reports -3 for a misaligned shard
    size_t last_nbytes = 0;
    int64_t last_n_filtered = 0;
    std::vector<cbl::PendingRx> queue;                 // rx submits since the last resolve
    std::vector<caribou_cuda_stream_result> sres;      // per-capture results of the last rx submit
    bool sres_valid = false;

    cbl::Shard* shard = nullptr;                       // multi-GPU time sharding (shard.cu)
};
