// caribou_cuda_unpack_noalloc: kernel (a) for callers that must not allocate (code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The context-based entry point (caribou_cuda_unpack_submit) keeps the per-chunk offsets, the per-tile flag counts,
// their scan and the work list in device buffers owned by the context.  Here the same kernels (unpack.cu) get that
// scratch from ONE file-scope device array of fixed size, so the call makes no cudaMalloc / cudaFree / cudaHostAlloc
// and creates no stream or event: memsets and kernel launches on the caller's stream, nothing else.
//
//   * The array exists once per GPU that loads the library and stays until the process exits (kScratchBytes each).
//   * In host code the array's NAME is only a host-side stand-in, never a device address.  The address is taken with
//     cudaGetSymbolAddress once per device, with that device current (device_scratch below), and that pointer is
//     what the kernels of unpack.cu receive.  Device code of THIS file (noalloc_result_kernel) names the array
//     directly.
//   * Every call shares the array and finds the previous call's contents in it: what has to start at zero is cleared
//     on the stream at the head of each call.  Two calls that overlap on the GPU (different streams, no event
//     between them) race on it.
#include "ctx.h"
#include <limits.h>
#include <mutex>

namespace cbl {

constexpr size_t   kNoallocMaxBytes  = CARIBOU_CUDA_NOALLOC_MAX_BYTES;
constexpr uint32_t kNoallocMaxChunks = CARIBOU_CUDA_NOALLOC_MAX_CHUNKS;
constexpr uint32_t kNoallocMaxTiles  = (uint32_t)((kNoallocMaxBytes / 4 + kUnpackTileWords - 1) / kUnpackTileWords);

// layout of the scratch array (byte offsets; every block 64-byte aligned)
struct NoallocHead {               // cleared / preset at the head of every call
    int32_t  first_bad;            // min chunk index with no alignment (preset to 0x7f7f7f7f)
    int32_t  n_misaligned;
    int32_t  any_misaligned;
    int32_t  lost;
    uint64_t totals_out[2];        // n_invalid, n_sync of this call
    uint64_t totals_in[2];         // zeros
    uint32_t n_work;
    uint32_t pad[3];
};
static_assert(sizeof(NoallocHead) == 64, "head is one 64-byte block");
constexpr size_t kOffCounts  = sizeof(NoallocHead);
constexpr size_t kOffOffsets = kOffCounts + (size_t)kNoallocMaxTiles * 8;
constexpr size_t kOffWork    = kOffOffsets + (size_t)kNoallocMaxTiles * 8;
constexpr size_t kOffOffs    = kOffWork + (((size_t)kNoallocMaxTiles * 4 + 63) & ~size_t(63));
constexpr size_t kScratchBytes = kOffOffs + (size_t)kNoallocMaxChunks * 4;
static_assert(kScratchBytes == CARIBOU_CUDA_NOALLOC_SCRATCH_BYTES, "caribou_cuda.h states the size of the scratch array");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

// what caribou_cuda_wait computes on the host for an unpack submit, written to the caller's device memory
__global__ void noalloc_result_kernel(caribou_cuda_result* res, uint64_t nbytes, uint32_t chunk_bytes)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const NoallocHead* h = reinterpret_cast<const NoallocHead*>(g_noalloc_scratch);
    const uint64_t cb = chunk_bytes, n_chunks = (nbytes + cb - 1) / cb;
    const int32_t fb = h->first_bad;
    const bool lost = fb >= 0 && (uint64_t)fb < n_chunks;          // caribou_smi.c:665-668
    caribou_cuda_result r;
    r.n_samples = lost ? (int64_t)CARIBOU_CUDA_ERR_SYNC : (int64_t)((nbytes / cb) * (cb / 4) + (nbytes % cb) / 4);   // :677
    r.n_samples_b = 0;
    r.n_filtered = 0;
    r.n_invalid = h->totals_out[0];
    r.n_sync = h->totals_out[1];
    r.first_bad_chunk = lost ? fb : -1;
    r.n_misaligned_chunks = h->n_misaligned;
    *res = r;
}

struct NoallocDevice { unsigned char* scratch = nullptr; int num_sms = 0; bool ready = false; };
static NoallocDevice g_noalloc_dev[kMaxDevices];
static std::mutex g_noalloc_mu;

// device address of the scratch array on `device` (which must be current), looked up once per device
static unsigned char* device_scratch(int device, int* num_sms)
{
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    NoallocDevice& d = g_noalloc_dev[device];
    if (!d.ready) {
        void* p = nullptr;
        if (!cuda_ok(cudaGetSymbolAddress(&p, g_noalloc_scratch), "cudaGetSymbolAddress(g_noalloc_scratch)")) return nullptr;
        if (!p) { set_error("cudaGetSymbolAddress returned a null address"); return nullptr; }
        if (!cuda_ok(cudaDeviceGetAttribute(&d.num_sms, cudaDevAttrMultiProcessorCount, device), "cudaDeviceGetAttribute")) return nullptr;
        d.scratch = static_cast<unsigned char*>(p);
        d.ready = true;
    }
    *num_sms = d.num_sms;
    return d.scratch;
}

}  // namespace cbl

using namespace cbl;

extern "C" int caribou_cuda_unpack_noalloc(int device, void* cuda_stream, int channel, size_t chunk_bytes,
                                           const void* d_bytes, size_t nbytes, const caribou_cuda_unpack_out* out,
                                           caribou_cuda_result* d_result)
{
    if (!out || (!d_bytes && nbytes)) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    if (device < 0 || device >= kMaxDevices) { set_error("bad device index %d", device); return CARIBOU_CUDA_ERR; }
    if (channel != CARIBOU_CUDA_CH_S1G && channel != CARIBOU_CUDA_CH_HIF) { set_error("bad channel"); return CARIBOU_CUDA_ERR; }
    if (chunk_bytes == 0) chunk_bytes = CARIBOU_CUDA_NATIVE_BATCH_LEN;
    if (chunk_bytes % 16 != 0 || chunk_bytes < 32 || chunk_bytes > (1u << 30)) {
        set_error("chunk_bytes must be a multiple of 16 in [32, 2^30]");
        return CARIBOU_CUDA_ERR;
    }
    if (nbytes > kNoallocMaxBytes) {
        set_error("caribou_cuda_unpack_noalloc: %zu bytes exceed the fixed scratch (at most %zu bytes per call)", nbytes, kNoallocMaxBytes);
        return CARIBOU_CUDA_ERR;
    }
    const uint64_t n_chunks = (nbytes + chunk_bytes - 1) / chunk_bytes;
    if (n_chunks > kNoallocMaxChunks) {
        set_error("caribou_cuda_unpack_noalloc: %llu chunks exceed the fixed scratch (at most %u chunks per call)",
                  (unsigned long long)n_chunks, kNoallocMaxChunks);
        return CARIBOU_CUDA_ERR;
    }
    if (out->cs16_b || out->meta_b || out->cf32_b) { set_error("caribou_cuda_unpack_noalloc: single-channel captures only"); return CARIBOU_CUDA_ERR; }
    if ((uintptr_t)d_bytes % 16) { set_error("SMI byte stream must be 16-byte aligned in device memory"); return CARIBOU_CUDA_ERR; }
    CBL_CHECK(cudaSetDevice(device));
    int num_sms = 0;
    unsigned char* scratch = device_scratch(device, &num_sms);
    if (!scratch) return CARIBOU_CUDA_ERR;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);

    NoallocHead* head = reinterpret_cast<NoallocHead*>(scratch);
    uint64_t* tile_counts = reinterpret_cast<uint64_t*>(scratch + kOffCounts);
    int32_t* offs = reinterpret_cast<int32_t*>(scratch + kOffOffs);
    const uint64_t n_raw_words = ((uint64_t)nbytes + 3) / 4;
    const uint32_t n_tiles = (uint32_t)((n_raw_words + kUnpackTileWords - 1) / kUnpackTileWords);

    // the previous call's leftovers: head and tile counts back to zero (one memset, they are adjacent), first_bad to "none"
    CBL_CHECK(cudaMemsetAsync(scratch, 0, kOffCounts + (size_t)n_tiles * 8, st));
    CBL_CHECK(cudaMemsetAsync(&head->first_bad, 0x7f, 4, st));

    SubView sv;
    sv.words = reinterpret_cast<const uint32_t*>(d_bytes);
    sv.nbytes = nbytes;
    sv.n_sub_words = n_raw_words;
    sv.B = 0;
    sv.sub = 0;

    OffsetsParams fp;
    memset(&fp, 0, sizeof(fp));
    fp.sv = sv;
    fp.stream_stride_bytes = 0;
    fp.n_streams = 1;
    fp.chunk_bytes = (uint32_t)chunk_bytes;
    fp.n_chunks = (uint32_t)n_chunks;
    fp.offs = offs;
    fp.offs_stride = kNoallocMaxChunks;
    fp.first_bad = &head->first_bad;
    fp.first_bad_stride = 1;
    fp.n_misaligned = &head->n_misaligned;
    fp.any_misaligned = &head->any_misaligned;
    fp.lost = &head->lost;
    launch_find_offsets(fp, st);

    if (n_tiles) {
        UnpackParams p;
        memset(&p, 0, sizeof(p));
        p.words = sv.words;
        p.nbytes = nbytes;
        p.B = 0;
        p.chunk_bytes = (uint32_t)chunk_bytes;
        p.spc = p.chunk_bytes / 4;
        p.n_subs = 1;
        p.sv[0] = sv;
        p.layout[0] = channel == CARIBOU_CUDA_CH_HIF ? 1 : 0;
        p.offs[0] = offs;
        p.first_bad[0] = &head->first_bad;
        p.cs16[0] = reinterpret_cast<uint32_t*>(out->cs16);
        p.meta[0] = reinterpret_cast<uint8_t*>(out->meta);
        p.cf32[0] = reinterpret_cast<float2*>(out->cf32);
        p.vec_ok = ((uintptr_t)p.cs16[0] % 16 || (uintptr_t)p.meta[0] % 4 || (uintptr_t)p.cf32[0] % 16) ? 0 : 1;
        p.want_flags = 1;
        p.inv_idx = out->invalid_idx; p.inv_cap = out->invalid_cap;
        p.syn_idx = out->sync_idx;    p.syn_cap = out->sync_cap;
        p.index_base = 0;
        p.tile_counts = tile_counts;
        p.tile_offsets = reinterpret_cast<uint64_t*>(scratch + kOffOffsets);
        p.work = reinterpret_cast<uint32_t*>(scratch + kOffWork);
        p.n_work = &head->n_work;
        p.totals_in = head->totals_in;
        p.totals_out = head->totals_out;
        p.n_tiles = n_tiles;
        launch_unpack(p, num_sms, st);
    }
    if (d_result) noalloc_result_kernel<<<1, 32, 0, st>>>(d_result, (uint64_t)nbytes, (uint32_t)chunk_bytes);
    CBL_CHECK(cudaGetLastError());
    return CARIBOU_CUDA_OK;
}
