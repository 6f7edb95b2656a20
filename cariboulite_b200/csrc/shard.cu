// Multi-GPU time sharding of ONE long stream (BASELINE.json configs[4]) with the hand-off on the
// device: no stream synchronisation and no host copy between the pass over the shard and the fix-up.
//
//   rank g of G owns samples [g n, (g+1) n) of every call
//   pass 1    fused kernel over the shard from ZERO state (all ranks at once) -> outputs, end vector f_g
//   hand-off  the 4K doubles per stream travel between the GPUs on the context's stream:
//               all-gather (default): ncclAllGather of every f_j, then each rank composes
//                   s_in(g) = sum_{j<g} M^(g-1-j) f_j + M^g carry ,  carry' = sum_{j<G} M^(G-1-j) f_j + M^G carry
//               chain:                ncclRecv s_in(g) from g-1, s_in(g+1) = M s_in(g) + f_g, ncclSend to g+1,
//                                     the last rank broadcasts the carry (G-1 serial hops)
//               peer (default when the GPUs can map each other's memory): every rank's exchange buffer is opened in
//                   every other rank's process (CUDA IPC, handles all-gathered once through NCCL); a one-CTA kernel
//                   stores f_g straight into all peers' buffers over NVLink and releases a sequence flag, the
//                   composition kernel acquires the flags - no collective launch on the path at all.  The
//                   composition and the fix-up of a call are queued one call late (after the pass over the next
//                   shard): nobody waits for the slowest rank's pass
//             M = A^n in kernel coordinates; its powers are computed on the host once per shard length
//   fix-up    the leading tiles a non-zero incoming state can still change above 2^-60 are filtered again from
//             s_in(g), read from device memory; the context's state becomes carry', so calls chain like on one GPU
// The reference behaviour this reproduces: one stream whose filter objects are never reset
// (LC/soapy_api/CaribouliteStream.hpp:124-131) fed MTU after MTU (LC/caribou_smi/caribou_smi.c:643-679).
//
// NCCL is bound at run time (dlopen "libnccl.so.2"): a process that already carries an NCCL - PyTorch's bundled
// one - shares it instead of getting a second copy with different symbol versions, and the library still loads
// on a box without NCCL (everything but the communicator then works).  cariboulite_b200/shard.py keeps the
// host-side (numpy, long double) statement of the same composition for the CPU tests.
#include <dlfcn.h>
#include <nccl.h>
#include <stdlib.h>
#include <string.h>
#include <mutex>
#include <new>
#include <vector>
#include "ctx.h"

namespace cbl {

// ------------------------------------------------------------------ NCCL, bound at run time
struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

static NcclApi* nccl_api()
{
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) return;
        auto sym = [&](const char* n) { return dlsym(h, n); };
        api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
        api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
        api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
        api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
        api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
        api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
        api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
        api.Broadcast = reinterpret_cast<decltype(api.Broadcast)>(sym("ncclBroadcast"));
        api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
        api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
        api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
        if (api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllGather && api.AllReduce && api.Send && api.Recv &&
            api.Broadcast && api.GroupStart && api.GroupEnd && api.GetErrorString)
            api.handle = h;
    });
    if (!api.handle) { set_error("NCCL is not available (dlopen of libnccl.so.2 failed or symbols are missing)"); return nullptr; }
    return &api;
}

static bool nccl_ok(NcclApi* a, ncclResult_t r, const char* what)
{
    if (r == ncclSuccess) return true;
    set_error("NCCL error %s in %s", a->GetErrorString(r), what);
    return false;
}
#define CBL_NCCL(a, call)                                                \
    do {                                                                 \
        if (!::cbl::nccl_ok((a), (call), #call)) return CARIBOU_CUDA_ERR; \
    } while (0)

// ------------------------------------------------------------------ per-context sharding state
struct ShardCall {                  // what one sharded call filters, and its sequence number (peer mode)
    int in_kind = 0; const void* in = nullptr; size_t n = 0, in_stride = 0; void* out = nullptr; size_t out_stride = 0;
    unsigned long long seq = 0;
};
struct Shard {
    int rank = 0, world = 1, mode = 0;
    ncclComm_t comm = nullptr;
    bool own_comm = false;
    size_t per_rank = 0;            // doubles per rank in the gather buffer: n_streams * 2S
    double* d_f = nullptr;          // [world][per_rank]  zero-state end vectors (slot `rank` is this rank's)
    double* d_sin = nullptr;        // [per_rank]  state entering this rank's shard
    double* d_work = nullptr;       // [per_rank]  scratch state of a whole-shard fix-up
    double* d_carry = nullptr;      // [per_rank]  state after the last rank's shard
    double* d_mpow = nullptr;       // [world + 1][S * S]  powers of M = A^n (kernel coordinates)
    uint64_t mpow_n = 0;            // shard length the powers were computed for (0: none)
    // peer exchange (mode 2): d_px = [2 parities][world][per_rank] doubles, then [2][world] sequence flags; peers store into it
    bool peer_ok = false;
    double* d_px = nullptr;
    double** d_peers = nullptr;     // [world] every rank's d_px in this process's address space
    std::vector<void*> peer_maps;   // what cudaIpcOpenMemHandle returned (to close)
    int* err_host = nullptr;        // device view of the control block's "a peer never arrived" slot in page-locked host memory
    unsigned long long seq = 0;     // sharded calls so far
    bool committed = false;         // the composition kernel of this call has written the carry into the context's state
    cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};   // profiling: start, after pass 1, end
    bool ev_valid = false;
    // the call between begin and finish
    bool open = false;
    ShardCall cur;
    // peer mode: the composition and the fix-up of a call are enqueued one call late, after the pass over the NEXT
    // shard - by then every peer's end vector has long arrived, so no rank waits for the slowest one's pass any more
    // (caribou_cuda_wait and every other entry point flush the pending hand-off first)
    bool deferred = false;
    ShardCall pend;
    int32_t* d_gate2 = nullptr;     // [2] per call parity: the "shard was misaligned" verdict of that call's alignment search
};

void shard_release(Shard* s)
{
    if (!s) return;
    if (s->comm && s->own_comm) { NcclApi* a = nccl_api(); if (a) a->CommDestroy(s->comm); }
    if (s->d_f) cudaFree(s->d_f);
    if (s->d_sin) cudaFree(s->d_sin);
    if (s->d_work) cudaFree(s->d_work);
    if (s->d_carry) cudaFree(s->d_carry);
    if (s->d_mpow) cudaFree(s->d_mpow);
    for (void* m : s->peer_maps) if (m) cudaIpcCloseMemHandle(m);
    if (s->d_px) cudaFree(s->d_px);
    if (s->d_peers) cudaFree(s->d_peers);
    if (s->d_gate2) cudaFree(s->d_gate2);
    for (cudaEvent_t e : s->ev) if (e) cudaEventDestroy(e);
    delete s;
}

// ------------------------------------------------------------------ composition kernels (one small CTA)
// thread t = (stream, rail, row): one component of the state vectors
struct ComposeParams {
    int S, world, rank;
    uint32_t n_comp;            // n_streams * 2S
    const double* f;            // [world][n_comp]
    const double* carry;        // [n_comp]  state entering shard 0 (the context's state)
    const double* mpow;         // [world + 1][S * S]
    double* s_in;               // [n_comp]
    double* carry_out;          // [n_comp]
    // peer mode: sequence flags of the ranks' end vectors (nullptr: f is complete when the kernel starts)
    const unsigned long long* flags;
    unsigned long long seq;
    int* err;
    int* err_host;              // the same verdict in the context's page-locked control block (what wait() reads)
    // single-CTA launches also commit the carry into the context's state (one launch less on the hand-off path),
    // unless *gate says the shard was misaligned
    double* commit;
    const int32_t* gate;
    // the fix-up launch that follows needs its look-back descriptors empty and its ticket zero: done here instead of
    // by two memsets on the stream (clear_n doubles of kDescEmpty; nullptr: the launch clears them itself)
    unsigned long long* clear_desc;
    uint32_t clear_n;
    uint32_t* clear_ticket;
};

// one system-scope acquire load / release store (the flags live in memory other GPUs write)
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// peer mode: this rank's end vector into slot `rank` of EVERY rank's exchange buffer (its own included), then the flag
// (one CTA)
__device__ __forceinline__ void shard_push(double* const* peers, const double* f_own, uint32_t n_comp, int world, int rank,
                                           size_t par_off, size_t flag_off, unsigned long long seq)
{
    for (int j = 0; j < world; j++) {
        double* dst = peers[j] + par_off + (size_t)rank * n_comp;
        for (uint32_t t = threadIdx.x; t < n_comp; t += blockDim.x) dst[t] = f_own[t];
    }
    __threadfence_system();
    __syncthreads();
    for (int j = threadIdx.x; j < world; j += blockDim.x)
        st_release_sys(reinterpret_cast<unsigned long long*>(peers[j] + flag_off) + rank, seq);
}

// peer mode with the hand-off one call late: the push right after the pass, and the verdict of the call's alignment search
// put aside (the next call's search resets the context's flag before this call's composition and fix-up run)
__global__ void shard_push_kernel(double* const* peers, const double* f_own, uint32_t n_comp, int world, int rank,
                                  size_t par_off, size_t flag_off, unsigned long long seq, const int32_t* gate_src, int32_t* gate_dst)
{
    if (threadIdx.x == 0 && gate_dst != nullptr) *gate_dst = gate_src != nullptr ? *gate_src : 0;
    shard_push(peers, f_own, n_comp, world, rank, par_off, flag_off, seq);
}

// The hand-off in one launch: wait for the peers' vectors (peer mode), compose s_in and the carry, empty the descriptors
// and the ticket of the fix-up launch, commit the carry (single-CTA launches).
__global__ void shard_compose_kernel(const ComposeParams p)
{
    if (p.clear_desc != nullptr) {
        for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < p.clear_n; t += gridDim.x * blockDim.x) p.clear_desc[t] = kDescEmpty;
        if (blockIdx.x == 0 && threadIdx.x == 0) *p.clear_ticket = 0;
    }
    if (p.flags != nullptr) {
        // wait until every rank's end vector of this call has landed (bounded: a peer that died must not hang the GPU)
        for (int j = threadIdx.x; j < p.world; j += blockDim.x) {
            unsigned long long spins = 0;
            while (ld_acquire_sys(p.flags + j) != p.seq) {
                __nanosleep(64);
                if (++spins > (1ull << 22)) { *p.err = 1; *p.err_host = 1; break; }     // a few seconds
            }
        }
        __syncthreads();
    }
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < p.n_comp; t += gridDim.x * blockDim.x) {
        const int row = (int)(t % (uint32_t)p.S);
        const uint32_t vec = t - (uint32_t)row;                 // first component of this (stream, rail) vector
        // M^e applied to vector v, component `row`
        auto mv = [&](int e, const double* v) {
            const double* m = p.mpow + (size_t)e * p.S * p.S + (size_t)row * p.S;
            double acc = 0.0;
            for (int j = 0; j < p.S; j++) acc = fma(m[j], v[vec + j], acc);
            return acc;
        };
        double sin = mv(p.rank, p.carry);
        for (int j = 0; j < p.rank; j++) sin += mv(p.rank - 1 - j, p.f + (size_t)j * p.n_comp);
        double co = mv(p.world, p.carry);
        for (int j = 0; j < p.world; j++) co += mv(p.world - 1 - j, p.f + (size_t)j * p.n_comp);
        p.s_in[t] = sin;
        p.carry_out[t] = co;
    }
    if (p.commit != nullptr && (p.gate == nullptr || *p.gate == 0)) {      // gridDim.x == 1: every read of the old carry is done
        __syncthreads();
        for (uint32_t t = threadIdx.x; t < p.n_comp; t += blockDim.x) p.commit[t] = p.carry_out[t];
    }
}

// chain mode: s_out = M s_in + f (one hop)
__global__ void shard_hop_kernel(int S, uint32_t n_comp, const double* m, const double* s_in, const double* f, double* s_out)
{
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < n_comp; t += gridDim.x * blockDim.x) {
        const int row = (int)(t % (uint32_t)S);
        const uint32_t vec = t - (uint32_t)row;
        double acc = f[t];
        for (int j = 0; j < S; j++) acc = fma(m[(size_t)row * S + j], s_in[vec + j], acc);
        s_out[t] = acc;
    }
}

static int shard_alloc(caribou_cuda_ctx* c, int rank, int world)
{
    if (!c || !c->has_filter) { set_error("time sharding needs a context with a filter"); return CARIBOU_CUDA_ERR; }
    if (world < 1 || rank < 0 || rank >= world || world > 4096) { set_error("bad rank/world"); return CARIBOU_CUDA_ERR; }
    CBL_CHECK(cudaSetDevice(c->device));
    if (resolve_queue(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (c->shard) { shard_release(c->shard); c->shard = nullptr; }
    Shard* s = new (std::nothrow) Shard();
    if (!s) { set_error("out of memory"); return CARIBOU_CUDA_ERR; }
    c->shard = s;
    s->rank = rank; s->world = world;
    const int S = c->plan.S;
    s->per_rank = (size_t)c->cfg.n_streams * 2 * S;
    if (!cuda_ok(cudaMalloc(&s->d_f, (size_t)world * s->per_rank * 8), "cudaMalloc") ||
        !cuda_ok(cudaMalloc(&s->d_sin, s->per_rank * 8), "cudaMalloc") ||
        !cuda_ok(cudaMalloc(&s->d_work, s->per_rank * 8), "cudaMalloc") ||
        !cuda_ok(cudaMalloc(&s->d_carry, s->per_rank * 8), "cudaMalloc") ||
        !cuda_ok(cudaMalloc(&s->d_mpow, (size_t)(world + 1) * S * S * 8), "cudaMalloc"))
        return CARIBOU_CUDA_ERR;
    CBL_CHECK(cudaMemsetAsync(s->d_f, 0, (size_t)world * s->per_rank * 8, c->stream));
    {
        // a composition kernel that gives up on a peer says so in the device control block (sticky, mirrored by every
        // later call) and straight in the page-locked host copy wait() reads (no copy on the hand-off path for it)
        void* dp = nullptr;
        CBL_CHECK(cudaHostGetDevicePointer(&dp, c->h_ctrl, 0));
        s->err_host = reinterpret_cast<int*>(static_cast<uint8_t*>(dp) +
                                             (reinterpret_cast<uint8_t*>(c->h_sflags + 3 * (size_t)c->cfg.n_streams + 1) - reinterpret_cast<uint8_t*>(c->h_ctrl)));
    }
    for (cudaEvent_t& e : s->ev) CBL_CHECK(cudaEventCreate(&e));
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    return CARIBOU_CUDA_OK;
}

// Peer exchange set-up (once per communicator): every rank's exchange buffer is mapped into every other rank's process.
// All ranks take the same decision: the outcome of the mapping is min-reduced over the communicator.
static int shard_peer_setup(caribou_cuda_ctx* c)
{
    Shard* s = c->shard;
    NcclApi* a = nccl_api();
    if (!a || !s->comm || s->world < 2 || getenv("CARIBOU_CUDA_NO_PEER")) return CARIBOU_CUDA_OK;
    const size_t doubles = 2 * (size_t)s->world * s->per_rank + 2 * (size_t)s->world;
    CBL_CHECK(cudaMalloc(&s->d_px, doubles * 8));
    CBL_CHECK(cudaMemsetAsync(s->d_px, 0, doubles * 8, c->stream));
    CBL_CHECK(cudaMalloc(&s->d_peers, (size_t)s->world * sizeof(double*)));
    CBL_CHECK(cudaMalloc(&s->d_gate2, 2 * sizeof(int32_t)));
    CBL_CHECK(cudaMemsetAsync(s->d_gate2, 0, 2 * sizeof(int32_t), c->stream));
    // handles: [world] x (64-byte handle + 1 int "ok so far")
    struct Slot { cudaIpcMemHandle_t h; int ok; int pad[15]; };
    static_assert(sizeof(Slot) == 128, "slot layout");
    std::vector<Slot> slots((size_t)s->world);
    memset(slots.data(), 0, slots.size() * sizeof(Slot));
    Slot& mine = slots[(size_t)s->rank];
    mine.ok = cudaIpcGetMemHandle(&mine.h, s->d_px) == cudaSuccess ? 1 : 0;
    cudaGetLastError();
    Slot* d_slots = nullptr;
    CBL_CHECK(cudaMalloc(&d_slots, slots.size() * sizeof(Slot)));
    auto fail = [&](void) { cudaFree(d_slots); return CARIBOU_CUDA_ERR; };
    if (!cuda_ok(cudaMemcpyAsync(d_slots + s->rank, &mine, sizeof(Slot), cudaMemcpyHostToDevice, c->stream), "H2D") ||
        !nccl_ok(a, a->AllGather(d_slots + s->rank, d_slots, sizeof(Slot), ncclChar, s->comm, c->stream), "ncclAllGather(handles)") ||
        !cuda_ok(cudaMemcpyAsync(slots.data(), d_slots, slots.size() * sizeof(Slot), cudaMemcpyDeviceToHost, c->stream), "D2H") ||
        !cuda_ok(cudaStreamSynchronize(c->stream), "cudaStreamSynchronize")) return fail();
    int ok = 1;
    std::vector<double*> ptrs((size_t)s->world, nullptr);
    for (int j = 0; j < s->world && ok; j++) {
        if (!slots[(size_t)j].ok) { ok = 0; break; }
        if (j == s->rank) { ptrs[(size_t)j] = s->d_px; continue; }
        void* m = nullptr;
        if (cudaIpcOpenMemHandle(&m, slots[(size_t)j].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = 0; break; }
        s->peer_maps.push_back(m);
        ptrs[(size_t)j] = static_cast<double*>(m);
    }
    // one verdict for all ranks
    int* d_ok = reinterpret_cast<int*>(d_slots);
    if (!cuda_ok(cudaMemcpyAsync(d_ok, &ok, 4, cudaMemcpyHostToDevice, c->stream), "H2D") ||
        !nccl_ok(a, a->AllReduce(d_ok, d_ok, 1, ncclInt, ncclMin, s->comm, c->stream), "ncclAllReduce(ok)") ||
        !cuda_ok(cudaMemcpyAsync(&ok, d_ok, 4, cudaMemcpyDeviceToHost, c->stream), "D2H") ||
        !cuda_ok(cudaStreamSynchronize(c->stream), "cudaStreamSynchronize")) return fail();
    cudaFree(d_slots);
    if (!ok) {       // some pair of GPUs cannot map each other: everybody stays on the NCCL modes
        for (void* m : s->peer_maps) if (m) cudaIpcCloseMemHandle(m);
        s->peer_maps.clear();
        return CARIBOU_CUDA_OK;
    }
    CBL_CHECK(cudaMemcpyAsync(s->d_peers, ptrs.data(), ptrs.size() * sizeof(double*), cudaMemcpyHostToDevice, c->stream));
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    s->peer_ok = true;
    s->mode = 2;
    return CARIBOU_CUDA_OK;
}

// powers M^0 .. M^world of the shard matrix for shards of n samples, uploaded once per n
static int shard_powers(caribou_cuda_ctx* c, uint64_t n)
{
    Shard* s = c->shard;
    if (s->mpow_n == n) return CARIBOU_CUDA_OK;
    const int S = c->plan.S;
    std::vector<double> m((size_t)(s->world + 1) * S * S);
    for (int e = 0; e <= s->world; e++) plan_power_kernel(c->plan, n * (uint64_t)e, m.data() + (size_t)e * S * S);   // A^(n e), e = 0: I
    // pageable source: ordered on the context's stream and waited for (once per shard length)
    CBL_CHECK(cudaMemcpyAsync(s->d_mpow, m.data(), m.size() * 8, cudaMemcpyHostToDevice, c->stream));
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    s->mpow_n = n;
    return CARIBOU_CUDA_OK;
}

static int shard_begin(caribou_cuda_ctx* c, int in_kind, const void* in, size_t n, size_t in_stride, void* out, size_t out_stride)
{
    if (!c || !c->shard) { set_error("call caribou_cuda_shard_init first"); return CARIBOU_CUDA_ERR; }
    if (!in || !out || in_kind < 0 || in_kind > 2 || n == 0) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    CBL_CHECK(cudaSetDevice(c->device));
    Shard* s = c->shard;
    const FilterPlan& P = c->plan;
    if (n % (size_t)P.D != 0) { set_error("n_samples must be a multiple of the decimation factor"); return CARIBOU_CUDA_ERR; }
    if (!c->queue.empty() && resolve_queue(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (ensure_scale(c, in_kind) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (s->mpow_n != n && shard_flush(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;     // a pending hand-off needs the old powers
    if (shard_powers(c, n) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    const size_t ns = (size_t)c->cfg.n_streams;
    if (ns == 1) { in_stride = (n + 3) & ~size_t(3); out_stride = n / P.D; }
    c->last_is_unpack = false;
    c->last_is_shard = true;
    c->sres_valid = false;
    c->last_nbytes = n * 4;
    c->last_ret = (int64_t)n;
    c->last_n_filtered = (int64_t)(n / P.D);
    const bool prof = c->profiling;
    s->ev_valid = false;
    if (prof) CBL_CHECK(cudaEventRecord(s->ev[0], c->stream));
    // pass 1 from zero state; a shard of raw words must consist of aligned chunks - the alignment search raises the
    // batch-level flag otherwise, every kernel of the call is then a no-op and wait reports -3
    double* f_own = s->d_f + (size_t)s->rank * s->per_rank;
    CBL_CHECK(cudaMemsetAsync(f_own, 0, s->per_rank * 8, c->stream));
    const int32_t* gate = nullptr;
    if (in_kind == IN_WORDS) {
        const size_t es = 4;
        const int mc = run_find_offsets(c, in, n * es, in_stride * es, false, n);
        if (mc < 0) return mc;
        gate = &c->d_counters->any_misaligned;
    } else {
        CBL_CHECK(cudaMemsetAsync(c->d_counters, 0, sizeof(Counters), c->stream));
    }
    c->mirror_in_tail = true;                   // pass 1's last launch mirrors the control block
    if (run_iir(c, in_kind, in, n, in_stride, out, out_stride, gate, 0, true, 0, f_own) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (c->mirror_in_tail) {
        c->mirror_in_tail = false;
        if (mirror_counters(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    }
    if (prof) CBL_CHECK(cudaEventRecord(s->ev[1], c->stream));
    s->open = true;
    s->cur.in_kind = in_kind; s->cur.in = in; s->cur.n = n; s->cur.in_stride = in_stride; s->cur.out = out; s->cur.out_stride = out_stride;
    return CARIBOU_CUDA_OK;
}

// what the fix-up of a call launches: the leading tiles only (partial), or the whole shard again
struct FixPlan { bool partial; uint32_t max_tiles; };
static FixPlan shard_fix_plan(caribou_cuda_ctx* c, const ShardCall& k)
{
    const size_t fix = caribou_cuda_fixup_samples(c, k.n);
    const int LTn = pick_plan(c, k.n).LT;               // the plan run_iir will take for this shard
    const uint32_t tile_samples = 32u * (uint32_t)LTn;
    const uint32_t max_tiles = (uint32_t)((fix + tile_samples - 1) / tile_samples);
    const uint64_t tiles_ps = (k.n / LTn + 31) / 32;
    FixPlan f;
    f.partial = !(max_tiles == 0 || max_tiles >= tiles_ps);
    f.max_tiles = max_tiles;
    return f;
}

// the "shard was misaligned" flag the kernels of call k look at: the context's own, or the copy put aside for a
// hand-off that runs one call late
static const int32_t* shard_gate(caribou_cuda_ctx* c, const ShardCall& k, bool late)
{
    if (k.in_kind != IN_WORDS) return nullptr;
    return late ? c->shard->d_gate2 + (k.seq & 1) : &c->d_counters->any_misaligned;
}

// fix-up of call k from d_sin and the new carry into the context's state
static int shard_fixup(caribou_cuda_ctx* c, const ShardCall& k, bool late)
{
    Shard* s = c->shard;
    const FilterPlan& P = c->plan;
    const int32_t* gate = shard_gate(c, k, late);
    const FixPlan fp = shard_fix_plan(c, k);
    if (!fp.partial) {
        // no useful contraction, or a short shard: the whole shard again (tail included) from a scratch copy of s_in
        CBL_CHECK(cudaMemcpyAsync(s->d_work, s->d_sin, s->per_rank * 8, cudaMemcpyDeviceToDevice, c->stream));
        if (run_iir(c, k.in_kind, k.in, k.n, k.in_stride, k.out, k.out_stride, gate, 0, true, 0, s->d_work) != CARIBOU_CUDA_OK)
            return CARIBOU_CUDA_ERR;
    } else {
        // (the composition kernel has emptied this launch's descriptors and ticket: c->precleared_ptr, consumed by run_iir)
        if (run_iir(c, k.in_kind, k.in, k.n, k.in_stride, k.out, k.out_stride, gate, 0, false, fp.max_tiles, s->d_sin) != CARIBOU_CUDA_OK)
            return CARIBOU_CUDA_ERR;
    }
    // the stream's state after this call: every rank holds the same carry
    if (!s->committed) {
        launch_commit_state(c->d_state, s->d_carry, (size_t)c->cfg.n_streams, (size_t)2 * P.S, gate, 0, c->stream);
        c->launches++;
    }
    c->pos += k.n * (uint64_t)s->world;
    CBL_CHECK(cudaGetLastError());
    return CARIBOU_CUDA_OK;
}

// end of a call's device work as the profile sees it
static int shard_mark_end(caribou_cuda_ctx* c)
{
    Shard* s = c->shard;
    if (c->profiling) { CBL_CHECK(cudaEventRecord(s->ev[2], c->stream)); s->ev_valid = true; }
    s->open = false;
    return CARIBOU_CUDA_OK;
}

// peer: the end vectors come through the exchange buffers (flags); late: the hand-off of an earlier call (shard_flush)
static int shard_compose(caribou_cuda_ctx* c, const ShardCall& k, bool peer = false, bool late = false)
{
    Shard* s = c->shard;
    const size_t ns = (size_t)c->cfg.n_streams;
    ComposeParams p;
    memset(&p, 0, sizeof(p));
    p.S = c->plan.S; p.world = s->world; p.rank = s->rank;
    p.n_comp = (uint32_t)s->per_rank;
    p.f = s->d_f; p.carry = c->d_state; p.mpow = s->d_mpow; p.s_in = s->d_sin; p.carry_out = s->d_carry;
    p.flags = nullptr; p.seq = 0;
    p.err = c->d_sflags + 3 * ns + 1;
    p.err_host = s->err_host;
    if (peer) {
        const size_t par = (size_t)(k.seq & 1);
        p.f = s->d_px + par * s->world * s->per_rank;
        p.flags = reinterpret_cast<const unsigned long long*>(s->d_px + 2 * (size_t)s->world * s->per_rank) + par * s->world;
        p.seq = k.seq;
    }
    // a fix-up of a few leading tiles: its descriptors and ticket are emptied here (pass 1 used more tiles of the same
    // plan, so the buffers exist and are large enough)
    const FixPlan fp = shard_fix_plan(c, k);
    if (fp.partial) {
        const FilterPlan& P = pick_plan(c, k.n);
        DevBuf& desc = P.chained ? c->desc_incl : c->desc_agg;
        const uint64_t doubles = (uint64_t)fp.max_tiles * c->win_cnt * 2 * P.S;
        if (desc.p && doubles * 8 <= desc.cap && doubles <= (1u << 16)) {
            p.clear_desc = desc.as<unsigned long long>();
            p.clear_n = (uint32_t)doubles;
            p.clear_ticket = c->d_ticket;
            c->precleared_ptr = desc.p; c->precleared_doubles = doubles;
        }
    }
    const unsigned threads = 128, grid = (unsigned)std::min<size_t>((s->per_rank + threads - 1) / threads, 64);
    p.commit = grid == 1 ? c->d_state : nullptr;
    p.gate = shard_gate(c, k, late);
    s->committed = grid == 1;
    shard_compose_kernel<<<grid, threads, 0, c->stream>>>(p);
    c->launches++;
    CBL_CHECK(cudaGetLastError());
    return CARIBOU_CUDA_OK;
}

// the hand-off a peer-mode call left for later (composition + fix-up of that call), now
int shard_flush(caribou_cuda_ctx* c)
{
    Shard* s = c ? c->shard : nullptr;
    if (!s || !s->deferred) return CARIBOU_CUDA_OK;
    s->deferred = false;
    if (shard_compose(c, s->pend, true, true) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    return shard_fixup(c, s->pend, true);
}

}  // namespace cbl

using namespace cbl;

extern "C" {

int caribou_cuda_shard_unique_id(void* id)
{
    if (!id) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    NcclApi* a = nccl_api();
    if (!a) return CARIBOU_CUDA_ERR;
    static_assert(sizeof(ncclUniqueId) == CARIBOU_CUDA_SHARD_ID_BYTES, "ncclUniqueId is 128 bytes");
    ncclUniqueId u;
    CBL_NCCL(a, a->GetUniqueId(&u));
    memcpy(id, &u, sizeof(u));
    return CARIBOU_CUDA_OK;
}

int caribou_cuda_shard_init(caribou_cuda_ctx* c, int rank, int world, const void* id)
{
    if (shard_alloc(c, rank, world) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (!id) return CARIBOU_CUDA_OK;          // the caller moves the end vectors itself
    NcclApi* a = nccl_api();
    if (!a) return CARIBOU_CUDA_ERR;
    ncclUniqueId u;
    memcpy(&u, id, sizeof(u));
    CBL_NCCL(a, a->CommInitRank(&c->shard->comm, world, u, rank));
    c->shard->own_comm = true;
    return shard_peer_setup(c);
}

int caribou_cuda_shard_attach_comm(caribou_cuda_ctx* c, int rank, int world, void* nccl_comm)
{
    if (!nccl_comm) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    if (!nccl_api()) return CARIBOU_CUDA_ERR;
    if (shard_alloc(c, rank, world) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    c->shard->comm = static_cast<ncclComm_t>(nccl_comm);
    c->shard->own_comm = false;
    return shard_peer_setup(c);
}

int caribou_cuda_shard_set_mode(caribou_cuda_ctx* c, int mode)
{
    if (!c || !c->shard || mode < 0 || mode > 2) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    if (mode == 2 && !c->shard->peer_ok) { set_error("peer exchange is not available on this communicator"); return CARIBOU_CUDA_ERR; }
    if (shard_flush(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    c->shard->mode = mode;
    return CARIBOU_CUDA_OK;
}

int caribou_cuda_shard_mode(caribou_cuda_ctx* c)
{
    if (!c || !c->shard) { set_error("call caribou_cuda_shard_init first"); return CARIBOU_CUDA_ERR; }
    return c->shard->mode;
}

int caribou_cuda_shard_begin(caribou_cuda_ctx* c, int in_kind, const void* in, size_t n, size_t in_stride, void* out, size_t out_stride)
{
    return shard_begin(c, in_kind, in, n, in_stride, out, out_stride);
}

void* caribou_cuda_shard_gather_buffer(caribou_cuda_ctx* c, size_t* bytes_per_rank)
{
    if (!c || !c->shard) { set_error("call caribou_cuda_shard_init first"); return nullptr; }
    if (bytes_per_rank) *bytes_per_rank = c->shard->per_rank * 8;
    return c->shard->d_f;
}

int caribou_cuda_shard_finish(caribou_cuda_ctx* c)
{
    if (!c || !c->shard || !c->shard->open) { set_error("caribou_cuda_shard_finish without caribou_cuda_shard_begin"); return CARIBOU_CUDA_ERR; }
    CBL_CHECK(cudaSetDevice(c->device));
    if (shard_compose(c, c->shard->cur) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (shard_fixup(c, c->shard->cur, false) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    return shard_mark_end(c);
}

int caribou_cuda_shard_submit(caribou_cuda_ctx* c, int in_kind, const void* in, size_t n, size_t in_stride, void* out, size_t out_stride)
{
    if (!c || !c->shard) { set_error("call caribou_cuda_shard_init first"); return CARIBOU_CUDA_ERR; }
    Shard* s = c->shard;
    if (s->world > 1 && !s->comm) { set_error("no communicator: use caribou_cuda_shard_begin / _finish and move the end vectors yourself"); return CARIBOU_CUDA_ERR; }
    if (s->mode != 2 && shard_flush(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;       // (the mode was switched)
    if (shard_begin(c, in_kind, in, n, in_stride, out, out_stride) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (s->world == 1) {
        if (shard_compose(c, s->cur) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
        if (shard_fixup(c, s->cur, false) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
        return shard_mark_end(c);
    }
    NcclApi* a = nccl_api();
    if (!a) return CARIBOU_CUDA_ERR;
    double* f_own = s->d_f + (size_t)s->rank * s->per_rank;
    if (s->mode == 2) {
        // pass 1 of this call is queued; now the hand-off of the PREVIOUS call (every peer's vector of that call arrived a
        // whole pass ago), then this call's push.  Order matters for the two-deep exchange buffers: a rank pushes call
        // k + 2 into the slots of call k only after its composition of call k + 1, which waited for the peers' pushes of
        // call k + 1, which they make after their compositions of call k.
        if (shard_flush(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
        s->seq++;
        s->cur.seq = s->seq;
        const size_t par = (size_t)(s->seq & 1);
        shard_push_kernel<<<1, 128, 0, c->stream>>>(s->d_peers, f_own, (uint32_t)s->per_rank, s->world, s->rank,
                                                    par * s->world * s->per_rank,
                                                    2 * (size_t)s->world * s->per_rank + par * s->world, s->seq,
                                                    in_kind == IN_WORDS ? &c->d_counters->any_misaligned : nullptr, s->d_gate2 + par);
        c->launches++;
        CBL_CHECK(cudaGetLastError());
        s->pend = s->cur;
        s->deferred = true;
        return shard_mark_end(c);
    }
    if (s->mode == 0) {
        CBL_NCCL(a, a->AllGather(f_own, s->d_f, s->per_rank, ncclDouble, s->comm, c->stream));
        if (shard_compose(c, s->cur) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
        if (shard_fixup(c, s->cur, false) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
        return shard_mark_end(c);
    }
    // chain: s_in travels rank to rank, the last rank broadcasts the state after its shard
    const int S = c->plan.S;
    const unsigned threads = 128, grid = (unsigned)std::min<size_t>((s->per_rank + threads - 1) / threads, 64);
    if (s->rank == 0) CBL_CHECK(cudaMemcpyAsync(s->d_sin, c->d_state, s->per_rank * 8, cudaMemcpyDeviceToDevice, c->stream));
    else CBL_NCCL(a, a->Recv(s->d_sin, s->per_rank, ncclDouble, s->rank - 1, s->comm, c->stream));
    shard_hop_kernel<<<grid, threads, 0, c->stream>>>(S, (uint32_t)s->per_rank, s->d_mpow + (size_t)S * S, s->d_sin, f_own, s->d_carry);
    c->launches++;
    if (s->rank < s->world - 1) CBL_NCCL(a, a->Send(s->d_carry, s->per_rank, ncclDouble, s->rank + 1, s->comm, c->stream));
    CBL_NCCL(a, a->Broadcast(s->d_carry, s->d_carry, s->per_rank, ncclDouble, s->world - 1, s->comm, c->stream));
    s->committed = false;
    if (shard_fixup(c, s->cur, false) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    return shard_mark_end(c);
}

// `world` contexts of ONE process (one per GPU, or several on one GPU): every context's end vector is copied
// into every other context's gather buffer, ordered after pass 1 on the source and before finish on the target
int caribou_cuda_shard_exchange_local(caribou_cuda_ctx* const* ctxs, int world)
{
    if (!ctxs || world < 1) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    for (int g = 0; g < world; g++) {
        if (!ctxs[g] || !ctxs[g]->shard || ctxs[g]->shard->world != world || ctxs[g]->shard->rank != g || !ctxs[g]->shard->open ||
            ctxs[g]->shard->per_rank != ctxs[0]->shard->per_rank) {
            set_error("exchange_local: context %d is not rank %d of %d between begin and finish", g, g, world);
            return CARIBOU_CUDA_ERR;
        }
    }
    const size_t bytes = ctxs[0]->shard->per_rank * 8;
    std::vector<cudaEvent_t> done((size_t)world, nullptr);
    auto cleanup = [&]() { for (cudaEvent_t e : done) if (e) cudaEventDestroy(e); };
    for (int g = 0; g < world; g++) {
        if (!cuda_ok(cudaSetDevice(ctxs[g]->device), "cudaSetDevice") ||
            !cuda_ok(cudaEventCreateWithFlags(&done[g], cudaEventDisableTiming), "cudaEventCreate") ||
            !cuda_ok(cudaEventRecord(done[g], ctxs[g]->stream), "cudaEventRecord")) { cleanup(); return CARIBOU_CUDA_ERR; }
    }
    for (int t = 0; t < world; t++) {            // target context pulls every other rank's slot on its own stream
        caribou_cuda_ctx* ct = ctxs[t];
        if (!cuda_ok(cudaSetDevice(ct->device), "cudaSetDevice")) { cleanup(); return CARIBOU_CUDA_ERR; }
        for (int g = 0; g < world; g++) {
            if (g == t) continue;
            const double* src = ctxs[g]->shard->d_f + (size_t)g * ctxs[g]->shard->per_rank;
            double* dst = ct->shard->d_f + (size_t)g * ct->shard->per_rank;
            if (!cuda_ok(cudaStreamWaitEvent(ct->stream, done[g], 0), "cudaStreamWaitEvent") ||
                !cuda_ok(cudaMemcpyPeerAsync(dst, ct->device, src, ctxs[g]->device, bytes, ct->stream), "cudaMemcpyPeerAsync")) {
                cleanup();
                return CARIBOU_CUDA_ERR;
            }
        }
    }
    cleanup();
    return CARIBOU_CUDA_OK;
}

int caribou_cuda_shard_profile(caribou_cuda_ctx* c, float* pass_ms, float* handoff_ms)
{
    if (!c || !c->shard || !c->shard->ev_valid) { set_error("no profiled sharded call (caribou_cuda_set_profiling)"); return CARIBOU_CUDA_ERR; }
    CBL_CHECK(cudaSetDevice(c->device));
    if (shard_flush(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    float a = 0, b = 0;
    CBL_CHECK(cudaEventElapsedTime(&a, c->shard->ev[0], c->shard->ev[1]));
    CBL_CHECK(cudaEventElapsedTime(&b, c->shard->ev[1], c->shard->ev[2]));
    if (pass_ms) *pass_ms = a;
    if (handoff_ms) *handoff_ms = b;
    return CARIBOU_CUDA_OK;
}

}  // extern "C"
