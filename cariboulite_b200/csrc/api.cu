// libcaribou_cuda C ABI (include/caribou_cuda.h): context, submit paths, state, utilities.
#include <limits.h>
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <new>
#include <vector>
#include "ctx.h"

namespace cbl {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
bool cuda_ok(cudaError_t e, const char* what)
{
    if (e == cudaSuccess) return true;
    set_error("CUDA error %s (%d) in %s", cudaGetErrorString(e), (int)e, what);
    return false;
}
bool DevBuf::reserve(size_t n)
{
    if (n <= cap) return true;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    const size_t want = (n + 4095) & ~size_t(4095);
    if (!cuda_ok(cudaMalloc(&p, want), "cudaMalloc")) return false;
    cap = want;
    return true;
}
void DevBuf::release()
{
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
}
bool HostBuf::reserve(size_t n)
{
    if (n <= cap) return true;
    if (p) cudaFreeHost(p);
    p = nullptr; cap = 0;
    const size_t want = (n + 65535) & ~size_t(65535);
    if (!cuda_ok(cudaMallocHost(&p, want), "cudaMallocHost")) return false;
    cap = want;
    return true;
}
void HostBuf::release()
{
    if (p) cudaFreeHost(p);
    p = nullptr; cap = 0;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no -lcuda needed)
typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static encode_tiled_fn get_encode()
{
    static encode_tiled_fn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<encode_tiled_fn>(p);
    }
    return fn;
}

// Tensor map over the runs of a batch of streams: [n_streams][n_runs][LT * elem_words] uint32,
// box = 128 bytes x 32 runs x 1 stream, 128B swizzle (bank-conflict-free row reads).
static bool make_run_tmap(CUtensorMap* tm, const void* base, int LT, int elem_words, uint64_t n_runs,
                          uint32_t n_streams, uint64_t stream_stride_bytes)
{
    encode_tiled_fn enc = get_encode();
    if (!enc) { set_error("cuTensorMapEncodeTiled not available"); return false; }
    const cuuint64_t row_words = (cuuint64_t)LT * elem_words;
    cuuint64_t dims[3] = {row_words, n_runs, n_streams};
    cuuint64_t strides[2] = {row_words * 4, n_streams > 1 ? stream_stride_bytes : std::max<cuuint64_t>(row_words * 4 * n_runs, 16)};
    cuuint32_t box[3] = {32, 32, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    if ((uintptr_t)base % 16 != 0 || strides[1] % 16 != 0) { set_error("input must be 16-byte aligned (base and stream stride)"); return false; }
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (%d)", (int)r); return false; }
    return true;
}

static size_t in_elem_bytes(int in_kind) { return in_kind == IN_CF32 ? 8 : 4; }

static bool resolve_out_mode(int parity, int fmt, int* mode, size_t* elem)
{
    if (parity == CARIBOU_CUDA_PARITY_F32) {
        if (fmt == CARIBOU_CUDA_FMT_CF32) { *mode = OUT_F32; *elem = 8; return true; }
        if (fmt == CARIBOU_CUDA_FMT_CF64) { *mode = OUT_F64; *elem = 16; return true; }
        return false;
    }
    switch (fmt) {
        case CARIBOU_CUDA_FMT_CS16: *mode = OUT_Q16_CS16; *elem = 4; return true;
        case CARIBOU_CUDA_FMT_CF32: *mode = OUT_Q16_CF32; *elem = 8; return true;
        case CARIBOU_CUDA_FMT_CF64: *mode = OUT_Q16_CF64; *elem = 16; return true;
        case CARIBOU_CUDA_FMT_CS8:  *mode = OUT_Q16_CS8;  *elem = 2; return true;
    }
    return false;
}

// kernel-coordinate <-> reference-coordinate state conversion, both rails (tables.cpp: Tkr / Trk)
static void state_to_kernel(const FilterPlan& P, const double* ref, double* ker)
{
    for (int r = 0; r < 2; r++) plan_state_to_kernel(P, ref + r * P.S, ker + r * P.S);
}
static void state_to_ref(const FilterPlan& P, const double* ker, double* ref)
{
    for (int r = 0; r < 2; r++) plan_state_to_ref(P, ker + r * P.S, ref + r * P.S);
}

static double scale_of(const caribou_cuda_ctx* c, int in_kind)
{
    return (c->cfg.parity == CARIBOU_CUDA_PARITY_F32 && in_kind != IN_CF32) ? 1.0 / 4096.0 : 1.0;
}
// The kernels filter the raw sample values and fold the input scale into the output gain (exact: a power
// of two).  If the input kind of a stream changes mid-life, its state is rescaled to the new raw units.
int ensure_scale(caribou_cuda_ctx* c, int in_kind)
{
    const double sc = scale_of(c, in_kind);
    if (sc == c->in_scale) return CARIBOU_CUDA_OK;
    if (resolve_queue(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    const size_t n = (size_t)c->cfg.n_streams * 2 * c->plan.S;
    std::vector<double> st(n);
    // copies ordered on the context's (non-blocking) stream, then waited for: the host vector is pageable
    CBL_CHECK(cudaMemcpyAsync(st.data(), c->d_state, n * 8, cudaMemcpyDeviceToHost, c->stream));
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    for (double& v : st) v *= c->in_scale / sc;
    CBL_CHECK(cudaMemcpyAsync(c->d_state, st.data(), n * 8, cudaMemcpyHostToDevice, c->stream));
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    c->in_scale = sc;
    return CARIBOU_CUDA_OK;
}

constexpr size_t kProfRing = 256;
// which = 0: start of the dominant kernel, 1: its end (closes the pair)
static void prof_mark(caribou_cuda_ctx* c, int which)
{
    if (!c->profiling || c->prof_ev.empty()) return;
    cudaEventRecord(c->prof_ev[2 * (c->prof_n % kProfRing) + which], c->stream);
    if (which == 1) c->prof_n++;
}

// ------------------------------------------------------------------ the IIR launch sequence
// Filters n samples per stream that sit in device memory (declaration and contract: ctx.h).
int run_iir(caribou_cuda_ctx* c, int in_kind, const void* d_in, size_t n, size_t in_stride,
            void* d_out, size_t out_stride, const int32_t* gate, uint32_t gate_stride, bool commit, uint32_t max_tiles,
            double* state)
{
    const FilterPlan& P = pick_plan(c, n);
    const bool small = &P == &c->plan_small;
    if (!state) state = c->d_state;
    const uint32_t ns = c->win_cnt;                               // streams of this launch (a window of the batch)
    const size_t st_off = (size_t)c->win0 * 2 * P.S;             // their filter states
    const int LT = P.LT;
    const uint64_t n_runs = n / LT;
    const uint64_t n_main = n_runs * LT;
    IirTables T = P.tables;
    T.gain = P.g[P.K] * scale_of(c, in_kind);   // x / 4096 first == y / 4096 last (exact)
    // integer inputs of the parallel form enter as x + bias (mantissa injection, iir.cuh)
    const double bias = !P.par ? 0.0 : (in_kind == IN_WORDS ? kBiasWords : (in_kind == IN_CS16 ? kBiasCs16 : 0.0));
    for (int i = 0; i < P.S; i++) T.sbias[i] = bias * P.sstar[i];
    T.out_bias = -bias * P.hn1 * T.gain;
    const int layout = c->cfg.channel == CARIBOU_CUDA_CH_HIF ? 1 : 0;
    const size_t es = in_elem_bytes(in_kind);
    bool need_commit = false;

    if (n_runs > 0) {
        uint64_t tiles_ps = (n_runs + 31) / 32;
        bool whole = true;
        if (max_tiles > 0 && tiles_ps > max_tiles) { tiles_ps = max_tiles; whole = false; }
        const uint64_t total = tiles_ps * ns;
        if (total > 0xFFFFFFF0ull) { set_error("too many tiles"); return CARIBOU_CUDA_ERR; }
        DevBuf& desc = P.chained ? c->desc_incl : c->desc_agg;     // the one the look-back reads
        if (!desc.reserve(total * 2 * P.S * 8) || !c->desc_incl.reserve(2 * P.S * 8) || !c->desc_agg.reserve(2 * P.S * 8))
            return CARIBOU_CUDA_ERR;
        CUtensorMap tm;
        if (!make_run_tmap(&tm, d_in, LT, (int)(es / 4), n_runs, ns, (uint64_t)in_stride * es)) return CARIBOU_CUDA_ERR;
        // an earlier kernel of this submit may have emptied exactly these descriptors and the ticket already (the
        // control-block preparation of the rx path, the sharded hand-off's composition kernel)
        const bool precleared = c->precleared_ptr == desc.p && c->precleared_doubles == total * 2 * P.S;
        c->precleared_ptr = nullptr;
        if (!precleared) {
            CBL_CHECK(cudaMemsetAsync(desc.p, 0xFF, total * 2 * P.S * 8, c->stream));      // kDescEmpty everywhere
            CBL_CHECK(cudaMemsetAsync(c->d_ticket, 0, 4, c->stream));
        }
        IirParams p;
        memset(&p, 0, sizeof(p));
        p.n_streams = ns;
        p.n_runs = (uint32_t)n_runs;
        p.tiles_per_stream = (uint32_t)tiles_ps;
        p.total_tiles = (uint32_t)total;
        p.last_tile_runs = (uint32_t)(n_runs - (tiles_ps - 1) * 32);
        if (!whole) p.last_tile_runs = 32;
        p.window = P.window;
        p.chained = P.chained ? 1 : 0;
        p.write_state = (commit && whole) ? 1 : 0;
        p.layout = layout;
        p.out_mode = c->out_mode;
        p.mtile_pow = small ? c->d_mtile_pow_small : c->d_mtile_pow;
        p.plane = small ? c->d_plane_small : c->d_plane;
        p.rr_boxes = (uint32_t)(LT / (in_kind == IN_CF32 ? 16 : 32));
        p.desc_agg = c->desc_agg.as<double>();
        p.desc_incl = c->desc_incl.as<double>();
        p.ticket = c->d_ticket;
        p.state_in = state + st_off;
        p.state_out = c->d_state_next + st_off;
        p.out = d_out;
        p.out_stride = out_stride;
        p.gate = gate;
        p.gate_stride = gate_stride;
        // (0x413 << 21) | sign bit of the 13-bit field at bits 13:1;  (0x414 << 21) | sign bit of the int16
        p.inj_mul = 1u << 31;
        p.inj_xor = in_kind == IN_WORDS ? 0x82602000u : 0x82808000u;
        p.scan_rounds = P.scan_rounds;
        p.out_vec = ((reinterpret_cast<uintptr_t>(d_out) & 31) == 0 && (out_stride & 3) == 0) ? 1 : 0;
        // CF32 output without int16 truncation: the float-staged kernel (half the shared memory per warp, more
        // warps per SM).  It stages gain * y_zs, so its R table carries the gain as well.
        const bool y32 = P.y32;        // decided with the plan (it fixes the run length)
        IirTables Tl = T;
        if (y32 && in_kind != IN_CF32) {
            // fixed-point staging of the integer-fed float-staged kernels: |gain * y_zs| <= |gain| * l1_zs * max|x|
            // must fit 31 bits below the binary point F (float input, whose range is unknown, is staged as floats)
            const double xmax = in_kind == IN_CS16 ? 32768.0 : 4096.0;
            const double bound = fabs(T.gain) * P.l1_zs * xmax * 1.0001;
            int e = 900;                                                  // a non-finite bound overflows the output anyway
            if (std::isfinite(bound)) { e = 0; if (bound > 0) frexp(bound, &e); e = std::min(std::max(e, -900), 900); }   // bound < 2^e
            const double magic = ldexp(1.5, 21 + e);                 // F = 31 - e: 1.5 * 2^(52-F)
            const double mo = magic + ldexp(1.0, e);                 // + 2^(31-F): staged words are non-negative
            unsigned long long mb;
            memcpy(&mb, &magic, 8);
            Tl.stage_mo = mo;
            Tl.stage_hi = (uint32_t)(mb >> 32);
            Tl.out_bias = T.out_bias + mo;
        }
        if (y32)
            for (int m = 0; m < kMaxNK; m++)
                for (int i = 0; i < kMaxS; i++) Tl.R[m][i] *= T.gain;
        // AP form of the rerun kernel's phase 3 (iir_rerun.cuh): the (1 + q)^2K numerator on the integer samples, then
        // the all-pole cascade (constants: tables.cpp plan_ap_launch)
        static const bool no_ap = getenv("CARIBOU_CUDA_NO_AP") != nullptr;
        if (P.rerun && P.ap && in_kind != IN_CF32 && !no_ap) {
            ApLaunch L;
            plan_ap_launch(P, in_kind == IN_WORDS, T.gain, &L);
            p.ap = 1;
            p.ap_exp = L.exp; p.ap_mul = L.mul; p.ap_xor = L.xorw;
            for (int i = 0; i < P.S; i++) {
                for (int k = 0; k < P.S; k++) Tl.R[i][k] = L.T[i][k];
                Tl.R[kApBiasRow][i] = L.bias[i];
            }
            Tl.gain = L.gain;
            Tl.out_bias = L.out_bias;
        }
        // rerun kernel: phase 1 takes the boxes that will stay in the ring last (iir_rerun.cuh); the end vectors of its
        // two parts combine through A^(samples of the first part)
        if (P.rerun && CBL_RR_RESIDENT) {
            const int nb = (int)p.rr_boxes, keep = rerun_keep(nb, rerun_ring(P.K, p.ap != 0), p.ap != 0);
            if (keep > 0 && keep < nb) {
                const uint64_t len_a = (uint64_t)(nb - keep) * (in_kind == IN_CF32 ? 16 : 32);
                auto& rp = c->rr_pow[small ? 1 : 0];
                if (rp.len != len_a) { plan_power_kernel(P, len_a, rp.m); rp.len = len_a; }
                for (int i = 0; i < P.S; i++)
                    for (int k = 0; k < P.S; k++) Tl.R[kRrPowRow + i][k] = rp.m[i * P.S + k];
            }
        }
        prof_mark(c, 0);
        const int kin = (in_kind == IN_WORDS && layout) ? IN_WORDS_HIF : in_kind;
        const int e = launch_iir(P.K, P.D, kin, P.par, y32, tm, p, Tl, c->num_sms, c->stream);
        prof_mark(c, 1);
        if (e != 0) { cuda_ok((cudaError_t)e, "iir_scan_kernel launch"); return CARIBOU_CUDA_ERR; }
        c->launches++;
        need_commit = p.write_state != 0;
    }
    // the remainder that fills no run, and the commit of the state the scan kernel left in its scratch array, are one
    // launch: the tail kernel starts from that array and writes the context's state (first == n: commit only)
    if ((n_main < n && max_tiles == 0) || need_commit) {
        IirTailParams t;
        memset(&t, 0, sizeof(t));
        t.n_streams = ns;
        t.in = d_in; t.in_stride = in_stride;
        t.first = n_main; t.n = max_tiles == 0 ? n : n_main;
        t.out = d_out; t.out_stride = out_stride;
        t.in_kind = in_kind; t.layout = layout; t.out_mode = c->out_mode; t.K = P.K; t.D = P.D; t.par = P.par ? 1 : 0;
        for (int k = 0; k < P.K; k++) { t.a1[k] = T.a1[k]; t.a2[k] = T.a2[k]; t.c1[k] = T.c1[k]; t.c2[k] = T.c2[k]; }
        t.gain = T.gain;
        t.state = state + st_off;
        t.state_src = need_commit ? c->d_state_next + st_off : nullptr;
        if (c->mirror_in_tail && c->h_ctrl_dev != nullptr && c->win0 == 0 && c->win_cnt == (uint32_t)c->cfg.n_streams) {
            t.mirror_src = reinterpret_cast<const uint32_t*>(c->d_ctrl);
            t.mirror_dst = reinterpret_cast<uint32_t*>(c->h_ctrl_dev);
            t.mirror_words = (uint32_t)(c->ctrl_bytes / 4);
            c->mirror_in_tail = false;          // done: the caller skips its copy
        }
        t.gate = gate;
        t.gate_stride = gate_stride;
        launch_iir_tail(t, c->stream);
        c->launches++;
    }
    CBL_CHECK(cudaGetLastError());
    return CARIBOU_CUDA_OK;
}

const FilterPlan& pick_plan(const caribou_cuda_ctx* c, size_t n)
{
    if (!c->has_small) return c->plan;
    const uint64_t tiles = ((uint64_t)(n / (size_t)c->plan.LT) + 31) / 32 * c->win_cnt;
    return tiles < (uint64_t)c->num_sms * 12 ? c->plan_small : c->plan;
}

// ------------------------------------------------------------------ alignment search for a batch
struct SubGeom { SubView sv; uint32_t n_chunks; };

static int n_subs_of(const caribou_cuda_ctx* c) { return c->cfg.interleave_block_words ? 2 : 1; }

static SubGeom sub_geom(const caribou_cuda_ctx* c, const void* d_bytes, size_t nbytes, int sub)
{
    SubGeom g;
    const uint32_t B = c->cfg.interleave_block_words;
    g.sv.words = reinterpret_cast<const uint32_t*>(d_bytes);
    g.sv.B = B;
    g.sv.sub = (uint32_t)sub;
    if (B == 0) {
        g.sv.nbytes = nbytes;
    } else {
        const uint64_t nw = nbytes / 4, per = 2ull * B;
        const uint64_t full = nw / per, rem = nw % per;
        uint64_t cnt = full * B;
        if (sub == 0) cnt += std::min<uint64_t>(rem, B);
        else cnt += rem > B ? rem - B : 0;
        g.sv.nbytes = cnt * 4;
    }
    g.sv.n_sub_words = (g.sv.nbytes + 3) / 4;
    g.n_chunks = (uint32_t)((g.sv.nbytes + c->cfg.chunk_bytes - 1) / c->cfg.chunk_bytes);
    return g;
}

// Everything the alignment search and the scan kernel that follows need reset, in one launch instead of five memsets:
// first-bad-chunk slots to "none", the counters and per-capture misalignment counts to zero, the scan kernel's
// look-back descriptors to "empty" and its ticket to zero.
struct PrepParams {
    int32_t* first_bad; uint32_t n_first_bad;
    uint32_t* counters; uint32_t n_counters;        // 32-bit words, nullptr: leave alone
    int32_t* nmis; uint32_t n_nmis;
    unsigned long long* desc; uint32_t n_desc;
    uint32_t* ticket;
};
__global__ void ctrl_prep_kernel(const PrepParams p)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x, nt = gridDim.x * blockDim.x;
    for (uint32_t i = t; i < p.n_first_bad; i += nt) p.first_bad[i] = 0x7f7f7f7f;
    if (p.counters) for (uint32_t i = t; i < p.n_counters; i += nt) p.counters[i] = 0u;
    if (p.nmis) for (uint32_t i = t; i < p.n_nmis; i += nt) p.nmis[i] = 0;
    if (p.desc) {
        for (uint32_t i = t; i < p.n_desc; i += nt) p.desc[i] = kDescEmpty;
        if (t == 0) *p.ticket = 0u;
    }
}

// iir_n > 0: a run_iir over iir_n samples per stream of the same window follows at once (whole streams, raw words):
// its descriptors and ticket are emptied here as well
int run_find_offsets(caribou_cuda_ctx* c, const void* d_bytes, size_t nbytes, size_t stride_bytes, bool with_gates, size_t iir_n)
{
    // d_bytes points at the first stream of the current window (c->win0, c->win_cnt)
    const uint32_t ns_all = (uint32_t)c->cfg.n_streams, ns = c->win_cnt, w0 = c->win0;
    const int nsub = n_subs_of(c);
    uint32_t max_chunks = 0;
    for (int s = 0; s < nsub; s++) max_chunks = std::max(max_chunks, sub_geom(c, d_bytes, nbytes, s).n_chunks);
    if (!c->offs.reserve((size_t)ns_all * 2 * std::max<uint32_t>(max_chunks, 1) * 4)) return CARIBOU_CUDA_ERR;
    // first_bad <- INT_MAX (0x7f7f7f7f is large enough) for the window; counters <- 0 at the first window
    PrepParams pp;
    memset(&pp, 0, sizeof(pp));
    pp.first_bad = c->d_first_bad + (size_t)w0 * 2; pp.n_first_bad = ns * 2;
    if (w0 == 0) { pp.counters = reinterpret_cast<uint32_t*>(c->d_counters); pp.n_counters = (uint32_t)(sizeof(Counters) / 4); }
    if (with_gates) { pp.nmis = c->d_sflags + 2 * (size_t)ns_all + w0; pp.n_nmis = ns; }
    c->precleared_ptr = nullptr;
    if (iir_n > 0 && c->has_filter) {
        const FilterPlan& P = pick_plan(c, iir_n);
        const uint64_t n_runs = iir_n / (size_t)P.LT, total = (n_runs + 31) / 32 * ns;
        DevBuf& desc = P.chained ? c->desc_incl : c->desc_agg;
        if (n_runs > 0 && total * 2 * P.S < (1ull << 31)) {
            if (!desc.reserve(total * 2 * P.S * 8)) return CARIBOU_CUDA_ERR;
            pp.desc = desc.as<unsigned long long>(); pp.n_desc = (uint32_t)(total * 2 * P.S);
            pp.ticket = c->d_ticket;
            c->precleared_ptr = desc.p; c->precleared_doubles = total * 2 * P.S;
        }
    }
    {
        const uint32_t work = std::max<uint32_t>(std::max(pp.n_first_bad, pp.n_desc), 256);
        const unsigned grid = (unsigned)std::min<uint32_t>((work + 1023) / 1024, (uint32_t)c->num_sms * 2);
        ctrl_prep_kernel<<<grid, 256, 0, c->stream>>>(pp);
        c->launches++;
    }
    for (int s = 0; s < nsub; s++) {
        const SubGeom g = sub_geom(c, d_bytes, nbytes, s);
        OffsetsParams p;
        p.sv = g.sv;
        p.stream_stride_bytes = stride_bytes;
        p.n_streams = ns;
        p.chunk_bytes = (uint32_t)c->cfg.chunk_bytes;
        p.n_chunks = g.n_chunks;
        p.offs = c->offs.as<int32_t>() + ((size_t)s * ns_all + w0) * max_chunks;
        p.offs_stride = max_chunks;
        p.first_bad = c->d_first_bad + (size_t)w0 * 2 + s;
        p.first_bad_stride = 2;
        p.n_misaligned = &c->d_counters->n_misaligned;
        p.any_misaligned = &c->d_counters->any_misaligned;
        p.gate = with_gates ? c->d_sflags + w0 : nullptr;
        p.gate_seq = with_gates ? c->d_sflags + ns_all + w0 : nullptr;
        p.nmis = with_gates ? c->d_sflags + 2 * (size_t)ns_all + w0 : nullptr;
        p.seq = (int32_t)c->queue.size();
        p.lost = c->d_sflags + 3 * (size_t)ns_all;
        launch_find_offsets(p, c->stream);
        if (g.n_chunks) c->launches++;
    }
    CBL_CHECK(cudaGetLastError());
    return (int)max_chunks;
}

// unpack kernel for every stream of the batch; all pointers are device pointers
// (streams [c->win0, c->win0 + c->win_cnt) of the batch; d_bytes and the outputs point at stream 0)
static int run_unpack(caribou_cuda_ctx* c, const void* d_bytes, size_t nbytes, size_t stride_bytes,
                      const caribou_cuda_unpack_out& o, uint32_t max_chunks, size_t out_stride_samples)
{
    const uint32_t ns = (uint32_t)c->cfg.n_streams;
    const uint32_t st0 = c->win0, st1 = c->win0 + c->win_cnt;
    const int nsub = n_subs_of(c);
    const uint64_t n_raw_words = (nbytes + 3) / 4;
    const uint32_t n_tiles = (uint32_t)((n_raw_words + kUnpackTileWords - 1) / kUnpackTileWords);
    if (n_tiles == 0) return CARIBOU_CUDA_OK;
    const bool want_flags = true;
    if (n_raw_words > (1ull << 31)) { set_error("unpack: at most 2^31 words per stream and submit"); return CARIBOU_CUDA_ERR; }
    // per-tile counts, their scan, work list (+ its length); running totals per stream launch
    if (n_raw_words * ns > (1ull << 32) && (o.invalid_idx || o.sync_idx)) {
        set_error("unpack: index lists are 32-bit - at most 2^32 words per batch when they are requested");
        return CARIBOU_CUDA_ERR;
    }
    if (!c->desc_inv.reserve((size_t)n_tiles * 20 + 32) || !c->desc_syn.reserve(((size_t)ns + 1) * 16)) return CARIBOU_CUDA_ERR;
    // running (n_invalid, n_sync): launch k reads what launch k - 1 wrote.  The first launch reads zeros from the 16
    // bytes in front of the tile counts (one memset covers both), the last one writes the context's counters.
    uint64_t* zero_totals = c->desc_inv.as<uint64_t>();
    uint64_t* run_totals = c->desc_syn.as<uint64_t>();
    const bool prof = c->profiling && !c->has_filter;
    if (prof) prof_mark(c, 0);
    for (uint32_t st = st0; st < st1; st++) {
        UnpackParams p;
        memset(&p, 0, sizeof(p));
        const uint8_t* sb = reinterpret_cast<const uint8_t*>(d_bytes) + (size_t)st * stride_bytes;
        p.words = reinterpret_cast<const uint32_t*>(sb);
        p.nbytes = nbytes;
        p.B = c->cfg.interleave_block_words;
        p.chunk_bytes = (uint32_t)c->cfg.chunk_bytes;
        p.spc = p.chunk_bytes / 4;
        p.n_subs = nsub;
        void* cs16[2] = {o.cs16, o.cs16_b};
        void* meta[2] = {o.meta, o.meta_b};
        void* cf32[2] = {o.cf32, o.cf32_b};
        int vec_ok = 1;
        for (int s = 0; s < nsub; s++) {
            const SubGeom g = sub_geom(c, sb, nbytes, s);
            p.sv[s] = g.sv;
            p.layout[s] = nsub == 2 ? s : (c->cfg.channel == CARIBOU_CUDA_CH_HIF ? 1 : 0);
            p.offs[s] = c->offs.as<int32_t>() + ((size_t)s * ns + st) * max_chunks;
            p.first_bad[s] = c->d_first_bad + (size_t)st * 2 + s;
            p.cs16[s] = cs16[s] ? reinterpret_cast<uint32_t*>(cs16[s]) + (size_t)st * out_stride_samples : nullptr;
            p.meta[s] = meta[s] ? reinterpret_cast<uint8_t*>(meta[s]) + (size_t)st * out_stride_samples : nullptr;
            p.cf32[s] = cf32[s] ? reinterpret_cast<float2*>(cf32[s]) + (size_t)st * out_stride_samples : nullptr;
            if ((uintptr_t)p.cs16[s] % 16 || (uintptr_t)p.meta[s] % 4 || (uintptr_t)p.cf32[s] % 16) vec_ok = 0;
        }
        if ((uintptr_t)p.words % 16) { set_error("SMI byte stream must be 16-byte aligned in device memory"); return CARIBOU_CUDA_ERR; }
        p.vec_ok = vec_ok;
        p.want_flags = want_flags ? 1 : 0;
        p.inv_idx = o.invalid_idx; p.inv_cap = o.invalid_cap;
        p.syn_idx = o.sync_idx; p.syn_cap = o.sync_cap;
        p.index_base = (uint64_t)st * (nbytes / 4);
        p.tile_counts = c->desc_inv.as<uint64_t>() + 2;
        p.tile_offsets = p.tile_counts + n_tiles;
        p.work = reinterpret_cast<uint32_t*>(p.tile_offsets + n_tiles);
        p.n_work = p.work + n_tiles;
        p.totals_in = st == st0 ? zero_totals : run_totals + 2 * (size_t)(st - st0 - 1);
        p.totals_out = st + 1 == st1 ? c->d_counters->totals : run_totals + 2 * (size_t)(st - st0);
        p.n_tiles = n_tiles;
        if (st == st0) CBL_CHECK(cudaMemsetAsync(zero_totals, 0, 16 + (size_t)n_tiles * 8, c->stream));
        else CBL_CHECK(cudaMemsetAsync(p.tile_counts, 0, (size_t)n_tiles * 8, c->stream));
        c->launches += (uint64_t)launch_unpack(p, c->num_sms, c->stream);
    }
    if (prof) prof_mark(c, 1);
    CBL_CHECK(cudaGetLastError());
    return CARIBOU_CUDA_OK;
}

int mirror_counters(caribou_cuda_ctx* c)
{
    CBL_CHECK(cudaMemcpyAsync(c->h_ctrl, c->d_ctrl, c->ctrl_bytes, cudaMemcpyDeviceToHost, c->stream));
    return CARIBOU_CUDA_OK;
}

// samples the reference's read loop delivers for one sub-stream of `nbytes` given the first bad chunk
static int64_t delivered_samples(const caribou_cuda_ctx* c, uint64_t sub_nbytes, int32_t first_bad)
{
    const uint64_t cb = c->cfg.chunk_bytes;
    const uint64_t n_chunks = (sub_nbytes + cb - 1) / cb;
    if (first_bad >= 0 && (uint64_t)first_bad < n_chunks) return CARIBOU_CUDA_ERR_SYNC;
    // every full chunk gives cb/4 samples, the last partial one floor(len/4) (caribou_smi.c:677)
    return (int64_t)((sub_nbytes / cb) * (cb / 4) + (sub_nbytes % cb) / 4);
}

}  // namespace cbl

using namespace cbl;

// ==================================================================== C ABI
extern "C" {

const char* caribou_cuda_last_error(void) { return g_err; }

caribou_cuda_ctx* caribou_cuda_create(const caribou_cuda_config* cfg_in)
{
    if (!cfg_in) { set_error("null config"); return nullptr; }
    caribou_cuda_ctx* c = new (std::nothrow) caribou_cuda_ctx();
    if (!c) { set_error("out of memory"); return nullptr; }
    c->cfg = *cfg_in;
    caribou_cuda_config& cfg = c->cfg;
    if (cfg.chunk_bytes == 0) cfg.chunk_bytes = CARIBOU_CUDA_NATIVE_BATCH_LEN;
    if (cfg.n_streams <= 0) cfg.n_streams = 1;
    if (cfg.decimation <= 0) cfg.decimation = 1;
    auto fail = [&](void) -> caribou_cuda_ctx* { caribou_cuda_destroy(c); return nullptr; };
    if (cfg.chunk_bytes % 16 != 0 || cfg.chunk_bytes < 32 || cfg.chunk_bytes > (1u << 30)) { set_error("chunk_bytes must be a multiple of 16 in [32, 2^30]"); return fail(); }
    const uint32_t B = cfg.interleave_block_words;
    if (!(B == 0 || B == 1 || B % 8 == 0)) { set_error("interleave_block_words must be 0, 1 or a multiple of 8"); return fail(); }
    if (B && cfg.n_sos) { set_error("filtering a dual-channel capture is done per channel: demux first"); return fail(); }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { set_error("no CUDA device: libcaribou_cuda has no CPU fallback"); return fail(); }
    if (cfg.device < 0 || cfg.device >= ndev) { set_error("bad device ordinal %d", cfg.device); return fail(); }
    c->device = cfg.device;
    if (!cuda_ok(cudaSetDevice(c->device), "cudaSetDevice")) return fail();
    cudaDeviceProp prop;
    if (!cuda_ok(cudaGetDeviceProperties(&prop, c->device), "cudaGetDeviceProperties")) return fail();
    if (prop.major != 10) { set_error("libcaribou_cuda is built for sm_100a only (device is sm_%d%d)", prop.major, prop.minor); return fail(); }
    c->num_sms = prop.multiProcessorCount;
    if (cfg.cuda_stream) { c->stream = (cudaStream_t)cfg.cuda_stream; c->own_stream = false; }
    else {
        if (!cuda_ok(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking), "cudaStreamCreate")) return fail();
        c->own_stream = true;
    }
    const size_t ns = (size_t)cfg.n_streams;
    c->win0 = 0; c->win_cnt = (uint32_t)ns;
    // one control block on the device and its page-locked mirror (one copy per submit brings all of it down):
    // [Counters][first_bad: ns x 2][sflags: 3 ns + 4]
    static_assert(sizeof(Counters) % 8 == 0, "control block layout");
    c->ctrl_bytes = sizeof(Counters) + ns * 2 * 4 + (ns * 3 + 4) * 4;
    if (!cuda_ok(cudaMalloc(&c->d_ticket, 256), "cudaMalloc") ||
        !cuda_ok(cudaMalloc(&c->d_ctrl, c->ctrl_bytes), "cudaMalloc") ||
        !cuda_ok(cudaMallocHost(&c->h_ctrl, c->ctrl_bytes), "cudaMallocHost")) return fail();
    c->d_counters = reinterpret_cast<Counters*>(c->d_ctrl);
    {
        void* dp = nullptr;
        if (cudaHostGetDevicePointer(&dp, c->h_ctrl, 0) == cudaSuccess) c->h_ctrl_dev = static_cast<uint8_t*>(dp);
        else cudaGetLastError();                // no device view: the mirror stays a copy operation
    }
    c->h_counters = reinterpret_cast<Counters*>(c->h_ctrl);
    c->d_first_bad = reinterpret_cast<int32_t*>(c->d_ctrl + sizeof(Counters));
    c->h_first_bad = reinterpret_cast<int32_t*>(c->h_ctrl + sizeof(Counters));
    c->d_sflags = c->d_first_bad + ns * 2;
    c->h_sflags = c->h_first_bad + ns * 2;
    cudaMemset(c->d_sflags, 0, (ns * 3 + 4) * 4);
    cudaMemset(c->d_sflags + ns, 0x7f, ns * 4);
    memset(c->h_sflags, 0, (ns * 3 + 4) * 4);
    c->sres.resize(ns);
    cudaMemset(c->d_ticket, 0, 256);
    cudaMemset(c->d_counters, 0, sizeof(Counters));
    cudaMemset(c->d_first_bad, 0x7f, ns * 2 * 4);
    memset(c->h_counters, 0, sizeof(Counters));
    for (size_t i = 0; i < ns * 2; i++) c->h_first_bad[i] = INT_MAX;

    if (cfg.n_sos > 0) {
        const char* msg = "";
        if (!resolve_out_mode(cfg.parity, cfg.out_format, &c->out_mode, &c->out_elem)) { set_error("unsupported parity/out_format combination"); return fail(); }
        if (!make_filter_plan(cfg.n_sos, cfg.sos, cfg.decimation, cfg.kernel_form == CARIBOU_CUDA_FORM_CASCADE, c->out_mode == OUT_F32, &c->plan, &msg)) { set_error("%s", msg); return fail(); }
        const FilterPlan& P = c->plan;
        if (!cuda_ok(cudaMalloc(&c->d_plane, P.plane.size() * 8), "cudaMalloc") ||
            !cuda_ok(cudaMalloc(&c->d_mtile_pow, P.mtile_pow.size() * 8), "cudaMalloc") ||
            !cuda_ok(cudaMalloc(&c->d_state, ns * 2 * P.S * 8), "cudaMalloc") ||
            !cuda_ok(cudaMalloc(&c->d_state_next, ns * 2 * P.S * 8), "cudaMalloc")) return fail();
        cudaMemcpy(c->d_plane, P.plane.data(), P.plane.size() * 8, cudaMemcpyHostToDevice);
        cudaMemcpy(c->d_mtile_pow, P.mtile_pow.data(), P.mtile_pow.size() * 8, cudaMemcpyHostToDevice);
        if (P.rerun && P.LT > kRerunSmallLT) {
            if (!make_filter_plan(cfg.n_sos, cfg.sos, cfg.decimation, false, c->out_mode == OUT_F32, &c->plan_small, &msg, kRerunSmallLT)) { set_error("%s", msg); return fail(); }
            const FilterPlan& Q = c->plan_small;
            if (!cuda_ok(cudaMalloc(&c->d_plane_small, Q.plane.size() * 8), "cudaMalloc") ||
                !cuda_ok(cudaMalloc(&c->d_mtile_pow_small, Q.mtile_pow.size() * 8), "cudaMalloc")) return fail();
            cudaMemcpy(c->d_plane_small, Q.plane.data(), Q.plane.size() * 8, cudaMemcpyHostToDevice);
            cudaMemcpy(c->d_mtile_pow_small, Q.mtile_pow.data(), Q.mtile_pow.size() * 8, cudaMemcpyHostToDevice);
            c->has_small = Q.rerun && Q.par == P.par;
        }
        cudaMemset(c->d_state, 0, ns * 2 * P.S * 8);
        cudaMemset(c->d_state_next, 0, ns * 2 * P.S * 8);
        c->has_filter = true;
    }
    if (!cuda_ok(cudaDeviceSynchronize(), "context set-up")) return fail();
    return c;
}

void caribou_cuda_destroy(caribou_cuda_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->shard) { shard_release(c->shard); c->shard = nullptr; }
    DevBuf* bufs[] = {&c->desc_agg, &c->desc_incl, &c->offs, &c->desc_inv, &c->desc_syn, &c->stage_in,
                      &c->stage_cs16[0], &c->stage_cs16[1], &c->stage_meta[0], &c->stage_meta[1], &c->stage_cf32[0],
                      &c->stage_cf32[1], &c->stage_filt, &c->stage_inv, &c->stage_syn, &c->scratch_cs16};
    for (DevBuf* b : bufs) b->release();
    if (c->d_plane) cudaFree(c->d_plane);
    if (c->d_mtile_pow) cudaFree(c->d_mtile_pow);
    if (c->d_plane_small) cudaFree(c->d_plane_small);
    if (c->d_mtile_pow_small) cudaFree(c->d_mtile_pow_small);
    if (c->d_state) cudaFree(c->d_state);
    if (c->d_state_next) cudaFree(c->d_state_next);
    for (cudaEvent_t e : c->pipe_ev) if (e) cudaEventDestroy(e);
    if (c->h2d_stream) cudaStreamDestroy(c->h2d_stream);
    if (c->d2h_stream) cudaStreamDestroy(c->d2h_stream);
    if (c->d_ticket) cudaFree(c->d_ticket);
    if (c->d_ctrl) cudaFree(c->d_ctrl);
    if (c->d_tone) cudaFree(c->d_tone);
    if (c->h_ctrl) cudaFreeHost(c->h_ctrl);
    c->h_bounce.release();
    for (cudaEvent_t e : c->prof_ev) if (e) cudaEventDestroy(e);
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

void* caribou_cuda_get_stream(caribou_cuda_ctx* c) { return c ? (void*)c->stream : nullptr; }
uint64_t caribou_cuda_launch_count(const caribou_cuda_ctx* c) { return c ? c->launches : 0; }

int caribou_cuda_set_profiling(caribou_cuda_ctx* c, int enable)
{
    if (!c) { set_error("null context"); return CARIBOU_CUDA_ERR; }
    CBL_CHECK(cudaSetDevice(c->device));
    if (enable && c->prof_ev.empty()) {
        c->prof_ev.resize(2 * kProfRing, nullptr);
        for (cudaEvent_t& e : c->prof_ev) CBL_CHECK(cudaEventCreate(&e));
    }
    c->profiling = enable != 0;
    c->prof_n = 0;
    return CARIBOU_CUDA_OK;
}

int caribou_cuda_read_profile(caribou_cuda_ctx* c, float* ms, int cap)
{
    if (!c || !c->profiling || (!ms && cap > 0)) { set_error("profiling is not enabled"); return CARIBOU_CUDA_ERR; }
    CBL_CHECK(cudaSetDevice(c->device));
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    const uint64_t have = std::min<uint64_t>(c->prof_n, kProfRing);
    const uint64_t first = c->prof_n - have;
    int n = 0;
    for (uint64_t k = first; k < c->prof_n && n < cap; k++, n++) {
        float t = 0;
        CBL_CHECK(cudaEventElapsedTime(&t, c->prof_ev[2 * (k % kProfRing)], c->prof_ev[2 * (k % kProfRing) + 1]));
        ms[n] = t;
    }
    c->prof_n = 0;
    return n;
}

void* caribou_cuda_host_alloc(size_t nbytes)
{
    void* p = nullptr;
    if (!cuda_ok(cudaMallocHost(&p, nbytes), "cudaMallocHost")) return nullptr;
    return p;
}
void caribou_cuda_host_free(void* p) { if (p) cudaFreeHost(p); }

int caribou_cuda_butter_lowpass(int order, double fs, double fc, double* sos)
{
    return butter_lowpass_sos(order, fs, fc, sos);
}

// ------------------------------------------------------------------ unpack
static int unpack_device(caribou_cuda_ctx* c, const void* d_bytes, size_t nbytes, size_t stride_bytes,
                         const caribou_cuda_unpack_out& o, size_t out_stride_samples)
{
    const int mc = run_find_offsets(c, d_bytes, nbytes, stride_bytes, false);
    if (mc < 0) return mc;
    return run_unpack(c, d_bytes, nbytes, stride_bytes, o, (uint32_t)mc, out_stride_samples);
}

int caribou_cuda_unpack_submit(caribou_cuda_ctx* c, const void* bytes, size_t nbytes, size_t stream_stride_bytes,
                               int mem, const caribou_cuda_unpack_out* out)
{
    if (!c || !out || (!bytes && nbytes)) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    CBL_CHECK(cudaSetDevice(c->device));
    const size_t ns = (size_t)c->cfg.n_streams;
    if (ns > 1 && stream_stride_bytes < nbytes) { set_error("stream stride smaller than the stream"); return CARIBOU_CUDA_ERR; }
    if (ns == 1) stream_stride_bytes = (nbytes + 15) & ~size_t(15);
    if (resolve_queue(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    c->last_is_unpack = true;
    c->last_is_shard = false;
    c->sres_valid = false;
    c->last_nbytes = nbytes;
    c->last_n_filtered = 0;
    const int nsub = n_subs_of(c);
    // per-sub-stream output capacity in samples
    size_t sub_samples[2];
    for (int s = 0; s < nsub; s++) sub_samples[s] = (size_t)(sub_geom(c, bytes, nbytes, s).sv.nbytes / 4);
    // one kind for everything, or CARIBOU_CUDA_MEM_IO(in, out)
    const int in_mem = (mem & 0x100) ? (mem & 0xF) : mem, out_mem = (mem & 0x100) ? ((mem >> 4) & 0xF) : mem;
    if (in_mem < 0 || in_mem > 2 || out_mem < 0 || out_mem > 2) { set_error("bad memory kind"); return CARIBOU_CUDA_ERR; }
    if (in_mem == CARIBOU_CUDA_MEM_DEVICE && out_mem == CARIBOU_CUDA_MEM_DEVICE) {
        if (ns > 1 && (stream_stride_bytes % 16)) { set_error("stream stride must be a multiple of 16 bytes"); return CARIBOU_CUDA_ERR; }
        const size_t ostride = (sub_samples[0] + 3) & ~size_t(3);
        if (unpack_device(c, bytes, nbytes, stream_stride_bytes, *out, ns > 1 ? ostride : 0) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
        return mirror_counters(c);
    }
    // host memory: stage through device buffers (single stream only).  The input and the outputs may be of
    // different kinds (CARIBOU_CUDA_MEM_IO): the shims read(2) into a page-locked segment buffer and deliver
    // into the caller's pageable arrays.
    if (ns != 1) { set_error("host-memory unpack supports n_streams == 1"); return CARIBOU_CUDA_ERR; }
    const void* d_src = bytes;
    if (in_mem != CARIBOU_CUDA_MEM_DEVICE) {
        if (!c->stage_in.reserve(nbytes + 64)) return CARIBOU_CUDA_ERR;
        CBL_CHECK(cudaMemcpyAsync(c->stage_in.p, bytes, nbytes, cudaMemcpyHostToDevice, c->stream));
        d_src = c->stage_in.p;
    }
    if (out_mem == CARIBOU_CUDA_MEM_DEVICE) {        // host bytes, device arrays: nothing comes back but the counters
        if (unpack_device(c, d_src, nbytes, stream_stride_bytes, *out, 0) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
        return mirror_counters(c);
    }
    const int mc = run_find_offsets(c, d_src, nbytes, stream_stride_bytes, false);
    if (mc < 0) return mc;
    caribou_cuda_unpack_out d = *out;
    void* hcs16[2] = {out->cs16, out->cs16_b};
    void* hmeta[2] = {out->meta, out->meta_b};
    void* hcf32[2] = {out->cf32, out->cf32_b};
    void** dcs16[2] = {&d.cs16, &d.cs16_b};
    void** dmeta[2] = {&d.meta, &d.meta_b};
    void** dcf32[2] = {&d.cf32, &d.cf32_b};
    for (int s = 0; s < 2; s++) {
        *dcs16[s] = *dmeta[s] = *dcf32[s] = nullptr;
        if (s >= nsub) continue;
        const size_t n = sub_samples[s];
        if (hcs16[s]) { if (!c->stage_cs16[s].reserve(n * 4 + 64)) return CARIBOU_CUDA_ERR; *dcs16[s] = c->stage_cs16[s].p; }
        if (hmeta[s]) { if (!c->stage_meta[s].reserve(n + 64)) return CARIBOU_CUDA_ERR; *dmeta[s] = c->stage_meta[s].p; }
        if (hcf32[s]) { if (!c->stage_cf32[s].reserve(n * 8 + 64)) return CARIBOU_CUDA_ERR; *dcf32[s] = c->stage_cf32[s].p; }
    }
    if (out->invalid_idx && out->invalid_cap) { if (!c->stage_inv.reserve(out->invalid_cap * 4)) return CARIBOU_CUDA_ERR; d.invalid_idx = c->stage_inv.as<uint32_t>(); }
    if (out->sync_idx && out->sync_cap) { if (!c->stage_syn.reserve(out->sync_cap * 4)) return CARIBOU_CUDA_ERR; d.sync_idx = c->stage_syn.as<uint32_t>(); }

    if (out_mem == CARIBOU_CUDA_MEM_HOST) {
        // ---- pageable outputs: everything is queued behind the alignment search without waiting for its verdict.
        // The results come down into a page-locked bounce buffer together with the per-chunk offsets; after the
        // one synchronisation the host copies them into the caller's arrays - all of them for an aligned capture,
        // for a misaligned one exactly the slots the reference writes (the others keep the caller's bytes).
        size_t off_cs16[2], off_meta[2], off_cf32[2], total = 0;
        auto carve = [&](size_t bytes_) { const size_t o = total; total += (bytes_ + 63) & ~size_t(63); return o; };
        for (int s = 0; s < nsub; s++) {
            const size_t n = sub_samples[s];
            off_cs16[s] = carve(hcs16[s] ? n * 4 : 0);
            off_meta[s] = carve(hmeta[s] ? n : 0);
            off_cf32[s] = carve(hcf32[s] ? n * 8 : 0);
        }
        const size_t n_offs = (size_t)nsub * std::max(mc, 1);
        const size_t off_offs = carve(n_offs * 4);
        if (!c->h_bounce.reserve(total)) return CARIBOU_CUDA_ERR;
        uint8_t* hb = c->h_bounce.as<uint8_t>();
        if (run_unpack(c, d_src, nbytes, stream_stride_bytes, d, (uint32_t)mc, 0) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
        if (mirror_counters(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
        if (mc > 0) CBL_CHECK(cudaMemcpyAsync(hb + off_offs, c->offs.p, n_offs * 4, cudaMemcpyDeviceToHost, c->stream));
        for (int s = 0; s < nsub; s++) {
            const size_t n = sub_samples[s];
            if (hcs16[s] && n) CBL_CHECK(cudaMemcpyAsync(hb + off_cs16[s], *dcs16[s], n * 4, cudaMemcpyDeviceToHost, c->stream));
            if (hmeta[s] && n) CBL_CHECK(cudaMemcpyAsync(hb + off_meta[s], *dmeta[s], n, cudaMemcpyDeviceToHost, c->stream));
            if (hcf32[s] && n) CBL_CHECK(cudaMemcpyAsync(hb + off_cf32[s], *dcf32[s], n * 8, cudaMemcpyDeviceToHost, c->stream));
        }
        CBL_CHECK(cudaStreamSynchronize(c->stream));
        const bool clean = c->h_counters->any_misaligned == 0;
        const int32_t* h_offs = reinterpret_cast<const int32_t*>(hb + off_offs);
        const size_t cb = c->cfg.chunk_bytes, spc = cb / 4;
        for (int s = 0; s < nsub; s++) {
            const size_t n = sub_samples[s];
            if (clean) {
                if (hcs16[s]) memcpy(hcs16[s], hb + off_cs16[s], n * 4);
                if (hmeta[s]) memcpy(hmeta[s], hb + off_meta[s], n);
                if (hcf32[s]) memcpy(hcf32[s], hb + off_cf32[s], n * 8);
                continue;
            }
            const SubGeom g = sub_geom(c, bytes, nbytes, s);
            const int32_t fb = c->h_first_bad[s];
            for (uint32_t ch = 0; ch < g.n_chunks; ch++) {
                if ((int64_t)ch >= (int64_t)fb) break;                        // caribou_smi.c:665-668: nothing after a lost sync
                const uint64_t cstart = (uint64_t)ch * cb, rem = g.sv.nbytes - cstart;
                const size_t len = (size_t)std::min<uint64_t>(rem, cb), nfull = len / 4;
                const int32_t offs = h_offs[(size_t)s * mc + ch];
                const size_t shorten = offs > 0 ? (size_t)offs / 4 + 1 : 0;     // caribou_smi.c:323
                const size_t nvalid = nfull > shorten ? nfull - shorten : 0;
                const size_t extra = (shorten > 0 && nvalid >= 2) ? 1 : 0;      // the extrapolated sample (:382-389), no meta
                const size_t j0 = (size_t)ch * spc;
                if (hcs16[s]) memcpy((uint8_t*)hcs16[s] + j0 * 4, hb + off_cs16[s] + j0 * 4, (nvalid + extra) * 4);
                if (hmeta[s]) memcpy((uint8_t*)hmeta[s] + j0, hb + off_meta[s] + j0, nvalid);
                if (hcf32[s]) memcpy((uint8_t*)hcf32[s] + j0 * 8, hb + off_cf32[s] + j0 * 8, (nvalid + extra) * 8);
            }
        }
        const size_t ni = (size_t)std::min<uint64_t>(c->h_counters->totals[0], out->invalid_cap);
        const size_t nsy = (size_t)std::min<uint64_t>(c->h_counters->totals[1], out->sync_cap);
        if (out->invalid_idx && ni) CBL_CHECK(cudaMemcpyAsync(out->invalid_idx, d.invalid_idx, ni * 4, cudaMemcpyDeviceToHost, c->stream));
        if (out->sync_idx && nsy) CBL_CHECK(cudaMemcpyAsync(out->sync_idx, d.sync_idx, nsy * 4, cudaMemcpyDeviceToHost, c->stream));
        if ((out->invalid_idx && ni) || (out->sync_idx && nsy)) CBL_CHECK(cudaStreamSynchronize(c->stream));
        return CARIBOU_CUDA_OK;
    }

    // ---- page-locked outputs: asynchronous copies straight into the caller's buffers.  Which slots get written is
    // only known after the alignment search, so its verdict is waited for; a misaligned capture starts from the
    // caller's buffer contents so that untouched slots come back unchanged.
    if (mirror_counters(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    const bool clean = c->h_counters->any_misaligned == 0;
    if (!clean) {
        for (int s = 0; s < nsub; s++) {
            const size_t n = sub_samples[s];
            if (hcs16[s]) CBL_CHECK(cudaMemcpyAsync(*dcs16[s], hcs16[s], n * 4, cudaMemcpyHostToDevice, c->stream));
            if (hmeta[s]) CBL_CHECK(cudaMemcpyAsync(*dmeta[s], hmeta[s], n, cudaMemcpyHostToDevice, c->stream));
            if (hcf32[s]) CBL_CHECK(cudaMemcpyAsync(*dcf32[s], hcf32[s], n * 8, cudaMemcpyHostToDevice, c->stream));
        }
    }
    if (run_unpack(c, d_src, nbytes, stream_stride_bytes, d, (uint32_t)mc, 0) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (mirror_counters(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    for (int s = 0; s < nsub; s++) {
        const size_t n = sub_samples[s];
        if (hcs16[s]) CBL_CHECK(cudaMemcpyAsync(hcs16[s], *dcs16[s], n * 4, cudaMemcpyDeviceToHost, c->stream));
        if (hmeta[s]) CBL_CHECK(cudaMemcpyAsync(hmeta[s], *dmeta[s], n, cudaMemcpyDeviceToHost, c->stream));
        if (hcf32[s]) CBL_CHECK(cudaMemcpyAsync(hcf32[s], *dcf32[s], n * 8, cudaMemcpyDeviceToHost, c->stream));
    }
    if (out->invalid_idx || out->sync_idx) {
        CBL_CHECK(cudaStreamSynchronize(c->stream));
        const size_t ni = (size_t)std::min<uint64_t>(c->h_counters->totals[0], out->invalid_cap);
        const size_t nsy = (size_t)std::min<uint64_t>(c->h_counters->totals[1], out->sync_cap);
        if (out->invalid_idx && ni) CBL_CHECK(cudaMemcpyAsync(out->invalid_idx, d.invalid_idx, ni * 4, cudaMemcpyDeviceToHost, c->stream));
        if (out->sync_idx && nsy) CBL_CHECK(cudaMemcpyAsync(out->sync_idx, d.sync_idx, nsy * 4, cudaMemcpyDeviceToHost, c->stream));
        CBL_CHECK(cudaStreamSynchronize(c->stream));
    }
    return CARIBOU_CUDA_OK;
}

// ------------------------------------------------------------------ filter
static int check_filter_args(caribou_cuda_ctx* c, size_t n)
{
    if (!c->has_filter) { set_error("context has no filter (n_sos == 0)"); return CARIBOU_CUDA_ERR; }
    if (n % (size_t)c->plan.D != 0) { set_error("n_samples must be a multiple of the decimation factor"); return CARIBOU_CUDA_ERR; }
    return CARIBOU_CUDA_OK;
}

}  // extern "C"

namespace cbl {
// single-stream filter pass on device memory, skipped on the device when *gate != 0 (compat.cu: a read that lost
// word sync delivers nothing and must not advance the Stream's filters)
int filter_device(caribou_cuda_ctx* c, int in_kind, const void* in, size_t n, void* out, const int32_t* gate)
{
    if (!c || !in || !out || c->cfg.n_streams != 1) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    CBL_CHECK(cudaSetDevice(c->device));
    if (check_filter_args(c, n) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (resolve_queue(c) != CARIBOU_CUDA_OK || ensure_scale(c, in_kind) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    c->last_is_unpack = false;
    c->last_is_shard = false;
    c->sres_valid = false;
    c->last_ret = (int64_t)n;
    c->last_n_filtered = (int64_t)(n / c->plan.D);
    if (n == 0) return CARIBOU_CUDA_OK;
    if (run_iir(c, in_kind, in, n, (n + 3) & ~size_t(3), out, n / c->plan.D, gate, 0, true, 0, nullptr) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    c->pos += n;
    return CARIBOU_CUDA_OK;
}
}  // namespace cbl

extern "C" {

int caribou_cuda_filter_submit(caribou_cuda_ctx* c, int in_kind, const void* in, size_t n, size_t in_stride,
                               int mem, void* out, size_t out_stride)
{
    if (!c || (!in && n) || (!out && n) || in_kind < 0 || in_kind > 2) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    CBL_CHECK(cudaSetDevice(c->device));
    if (check_filter_args(c, n) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (ensure_scale(c, in_kind) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    const size_t ns = (size_t)c->cfg.n_streams;
    const size_t es = in_elem_bytes(in_kind);
    const size_t nout = n / c->plan.D;
    if (resolve_queue(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    c->last_is_unpack = false;
    c->last_is_shard = false;
    c->sres_valid = false;
    c->last_ret = (int64_t)n;
    c->last_n_filtered = (int64_t)nout;
    if (n == 0) return CARIBOU_CUDA_OK;
    if (ns == 1) { in_stride = (n + 3) & ~size_t(3); out_stride = nout; }
    if (mem == CARIBOU_CUDA_MEM_DEVICE) {
        if (run_iir(c, in_kind, in, n, in_stride, out, out_stride, nullptr, 0, true, 0, nullptr) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
        c->pos += n;
        return CARIBOU_CUDA_OK;
    }
    if (ns != 1) { set_error("host-memory filter supports n_streams == 1"); return CARIBOU_CUDA_ERR; }
    if (!c->stage_in.reserve(n * es + 64) || !c->stage_filt.reserve(nout * c->out_elem + 64)) return CARIBOU_CUDA_ERR;
    CBL_CHECK(cudaMemcpyAsync(c->stage_in.p, in, n * es, cudaMemcpyHostToDevice, c->stream));
    if (run_iir(c, in_kind, c->stage_in.p, n, in_stride, c->stage_filt.p, out_stride, nullptr, 0, true, 0, nullptr) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    CBL_CHECK(cudaMemcpyAsync(out, c->stage_filt.p, nout * c->out_elem, cudaMemcpyDeviceToHost, c->stream));
    c->pos += n;
    if (mem == CARIBOU_CUDA_MEM_HOST) CBL_CHECK(cudaStreamSynchronize(c->stream));
    return CARIBOU_CUDA_OK;
}

int caribou_cuda_filter_fixup(caribou_cuda_ctx* c, int in_kind, const void* in, size_t n, size_t in_stride, int mem,
                              void* out, size_t out_stride, const double* state_in)
{
    if (!c || !in || !out || !state_in || in_kind < 0 || in_kind > 2) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    if (mem != CARIBOU_CUDA_MEM_DEVICE) { set_error("fix-up works on device memory"); return CARIBOU_CUDA_ERR; }
    CBL_CHECK(cudaSetDevice(c->device));
    if (check_filter_args(c, n) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (resolve_queue(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (ensure_scale(c, in_kind) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    const FilterPlan& P = c->plan;
    const size_t ns = (size_t)c->cfg.n_streams;
    if (ns == 1) { in_stride = (n + 3) & ~size_t(3); out_stride = n / P.D; }
    // the shard was filtered from zero state and left f in d_state; its true end state is M_shard s_in + f
    std::vector<double> f(ns * 2 * P.S), ker(ns * 2 * P.S), m((size_t)P.S * P.S);
    CBL_CHECK(cudaMemcpyAsync(f.data(), c->d_state, f.size() * 8, cudaMemcpyDeviceToHost, c->stream));
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    for (size_t s = 0; s < ns; s++) state_to_kernel(P, state_in + s * 2 * P.S, ker.data() + s * 2 * P.S);
    for (double& v : ker) v /= c->in_scale;
    plan_power_kernel(P, n, m.data());     // A^n in kernel coordinates
    std::vector<double> end(f);
    for (size_t s = 0; s < ns; s++)
        for (int r = 0; r < 2; r++)
            for (int i = 0; i < P.S; i++) {
                long double acc = f[s * 2 * P.S + r * P.S + i];
                for (int j = 0; j < P.S; j++) acc += (long double)m[(size_t)i * P.S + j] * ker[s * 2 * P.S + r * P.S + j];
                end[s * 2 * P.S + r * P.S + i] = (double)acc;
            }
    // re-run the leading tiles from the true incoming state
    CBL_CHECK(cudaMemcpyAsync(c->d_state, ker.data(), ker.size() * 8, cudaMemcpyHostToDevice, c->stream));
    CBL_CHECK(cudaStreamSynchronize(c->stream));       // `ker` is pageable: the copy has landed before it is reused
    const size_t fix = caribou_cuda_fixup_samples(c, n);
    const int LTn = pick_plan(c, n).LT;                  // the plan run_iir will take for this shard
    const uint32_t tile_samples = 32u * (uint32_t)LTn;
    uint32_t max_tiles = (uint32_t)((fix + tile_samples - 1) / tile_samples);
    const uint64_t tiles_ps = (n / LTn + 31) / 32;
    if (max_tiles == 0 || max_tiles >= tiles_ps) {
        // no useful contraction (or a short shard): filter the whole shard again from the true state;
        // the kernels then leave the true end state in d_state themselves
        if (run_iir(c, in_kind, in, n, in_stride, out, out_stride, nullptr, 0, true, 0, nullptr) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
        CBL_CHECK(cudaStreamSynchronize(c->stream));
        return CARIBOU_CUDA_OK;
    }
    if (run_iir(c, in_kind, in, n, in_stride, out, out_stride, nullptr, 0, false, max_tiles, nullptr) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    CBL_CHECK(cudaMemcpyAsync(c->d_state, end.data(), end.size() * 8, cudaMemcpyHostToDevice, c->stream));
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    return CARIBOU_CUDA_OK;
}

size_t caribou_cuda_fixup_samples(caribou_cuda_ctx* c, size_t n_samples)
{
    if (!c || !c->has_filter) return 0;
    if (c->plan.chained || c->plan.decay_samples == 0) return n_samples;
    return (size_t)std::min<uint64_t>(c->plan.decay_samples, n_samples);
}

int caribou_cuda_shard_matrix(caribou_cuda_ctx* c, uint64_t n_samples, double* m)
{
    if (!c || !c->has_filter || !m) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    plan_shard_matrix(c->plan, n_samples, m);
    return CARIBOU_CUDA_OK;
}

// ------------------------------------------------------------------ fused rx path
#ifndef CBL_RX_GROUPS
#define CBL_RX_GROUPS 32         // groups of streams a page-locked host batch is pipelined in (fill and drain are 1/groups)
#endif
// Every capture of a batch is an independent read.  The fused kernel is launched optimistically for all of
// them; the alignment search that runs just before it raises a per-capture gate (sticky until the queue is
// resolved) for a capture with a misaligned or un-alignable chunk, and the fused kernel, the state commit and
// the tail kernel skip exactly those captures.  Submits are queued; caribou_cuda_wait (or the next call that
// needs a quiet context) redoes the gated captures - and only those - on the slow path for every queued
// submit from the one that raised the gate on, in order, so their filter state advances as it would have.
int caribou_cuda_rx_submit(caribou_cuda_ctx* c, const void* bytes, size_t nbytes, size_t stride_bytes, int mem,
                           const caribou_cuda_unpack_out* uo, void* filt_out, size_t filt_stride)
{
    if (!c || (!bytes && nbytes)) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    CBL_CHECK(cudaSetDevice(c->device));
    if (!c->has_filter || !filt_out) { set_error("rx_submit needs a filter and an output buffer; use unpack_submit otherwise"); return CARIBOU_CUDA_ERR; }
    const size_t n = nbytes / 4;
    if (check_filter_args(c, n) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (shard_flush(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (ensure_scale(c, IN_WORDS) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (c->queue.size() >= kMaxQueuedRx && resolve_queue(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    const FilterPlan& P = c->plan;
    const size_t ns = (size_t)c->cfg.n_streams;
    const size_t nout = n / P.D;
    if (ns == 1) { stride_bytes = (nbytes + 15) & ~size_t(15); filt_stride = nout; }
    if (stride_bytes % 16) { set_error("stream stride must be a multiple of 16 bytes"); return CARIBOU_CUDA_ERR; }
    c->last_is_unpack = false;
    c->last_is_shard = false;
    c->sres_valid = false;
    c->last_nbytes = nbytes;
    c->last_ret = (int64_t)n;
    c->last_n_filtered = (int64_t)nout;
    if (n == 0) return CARIBOU_CUDA_OK;

    const void* d_bytes = bytes;
    void* d_out = filt_out;
    const int32_t* gates = c->d_sflags;
    PendingRx q;
    q.seq = (int32_t)c->queue.size();
    q.bytes = bytes; q.nbytes = nbytes; q.stride = stride_bytes; q.mem = mem;
    q.filt_out = filt_out; q.filt_stride = filt_stride;
    if (mem != CARIBOU_CUDA_MEM_DEVICE) {
        if (uo) { set_error("rx_submit with unpack outputs needs device memory"); return CARIBOU_CUDA_ERR; }
        if (ns > 1 && (stride_bytes < nbytes || filt_stride < nout)) { set_error("stream stride smaller than the stream"); return CARIBOU_CUDA_ERR; }
        if (!c->stage_in.reserve(ns * stride_bytes + 64) || !c->stage_filt.reserve(ns * filt_stride * c->out_elem + 64)) return CARIBOU_CUDA_ERR;
        d_bytes = c->stage_in.p;
        d_out = c->stage_filt.p;
        q.d_bytes = d_bytes; q.d_out = d_out;
        if (mem == CARIBOU_CUDA_MEM_HOST_PINNED && ns >= 4) {
            // ---- pipelined: the batch goes through in groups of streams; group g+1 is on its way up while
            // group g is being filtered and group g-1 is on its way down (both copy engines and the SMs busy)
            const int kGroups = CBL_RX_GROUPS;
            const size_t per = (ns + kGroups - 1) / kGroups;
            const int ng = (int)((ns + per - 1) / per);
            if (!c->h2d_stream) {
                CBL_CHECK(cudaStreamCreateWithFlags(&c->h2d_stream, cudaStreamNonBlocking));
                CBL_CHECK(cudaStreamCreateWithFlags(&c->d2h_stream, cudaStreamNonBlocking));
                c->pipe_ev.resize(2 * kGroups + 2, nullptr);
                for (cudaEvent_t& e : c->pipe_ev) CBL_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            }
            cudaEvent_t ev_start = c->pipe_ev[2 * kGroups], ev_done = c->pipe_ev[2 * kGroups + 1];
            // the staging buffers may still be in use by the previous submit on the main stream
            CBL_CHECK(cudaEventRecord(ev_start, c->stream));
            CBL_CHECK(cudaStreamWaitEvent(c->h2d_stream, ev_start, 0));
            CBL_CHECK(cudaStreamWaitEvent(c->d2h_stream, ev_start, 0));
            const uint8_t* hb = reinterpret_cast<const uint8_t*>(bytes);
            uint8_t* ho = reinterpret_cast<uint8_t*>(filt_out);
            uint8_t* db = reinterpret_cast<uint8_t*>(c->stage_in.p);
            uint8_t* dob = reinterpret_cast<uint8_t*>(c->stage_filt.p);
            int rc = CARIBOU_CUDA_OK;
            for (int g = 0; g < ng && rc == CARIBOU_CUDA_OK; g++) {
                const size_t s0 = (size_t)g * per, cnt = std::min(per, ns - s0);
                CBL_CHECK(cudaMemcpyAsync(db + s0 * stride_bytes, hb + s0 * stride_bytes, (cnt - 1) * stride_bytes + nbytes,
                                          cudaMemcpyHostToDevice, c->h2d_stream));
                CBL_CHECK(cudaEventRecord(c->pipe_ev[2 * g], c->h2d_stream));
                CBL_CHECK(cudaStreamWaitEvent(c->stream, c->pipe_ev[2 * g], 0));
                c->win0 = (uint32_t)s0; c->win_cnt = (uint32_t)cnt;
                const int mcg = run_find_offsets(c, db + s0 * stride_bytes, nbytes, stride_bytes, true, n);
                if (mcg < 0) { rc = mcg; break; }
                rc = run_iir(c, IN_WORDS, db + s0 * stride_bytes, n, stride_bytes / 4, dob + s0 * filt_stride * c->out_elem, filt_stride,
                             gates + s0, 1, true, 0, nullptr);
                if (rc != CARIBOU_CUDA_OK) break;
                CBL_CHECK(cudaEventRecord(c->pipe_ev[2 * g + 1], c->stream));
                CBL_CHECK(cudaStreamWaitEvent(c->d2h_stream, c->pipe_ev[2 * g + 1], 0));
                CBL_CHECK(cudaMemcpyAsync(ho + s0 * filt_stride * c->out_elem, dob + s0 * filt_stride * c->out_elem,
                                          ((cnt - 1) * filt_stride + nout) * c->out_elem, cudaMemcpyDeviceToHost, c->d2h_stream));
            }
            c->win0 = 0; c->win_cnt = (uint32_t)ns;
            if (rc != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
            CBL_CHECK(cudaEventRecord(ev_done, c->d2h_stream));
            CBL_CHECK(cudaStreamWaitEvent(c->stream, ev_done, 0));      // waiting for the context now covers the copies
            if (mirror_counters(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
            c->queue.push_back(q);
            c->pos += n;
            return CARIBOU_CUDA_OK;
        }
        // the batch keeps its host layout (same strides) in the device staging buffers
        const size_t in_span = (ns - 1) * stride_bytes + nbytes;
        CBL_CHECK(cudaMemcpyAsync(c->stage_in.p, bytes, in_span, cudaMemcpyHostToDevice, c->stream));
    }
    q.d_bytes = d_bytes; q.d_out = d_out;
    const int mc = run_find_offsets(c, d_bytes, nbytes, stride_bytes, true, n);
    if (mc < 0) return mc;
    // optimistic fused launch: captures whose gate the search just raised are skipped on the device
    c->mirror_in_tail = uo == nullptr;          // the submit's last launch also mirrors the control block
    if (run_iir(c, IN_WORDS, d_bytes, n, stride_bytes / 4, d_out, filt_stride, gates, 1, true, 0, nullptr) != CARIBOU_CUDA_OK)
        return CARIBOU_CUDA_ERR;
    if (uo) {
        const size_t ostride = ns > 1 ? ((n + 3) & ~size_t(3)) : 0;
        if (run_unpack(c, d_bytes, nbytes, stride_bytes, *uo, (uint32_t)mc, ostride) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    }
    if (c->mirror_in_tail || uo) {              // (no tail launch happened, or the unpack kernels counted after it)
        c->mirror_in_tail = false;
        if (mirror_counters(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    }
    if (mem != CARIBOU_CUDA_MEM_DEVICE)
        CBL_CHECK(cudaMemcpyAsync(filt_out, d_out, ((ns - 1) * filt_stride + nout) * c->out_elem, cudaMemcpyDeviceToHost, c->stream));
    c->queue.push_back(q);
    c->pos += n;
    if (mem == CARIBOU_CUDA_MEM_HOST) return resolve_queue(c);
    return CARIBOU_CUDA_OK;
}

}  // extern "C"

namespace cbl {

// One capture (stream s) of one queued submit on the slow path: decode into a zero-filled CS16 scratch with the
// reference's chunk semantics, then filter what the reference's read loop would have delivered.  Synchronous.
static int rx_replay_stream(caribou_cuda_ctx* c, const PendingRx& q, size_t s, caribou_cuda_stream_result* r)
{
    const FilterPlan& P = c->plan;
    const size_t ns = (size_t)c->cfg.n_streams;
    const size_t n = q.nbytes / 4;
    const size_t sstride = (n + 3) & ~size_t(3);
    if (!c->scratch_cs16.reserve(sstride * 4 + 64)) return CARIBOU_CUDA_ERR;
    const uint8_t* d_bytes = reinterpret_cast<const uint8_t*>(q.d_bytes);
    uint8_t* d_out = reinterpret_cast<uint8_t*>(q.d_out);
    if (q.mem != CARIBOU_CUDA_MEM_DEVICE) {
        // later submits have reused the staging buffers: fetch this capture again
        if (!c->stage_in.reserve(ns * q.stride + 64) || !c->stage_filt.reserve(ns * q.filt_stride * c->out_elem + 64)) return CARIBOU_CUDA_ERR;
        d_bytes = c->stage_in.as<uint8_t>();
        d_out = c->stage_filt.as<uint8_t>();
        CBL_CHECK(cudaMemcpyAsync(c->stage_in.as<uint8_t>() + s * q.stride, reinterpret_cast<const uint8_t*>(q.bytes) + s * q.stride,
                                  q.nbytes, cudaMemcpyHostToDevice, c->stream));
    }
    CBL_CHECK(cudaMemsetAsync(c->scratch_cs16.p, 0, sstride * 4, c->stream));
    CBL_CHECK(cudaMemsetAsync(c->d_counters, 0, sizeof(Counters), c->stream));
    c->win0 = (uint32_t)s; c->win_cnt = 1;
    auto done = [&](int rc) { c->win0 = 0; c->win_cnt = (uint32_t)ns; return rc; };
    caribou_cuda_unpack_out o;
    memset(&o, 0, sizeof(o));
    // run_unpack addresses stream s of the batch at out + s * out_stride: hand it a base that lands on the scratch
    o.cs16 = c->scratch_cs16.as<uint32_t>() - s * sstride;
    const int mc = run_find_offsets(c, d_bytes + s * q.stride, q.nbytes, q.stride, false);
    if (mc < 0) return done(mc);
    if (run_unpack(c, d_bytes, q.nbytes, q.stride, o, (uint32_t)mc, sstride) != CARIBOU_CUDA_OK) return done(CARIBOU_CUDA_ERR);
    if (mirror_counters(c) != CARIBOU_CUDA_OK) return done(CARIBOU_CUDA_ERR);
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    // samples delivered before a lost sync (caribou_smi.c:664-677)
    int64_t deliver = (int64_t)n;
    bool lost = false;
    const int32_t fb = c->h_first_bad[s * 2];
    const uint64_t cb = c->cfg.chunk_bytes;
    if (fb != INT_MAX && (uint64_t)fb < (q.nbytes + cb - 1) / cb) {
        lost = true;
        deliver = (int64_t)fb * (int64_t)(cb / 4);
    }
    deliver -= deliver % P.D;
    if (deliver > 0) {
        if (run_iir(c, IN_CS16, c->scratch_cs16.p, (size_t)deliver, sstride, d_out + s * q.filt_stride * c->out_elem, q.filt_stride,
                    nullptr, 0, true, 0, nullptr) != CARIBOU_CUDA_OK)
            return done(CARIBOU_CUDA_ERR);
        if (q.mem != CARIBOU_CUDA_MEM_DEVICE)
            CBL_CHECK(cudaMemcpyAsync(reinterpret_cast<uint8_t*>(q.filt_out) + s * q.filt_stride * c->out_elem,
                                      d_out + s * q.filt_stride * c->out_elem, (size_t)deliver / P.D * c->out_elem,
                                      cudaMemcpyDeviceToHost, c->stream));
    }
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    if (r) {
        r->n_samples = lost ? CARIBOU_CUDA_ERR_SYNC : deliver;
        r->n_filtered = deliver / P.D;
        r->first_bad_chunk = lost ? fb : -1;
        r->n_misaligned_chunks = c->h_counters->n_misaligned;
    }
    return done(CARIBOU_CUDA_OK);
}

int resolve_queue(caribou_cuda_ctx* c)
{
    if (shard_flush(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (c->queue.empty()) return CARIBOU_CUDA_OK;
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    const size_t ns = (size_t)c->cfg.n_streams;
    const PendingRx last = c->queue.back();
    const size_t n = last.nbytes / 4;
    // results of the last submit as the fused path left them (a gated capture is overwritten by its replay)
    std::vector<int32_t> gate(c->h_sflags, c->h_sflags + ns), gseq(c->h_sflags + ns, c->h_sflags + 2 * ns);
    Counters cnt = *c->h_counters;
    bool any = false;
    for (size_t s = 0; s < ns; s++) {
        c->sres[s].n_samples = (int64_t)n;
        c->sres[s].n_filtered = (int64_t)(n / c->plan.D);
        c->sres[s].first_bad_chunk = -1;
        c->sres[s].n_misaligned_chunks = c->h_sflags[2 * ns + s];
        any = any || gate[s] != 0;
    }
    std::vector<PendingRx> queue;
    queue.swap(c->queue);              // run_find_offsets stamps submits with the queue length: empty from here on
    if (any) {
        for (size_t s = 0; s < ns; s++) {
            if (!gate[s]) continue;
            for (const PendingRx& q : queue) {
                if (q.seq < gseq[s]) continue;
                if (rx_replay_stream(c, q, s, q.seq == last.seq ? &c->sres[s] : nullptr) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
            }
        }
        CBL_CHECK(cudaMemsetAsync(c->d_sflags, 0, ns * 4, c->stream));
        CBL_CHECK(cudaMemsetAsync(c->d_sflags + ns, 0x7f, ns * 4, c->stream));
        CBL_CHECK(cudaStreamSynchronize(c->stream));
        memset(c->h_sflags, 0, ns * 4);
        // the batch-level counters describe the last submit as a whole again
        *c->h_counters = cnt;
    }
    int64_t ret = (int64_t)n, nf = (int64_t)(n / c->plan.D);
    for (size_t s = 0; s < ns; s++) {
        if (c->sres[s].n_samples == CARIBOU_CUDA_ERR_SYNC) ret = CARIBOU_CUDA_ERR_SYNC;
        else if (ret >= 0) ret = std::min(ret, c->sres[s].n_samples);
        nf = std::min(nf, c->sres[s].n_filtered);
    }
    c->last_ret = ret;
    c->last_n_filtered = nf;
    c->sres_valid = true;
    return CARIBOU_CUDA_OK;
}

}  // namespace cbl

extern "C" {

int64_t caribou_cuda_wait(caribou_cuda_ctx* c, caribou_cuda_result* res)
{
    if (!c) { set_error("null context"); return CARIBOU_CUDA_ERR; }
    if (!cuda_ok(cudaSetDevice(c->device), "cudaSetDevice") || shard_flush(c) != CARIBOU_CUDA_OK ||
        !cuda_ok(cudaStreamSynchronize(c->stream), "cudaStreamSynchronize"))
        return CARIBOU_CUDA_ERR;
    if (resolve_queue(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    int64_t ret = c->last_ret, ret_b = 0;
    if (c->last_is_shard && c->h_counters->any_misaligned != 0) ret = CARIBOU_CUDA_ERR_SYNC;   // shards are aligned chunks or nothing
    if (c->last_is_shard && c->h_sflags[3 * (size_t)c->cfg.n_streams + 1] != 0) {
        set_error("time sharding: a peer's end vector never arrived");
        return CARIBOU_CUDA_ERR;
    }
    if (c->last_is_unpack) {
        const int nsub = n_subs_of(c);
        int64_t r[2] = {0, 0};
        for (int s = 0; s < nsub; s++) {
            const SubGeom g = sub_geom(c, nullptr, c->last_nbytes, s);
            int32_t fb = INT_MAX;
            for (int st = 0; st < c->cfg.n_streams; st++) fb = std::min(fb, c->h_first_bad[st * 2 + s]);
            r[s] = delivered_samples(c, g.sv.nbytes, fb == INT_MAX ? -1 : fb);
        }
        ret = r[0]; ret_b = r[1];
        if (nsub == 2 && ret_b < 0) ret = ret_b;
        c->last_ret = ret;
    }
    if (res) {
        memset(res, 0, sizeof(*res));
        res->n_samples = ret;
        res->n_samples_b = ret_b;
        res->n_filtered = c->last_n_filtered;
        res->n_invalid = c->h_counters->totals[0];
        res->n_sync = c->h_counters->totals[1];
        res->first_bad_chunk = -1;
        res->n_misaligned_chunks = c->h_counters->n_misaligned;
        if (c->sres_valid) {
            int32_t nm = 0;
            for (const caribou_cuda_stream_result& r : c->sres) {
                if (r.first_bad_chunk >= 0 && (res->first_bad_chunk < 0 || r.first_bad_chunk < res->first_bad_chunk)) res->first_bad_chunk = r.first_bad_chunk;
                nm += r.n_misaligned_chunks;
            }
            res->n_misaligned_chunks = nm;
        } else {
            int32_t fb = INT_MAX;
            for (int i = 0; i < c->cfg.n_streams * 2; i++) fb = std::min(fb, c->h_first_bad[i]);
            const uint64_t cb = c->cfg.chunk_bytes;
            res->first_bad_chunk = (fb != INT_MAX && (uint64_t)fb < (c->last_nbytes + cb - 1) / cb) ? fb : -1;
        }
    }
    return ret;
}

int caribou_cuda_stream_results(caribou_cuda_ctx* c, caribou_cuda_stream_result* out)
{
    if (!c || !out) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    if (caribou_cuda_wait(c, nullptr) == CARIBOU_CUDA_ERR) return CARIBOU_CUDA_ERR;
    if (!c->sres_valid) { set_error("per-capture results exist after caribou_cuda_rx_submit only"); return CARIBOU_CUDA_ERR; }
    for (int s = 0; s < c->cfg.n_streams; s++) out[s] = c->sres[s];
    return CARIBOU_CUDA_OK;
}


int caribou_cuda_filter_info(caribou_cuda_ctx* c, caribou_cuda_filter_info_t* info)
{
    if (!c || !info) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    memset(info, 0, sizeof(*info));
    if (!c->has_filter) return CARIBOU_CUDA_OK;
    const FilterPlan& P = c->plan;
    info->n_sos = P.K; info->decimation = P.D; info->run_len = P.LT; info->tile_samples = 32 * P.LT;
    info->form = P.par ? CARIBOU_CUDA_FORM_PARALLEL : CARIBOU_CUDA_FORM_CASCADE;
    info->lookback_window = P.window; info->chained = P.chained ? 1 : 0;
    info->decay_samples = P.decay_samples;
    info->parallel_form_error = P.par_err;
    if (P.rerun && P.ap)
        snprintf(info->note, sizeof(info->note), "integer inputs: numerator as integer two-tap sums + all-pole cascade (%.1e)", P.ap_err);
    else
        snprintf(info->note, sizeof(info->note), "%s", P.form_note);
    return CARIBOU_CUDA_OK;
}

// ------------------------------------------------------------------ state
int caribou_cuda_get_state(caribou_cuda_ctx* c, int stream, double* state)
{
    if (!c || !c->has_filter || !state || stream < 0 || stream >= c->cfg.n_streams) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    if (caribou_cuda_wait(c, nullptr) == CARIBOU_CUDA_ERR) return CARIBOU_CUDA_ERR;
    const FilterPlan& P = c->plan;
    double ker[2 * kMaxS];
    CBL_CHECK(cudaMemcpyAsync(ker, c->d_state + (size_t)stream * 2 * P.S, (size_t)2 * P.S * 8, cudaMemcpyDeviceToHost, c->stream));
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    state_to_ref(P, ker, state);
    for (int i = 0; i < 2 * P.S; i++) state[i] *= c->in_scale;
    return CARIBOU_CUDA_OK;
}

int caribou_cuda_set_state(caribou_cuda_ctx* c, int stream, const double* state)
{
    if (!c || !c->has_filter || !state || stream < 0 || stream >= c->cfg.n_streams) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    if (caribou_cuda_wait(c, nullptr) == CARIBOU_CUDA_ERR) return CARIBOU_CUDA_ERR;
    const FilterPlan& P = c->plan;
    double ker[2 * kMaxS];
    state_to_kernel(P, state, ker);
    for (int i = 0; i < 2 * P.S; i++) ker[i] /= c->in_scale;
    // ordered on the context's stream (the kernels that read the state run there), waited for because `ker` is a local
    CBL_CHECK(cudaMemcpyAsync(c->d_state + (size_t)stream * 2 * P.S, ker, (size_t)2 * P.S * 8, cudaMemcpyHostToDevice, c->stream));
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    return CARIBOU_CUDA_OK;
}

int caribou_cuda_reset(caribou_cuda_ctx* c)
{
    if (!c) { set_error("null context"); return CARIBOU_CUDA_ERR; }
    CBL_CHECK(cudaSetDevice(c->device));
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    // whatever was queued is dropped with the state it would have advanced
    const size_t ns = (size_t)c->cfg.n_streams;
    c->queue.clear();
    CBL_CHECK(cudaMemsetAsync(c->d_sflags, 0, ns * 4, c->stream));
    CBL_CHECK(cudaMemsetAsync(c->d_sflags + ns, 0x7f, ns * 4, c->stream));
    if (c->has_filter) CBL_CHECK(cudaMemsetAsync(c->d_state, 0, ns * 2 * c->plan.S * 8, c->stream));
    CBL_CHECK(cudaStreamSynchronize(c->stream));
    memset(c->h_sflags, 0, ns * 3 * 4);
    c->pos = 0;
    return CARIBOU_CUDA_OK;
}

// ------------------------------------------------------------------ TX direction (kernel (c), tx.cu)
int caribou_cuda_tx_pack_submit(caribou_cuda_ctx* c, int in_format, const void* samples, size_t n, int mem, void* words)
{
    if (!c || (n && (!samples || !words))) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    size_t es = 0;
    switch (in_format) {
        case CARIBOU_CUDA_FMT_CS16: es = 4; break;
        case CARIBOU_CUDA_FMT_CS8:  es = 2; break;
        case CARIBOU_CUDA_FMT_CF32: es = 8; break;
        case CARIBOU_CUDA_FMT_CF64: es = 16; break;
        default: set_error("unknown sample format"); return CARIBOU_CUDA_ERR;
    }
    CBL_CHECK(cudaSetDevice(c->device));
    if (resolve_queue(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    c->last_is_unpack = false;
    c->last_is_shard = false;
    c->sres_valid = false;
    c->last_n_filtered = 0;
    c->last_ret = (int64_t)n;
    if (n == 0) return CARIBOU_CUDA_OK;
    TxParams p;
    memset(&p, 0, sizeof(p));
    p.n = n;
    p.fmt = in_format;
    if (mem == CARIBOU_CUDA_MEM_DEVICE) {
        if ((reinterpret_cast<uintptr_t>(samples) | reinterpret_cast<uintptr_t>(words)) & 15) {
            set_error("device buffers must be 16-byte aligned");
            return CARIBOU_CUDA_ERR;
        }
        p.in = samples;
        p.out = static_cast<uint32_t*>(words);
    } else {
        if (!c->stage_in.reserve(n * es + 64) || !c->stage_filt.reserve(n * 4 + 64)) return CARIBOU_CUDA_ERR;
        CBL_CHECK(cudaMemcpyAsync(c->stage_in.p, samples, n * es, cudaMemcpyHostToDevice, c->stream));
        p.in = c->stage_in.p;
        p.out = c->stage_filt.as<uint32_t>();
    }
    prof_mark(c, 0);
    const int e = launch_tx_pack(p, c->num_sms, c->stream);
    prof_mark(c, 1);
    if (e != 0) { cuda_ok((cudaError_t)e, "tx_pack_kernel launch"); return CARIBOU_CUDA_ERR; }
    c->launches++;
    if (mem != CARIBOU_CUDA_MEM_DEVICE) {
        CBL_CHECK(cudaMemcpyAsync(words, p.out, n * 4, cudaMemcpyDeviceToHost, c->stream));
        if (mem == CARIBOU_CUDA_MEM_HOST) CBL_CHECK(cudaStreamSynchronize(c->stream));
    }
    return CARIBOU_CUDA_OK;
}

// ------------------------------------------------------------------ synthetic words
int caribou_cuda_synth_words(caribou_cuda_ctx* c, void* dev_words, uint64_t n0, size_t n, int hif_layout, uint64_t seed,
                             const int16_t* tone_i, const int16_t* tone_q)
{
    if (!c || !dev_words || !tone_i || !tone_q) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    CBL_CHECK(cudaSetDevice(c->device));
    if (!c->d_tone) CBL_CHECK(cudaMalloc(&c->d_tone, 1600 * sizeof(int16_t)));
    CBL_CHECK(cudaMemcpyAsync(c->d_tone, tone_i, 800 * 2, cudaMemcpyHostToDevice, c->stream));
    CBL_CHECK(cudaMemcpyAsync(c->d_tone + 800, tone_q, 800 * 2, cudaMemcpyHostToDevice, c->stream));
    launch_synth(reinterpret_cast<uint32_t*>(dev_words), n0, n, hif_layout, seed, c->d_tone, c->d_tone + 800, c->num_sms, c->stream);
    c->launches++;
    CBL_CHECK(cudaGetLastError());
    return CARIBOU_CUDA_OK;
}

}  // extern "C"
