// Drop-in shims: caribou_smi_read / caribou_smi_get_native_batch_samples shaped entry points fed from a
// file descriptor, and the RX half of SoapySDR::Stream (format switch, MTU clamp, never-reset filters).
// All sample arithmetic runs in the CUDA kernels; the host side only moves bytes and keeps the
// reference's control flow (LC/caribou_smi/caribou_smi.c:396-492,624-682; LC/soapy_api/CaribouliteStream.cpp).
#include <errno.h>
#include <limits.h>
#include <poll.h>
#include <string.h>
#include <unistd.h>
#include <algorithm>
#include <functional>
#include <new>
#include "ctx.h"

using namespace cbl;

struct caribou_cuda_smi {
    int fd = -1;
    size_t native_batch_len = CARIBOU_CUDA_NATIVE_BATCH_LEN;
    int device = 0;
    uint32_t sample_rate = 4000000;            // CARIBOU_SMI_SAMPLE_RATE
    caribou_cuda_ctx* ctx[2] = {nullptr, nullptr};   // one unpack context per channel layout
    uint8_t* h_stage = nullptr;                 // pinned: the bytes read(2) delivers
    size_t h_cap = 0;
    cudaEvent_t stage_ev = nullptr;             // the last asynchronous upload out of h_stage
    bool stage_busy = false;
    // CaribouLiteRadio's _read_samples / _read_metadata (LC/CaribouLite.hpp): persistent native buffers of the C++ API
    DevBuf radio_cs16, radio_meta, radio_cf32;
    HostBuf radio_bounce;
};

namespace {

// caribou_smi_calc_read_timeout, caribou_smi.c:624-629
int calc_read_timeout(uint32_t sample_rate, size_t len)
{
    uint32_t to = (uint32_t)((2 * len * 1000) / sample_rate);
    if (to < 1) to = 1;
    return (int)(to * 2);
}

// caribou_smi_poll, caribou_smi.c:396-441: 1 ready, 0 timeout, -1 error
int smi_poll(int fd, int timeout_ms, short events)
{
    struct pollfd fds;
    fds.fd = fd; fds.events = events; fds.revents = 0;
    int pr;
    do { pr = poll(&fds, 1, timeout_ms); } while (pr == -1 && (errno == EINTR || errno == EAGAIN));
    if (pr < 0) return -1;
    if (pr == 0) return 0;
    return ((fds.revents & POLLIN) || (fds.revents & POLLOUT)) ? 1 : 0;
}

// caribou_smi_timeout_read, caribou_smi.c:466-492
int timeout_read(int fd, uint8_t* buf, size_t len, int timeout_ms)
{
    int ret = (int)read(fd, buf, len);
    if (ret <= 0) {
        const int pr = smi_poll(fd, timeout_ms, POLLIN);
        if (pr < 0) return -1;
        if (pr == 0) return 0;
        return (int)read(fd, buf, len);
    }
    return ret;
}

// caribou_smi_timeout_write, caribou_smi.c:443-464
int timeout_write(int fd, const uint8_t* buf, size_t len, int timeout_ms)
{
    const int pr = smi_poll(fd, timeout_ms, POLLOUT);
    if (pr < 0) return -1;
    if (pr == 0) return 0;
    return (int)write(fd, buf, len);
}

caribou_cuda_ctx* smi_ctx(caribou_cuda_smi* d, int channel)
{
    const int ch = channel == CARIBOU_CUDA_CH_HIF ? 1 : 0;
    if (!d->ctx[ch]) {
        caribou_cuda_config cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.device = d->device;
        cfg.channel = ch;
        cfg.chunk_bytes = d->native_batch_len;
        d->ctx[ch] = caribou_cuda_create(&cfg);
    }
    return d->ctx[ch];
}

bool stage_reserve(caribou_cuda_smi* d, size_t n)
{
    if (n <= d->h_cap) return true;
    uint8_t* p = nullptr;
    const size_t want = (n + 65535) & ~size_t(65535);
    if (!cuda_ok(cudaMallocHost(&p, want), "cudaMallocHost")) return false;
    if (d->stage_busy) { cudaEventSynchronize(d->stage_ev); d->stage_busy = false; }
    if (d->h_stage) { memcpy(p, d->h_stage, d->h_cap); cudaFreeHost(d->h_stage); }
    d->h_stage = p; d->h_cap = want;
    return true;
}

// Does candidate offset 0 of a chunk carry four valid headers (caribou_smi.c:248-262, the aligned case)?  A look at 16
// bytes of what read(2) just delivered: it decides where a segment ends, nothing else - the alignment search itself
// runs on the device.
bool chunk_starts_aligned(const uint8_t* p, size_t len)
{
    if (len <= 16) return true;                                // caribou_smi.c:240-243
    for (int k = 0; k < 4; k++) {
        uint32_t w;
        memcpy(&w, p + 4 * k, 4);
        if (!hdr_ok(w)) return false;
    }
    return true;
}

// The reference's read loop (caribou_smi.c:632-682).  read(2) results are gathered into segments
// (a run of full chunks closed by at most one short chunk) and each segment is handed to `flush`
// (segment bytes, byte count, index of its first output sample), which returns 0 or an error code.
// *got receives the samples handed to flush so far, also when the loop ends in an error: like the reference,
// which has decoded every chunk it read before the failing read(2) (caribou_smi.c:653-664).
// The reference stops reading at the first chunk without alignment (caribou_smi.c:665-668), so the descriptor must
// not be read past such a chunk: a chunk that does not start aligned closes its segment as well, and the device's
// verdict on it is waited for before the next read(2) - `flush` returns it (CARIBOU_CUDA_ERR_SYNC) or, where the
// flush only queues work, `lost_now` synchronises and tells.  Aligned captures never take that turn.
int read_loop(caribou_cuda_smi* d, size_t length_samples,
              const std::function<int(const uint8_t*, size_t, size_t)>& flush, size_t* got,
              const std::function<bool()>& lost_now = std::function<bool()>())
{
    size_t left = length_samples * CARIBOU_CUDA_BYTES_PER_SAMPLE;
    size_t read_so_far = 0, seg_bytes = 0, seg_first = 0;
    const size_t nbl = d->native_batch_len;
    if (got) *got = 0;
    while (left) {
        const size_t cur = left > nbl ? nbl : left;
        if (!stage_reserve(d, seg_bytes + cur + 64)) return CARIBOU_CUDA_ERR;
        // a new segment reuses the buffer an asynchronous upload may still be reading
        if (seg_bytes == 0 && d->stage_busy) { cudaEventSynchronize(d->stage_ev); d->stage_busy = false; }
        const int to = calc_read_timeout(d->sample_rate, cur);
        const int ret = timeout_read(d->fd, d->h_stage + seg_bytes, cur, to);
        if (ret < 0) {                                         // I/O error: what was read before it is still decoded
            if (seg_bytes) { const int rc = flush(d->h_stage, seg_bytes, seg_first); if (rc < 0) return rc; if (got) *got = read_so_far; }
            return CARIBOU_CUDA_ERR;
        }
        if (ret == 0) break;                                   // timeout: deliver what we have
        const bool aligned0 = chunk_starts_aligned(d->h_stage + seg_bytes, (size_t)ret);
        seg_bytes += (size_t)ret;
        read_so_far += (size_t)ret / CARIBOU_CUDA_BYTES_PER_SAMPLE;
        left -= (size_t)ret;
        if ((size_t)ret != nbl || !aligned0) {                 // a short chunk closes the segment; so does a suspect one
            const int rc = flush(d->h_stage, seg_bytes, seg_first);
            if (rc < 0) return rc;
            if (got) *got = read_so_far;
            seg_bytes = 0; seg_first = read_so_far;
            if (!aligned0 && lost_now && lost_now()) break;    // no alignment in that chunk: nothing is read after it
        }
    }
    if (seg_bytes) {
        const int rc = flush(d->h_stage, seg_bytes, seg_first);
        if (rc < 0) return rc;
        if (got) *got = read_so_far;
    }
    return (int)read_so_far;
}

int32_t* lost_flag(caribou_cuda_ctx* c) { return c->d_sflags + 3 * (size_t)c->cfg.n_streams; }
bool lost_seen(caribou_cuda_ctx* c) { return c->h_sflags[3 * (size_t)c->cfg.n_streams] != 0; }

// segment -> device CS16 (+meta) at sample offset `first`, queued on the context's stream and not waited for
int unpack_segment_to_device(caribou_cuda_smi* d, caribou_cuda_ctx* c, const uint8_t* h, size_t nbytes, uint32_t* d_cs16,
                             uint8_t* d_meta, size_t first)
{
    // vector stores need 16-byte aligned slots; a segment that starts at an odd sample offset falls back to scalar stores
    caribou_cuda_unpack_out o;
    memset(&o, 0, sizeof(o));
    o.cs16 = d_cs16 ? d_cs16 + first : nullptr;
    o.meta = d_meta ? d_meta + first : nullptr;
    if (!c->stage_in.reserve(nbytes + 64)) return CARIBOU_CUDA_ERR;
    if (!cuda_ok(cudaMemcpyAsync(c->stage_in.p, h, nbytes, cudaMemcpyHostToDevice, c->stream), "H2D")) return CARIBOU_CUDA_ERR;
    if (!cuda_ok(cudaEventRecord(d->stage_ev, c->stream), "cudaEventRecord")) return CARIBOU_CUDA_ERR;
    d->stage_busy = true;
    if (caribou_cuda_unpack_submit(c, c->stage_in.p, nbytes, 0, CARIBOU_CUDA_MEM_DEVICE, &o) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    return 0;
}

// read + unpack of `n` samples into device buffers, everything queued on the channel context's stream; the
// sticky lost-sync flag of that context says afterwards (mirror_counters + synchronise) whether a chunk had no
// alignment.  Returns what read_loop returns; *got = samples whose decode has been queued.
int read_to_device(caribou_cuda_smi* d, caribou_cuda_ctx* c, uint32_t* d_cs16, uint8_t* d_meta, size_t n, size_t* got)
{
    if (!cuda_ok(cudaMemsetAsync(lost_flag(c), 0, 4, c->stream), "cudaMemsetAsync")) return CARIBOU_CUDA_ERR;
    auto flush = [&](const uint8_t* h, size_t nbytes, size_t first) -> int {
        return unpack_segment_to_device(d, c, h, nbytes, d_cs16, d_meta, first);
    };
    auto lost_now = [&]() -> bool {                            // (only after a chunk that did not start aligned)
        if (mirror_counters(c) != CARIBOU_CUDA_OK || !cuda_ok(cudaStreamSynchronize(c->stream), "cudaStreamSynchronize")) return true;
        d->stage_busy = false;
        return lost_seen(c);
    };
    return read_loop(d, n, flush, got, lost_now);
}

}  // namespace

// ------------------------------------------------------------------ format conversion kernel
// LC/soapy_api/CaribouliteStream.cpp:315-321 (CF32), :337-343 (CF64), :360-364 (CS8); LC/CaribouLiteRadioCpp.cpp:91 (CF32)
__global__ void convert_cs16_kernel(const uint32_t* __restrict__ in, void* __restrict__ out, size_t n, int fmt)
{
    for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (size_t)gridDim.x * blockDim.x) {
        const uint32_t w = in[k];
        const int i = (int)(int16_t)(w & 0xFFFFu), q = (int32_t)w >> 16;
        if (fmt == CARIBOU_CUDA_FMT_CF32)
            reinterpret_cast<float2*>(out)[k] = make_float2((float)i / 4096.0f, (float)q / 4096.0f);
        else if (fmt == CARIBOU_CUDA_FMT_CF64)
            reinterpret_cast<double2*>(out)[k] = make_double2((double)i / 4096.0, (double)q / 4096.0);
        else
            reinterpret_cast<uint16_t*>(out)[k] = (uint16_t)(((uint32_t)(i >> 5) & 0xFFu) | (((uint32_t)(q >> 5) & 0xFFu) << 8));
    }
}
static bool launch_convert(caribou_cuda_ctx* c, const uint32_t* in, void* out, size_t n, int fmt)
{
    if (n == 0) return true;
    const unsigned grid = (unsigned)std::min<size_t>((n + 255) / 256, (size_t)c->num_sms * 8);
    convert_cs16_kernel<<<grid, 256, 0, c->stream>>>(in, out, n, fmt);
    c->launches++;
    return cuda_ok(cudaGetLastError(), "convert_cs16_kernel");
}

struct caribou_cuda_rxstream {
    caribou_cuda_smi* dev = nullptr;
    int channel = 0;
    int format = CARIBOU_CUDA_FMT_CS16;      // CaribouliteStream.cpp:77
    int filter_type = 0;                     // DigitalFilter_None
    size_t mtu = 0;
    caribou_cuda_ctx* filt[4] = {nullptr, nullptr, nullptr, nullptr};   // [1..3]: 20k/50k/100k, never reset
    DevBuf interm;                           // interm_native_buffer2: MTU+1 CS16, persistent
    DevBuf user;                             // device image of the caller's CS16 buffer
    DevBuf conv;                             // converted output
    HostBuf bounce;                          // page-locked: the caller's CS16 buffer on its way up, the result on its way down
};

extern "C" {

caribou_cuda_smi* caribou_cuda_smi_open(int fd, size_t native_batch_len, int device)
{
    caribou_cuda_smi* d = new (std::nothrow) caribou_cuda_smi();
    if (!d) { set_error("out of memory"); return nullptr; }
    d->fd = fd;
    d->native_batch_len = native_batch_len ? native_batch_len : CARIBOU_CUDA_NATIVE_BATCH_LEN;
    d->device = device;
    if (!smi_ctx(d, CARIBOU_CUDA_CH_S1G)) { delete d; return nullptr; }     // fails loudly without a B200
    if (!cuda_ok(cudaEventCreateWithFlags(&d->stage_ev, cudaEventDisableTiming), "cudaEventCreate")) { caribou_cuda_smi_close(d); return nullptr; }
    return d;
}

void caribou_cuda_smi_close(caribou_cuda_smi* d)
{
    if (!d) return;
    cudaSetDevice(d->device);
    for (int i = 0; i < 2; i++) if (d->ctx[i]) caribou_cuda_destroy(d->ctx[i]);
    if (d->stage_ev) cudaEventDestroy(d->stage_ev);
    if (d->h_stage) cudaFreeHost(d->h_stage);
    d->radio_cs16.release(); d->radio_meta.release(); d->radio_cf32.release(); d->radio_bounce.release();
    delete d;
}

size_t caribou_cuda_smi_get_native_batch_samples(caribou_cuda_smi* d)
{
    return d ? d->native_batch_len / CARIBOU_CUDA_BYTES_PER_SAMPLE : 0;
}

int caribou_cuda_smi_read(caribou_cuda_smi* d, int channel, caribou_cuda_sample_complex_int16* buffer,
                          caribou_cuda_sample_meta* metadata, size_t length_samples)
{
    if (!d) { set_error("null device"); return CARIBOU_CUDA_ERR; }
    caribou_cuda_ctx* c = smi_ctx(d, channel);
    if (!c) return CARIBOU_CUDA_ERR;
    // the segment sits in the page-locked stage, the caller's arrays are pageable: one synchronisation per segment
    // (api.cu, pageable-output path), after which the stage is free again
    auto flush = [&](const uint8_t* h, size_t nbytes, size_t first) -> int {
        caribou_cuda_unpack_out o;
        memset(&o, 0, sizeof(o));
        o.cs16 = buffer ? (void*)(buffer + first) : nullptr;
        o.meta = metadata ? (void*)(metadata + first) : nullptr;
        if (caribou_cuda_unpack_submit(c, h, nbytes, 0, CARIBOU_CUDA_MEM_IO(CARIBOU_CUDA_MEM_HOST_PINNED, CARIBOU_CUDA_MEM_HOST), &o) != CARIBOU_CUDA_OK)
            return CARIBOU_CUDA_ERR;
        const int64_t r = caribou_cuda_wait(c, nullptr);
        if (r == CARIBOU_CUDA_ERR_SYNC) return CARIBOU_CUDA_ERR_SYNC;
        return r < 0 ? CARIBOU_CUDA_ERR : 0;
    };
    return read_loop(d, length_samples, flush, nullptr);
}

// CaribouLiteRadio::ReadSamples(std::complex<float>*, size_t, uint8_t*) - LC/CaribouLite.hpp:188-189,
// LC/CaribouLiteRadioCpp.cpp:72-95 with :98-130 underneath: the native read goes into the radio's own persistent
// _read_samples / _read_metadata buffers, then samples[i] = _read_samples[i] / 4096.0f and the metadata is copied
// out, `ret` of each.  Here those buffers live on the device (slots a misaligned chunk leaves unwritten keep the
// previous read's values, as they do in the reference), the conversion is a kernel, and CF32 + meta come back
// in one synchronisation.  Returns what cariboulite_radio_read_samples returns (<= 0: nothing is written).
int caribou_cuda_radio_read_cf32(caribou_cuda_smi* d, int channel, float* samples, uint8_t* meta, size_t num_to_read)
{
    if (!d) { set_error("null device"); return CARIBOU_CUDA_ERR; }
    if (!samples || num_to_read == 0) return 0;                              // CaribouLiteRadioCpp.cpp:74-78,100-105
    caribou_cuda_ctx* c = smi_ctx(d, channel);
    if (!c) return CARIBOU_CUDA_ERR;
    cudaSetDevice(d->device);
    const size_t n = num_to_read;
    if (n * 4 + 64 > d->radio_cs16.cap) {
        // first use or a longer read than before: fresh zeroed buffers, like the radio's calloc'ed ones
        if (!d->radio_cs16.reserve(n * 4 + 64) || !d->radio_meta.reserve(n + 64) || !d->radio_cf32.reserve(n * 8 + 64)) return CARIBOU_CUDA_ERR;
        cudaMemsetAsync(d->radio_cs16.p, 0, d->radio_cs16.cap, c->stream);
        cudaMemsetAsync(d->radio_meta.p, 0, d->radio_meta.cap, c->stream);
    }
    const size_t off_meta = (n * 8 + 63) & ~size_t(63);
    if (!d->radio_bounce.reserve(off_meta + n + 64)) return CARIBOU_CUDA_ERR;
    size_t got = 0;
    const int res = read_to_device(d, c, d->radio_cs16.as<uint32_t>(), d->radio_meta.as<uint8_t>(), n, &got);
    if (res > 0) {
        if (!launch_convert(c, d->radio_cs16.as<uint32_t>(), d->radio_cf32.p, (size_t)res, CARIBOU_CUDA_FMT_CF32)) return CARIBOU_CUDA_ERR;
        uint8_t* hb = d->radio_bounce.as<uint8_t>();
        if (!cuda_ok(cudaMemcpyAsync(hb, d->radio_cf32.p, (size_t)res * 8, cudaMemcpyDeviceToHost, c->stream), "D2H")) return CARIBOU_CUDA_ERR;
        if (meta && !cuda_ok(cudaMemcpyAsync(hb + off_meta, d->radio_meta.p, (size_t)res, cudaMemcpyDeviceToHost, c->stream), "D2H")) return CARIBOU_CUDA_ERR;
    }
    if (mirror_counters(c) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (!cuda_ok(cudaStreamSynchronize(c->stream), "cudaStreamSynchronize")) return CARIBOU_CUDA_ERR;
    d->stage_busy = false;
    if (res <= 0) return res;
    if (lost_seen(c)) return CARIBOU_CUDA_ERR_SYNC;
    memcpy(samples, d->radio_bounce.p, (size_t)res * 8);
    if (meta) memcpy(meta, d->radio_bounce.as<uint8_t>() + off_meta, (size_t)res);
    return res;
}

// ------------------------------------------------------------------ Stream
caribou_cuda_rxstream* caribou_cuda_rxstream_create(caribou_cuda_smi* dev, int channel)
{
    if (!dev) { set_error("null device"); return nullptr; }
    caribou_cuda_ctx* c = smi_ctx(dev, channel);
    if (!c) return nullptr;
    caribou_cuda_rxstream* s = new (std::nothrow) caribou_cuda_rxstream();
    if (!s) { set_error("out of memory"); return nullptr; }
    s->dev = dev;
    s->channel = channel;
    s->mtu = caribou_cuda_smi_get_native_batch_samples(dev);           // CaribouliteStream.cpp:66,118-121
    // filt20/50/100: LowPass<6>.setup(4e6, {20e3,50e3,100e3}/2), CaribouliteStream.cpp:85-91.  The filter contexts
    // launch on the channel context's stream: read -> unpack -> filter -> convert -> copy is one queue, nothing waits
    // in between.
    const double fc[4] = {0, 20e3 / 2, 50e3 / 2, 100e3 / 2};
    for (int t = 1; t <= 3; t++) {
        caribou_cuda_config cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.device = dev->device;
        cfg.n_sos = caribou_cuda_butter_lowpass(6, 4e6, fc[t], &cfg.sos[0][0]);
        cfg.decimation = 1;
        cfg.parity = CARIBOU_CUDA_PARITY_Q16;
        cfg.out_format = CARIBOU_CUDA_FMT_CS16;       // filtered in place as int16, CaribouliteStream.cpp:291-298
        cfg.cuda_stream = (void*)c->stream;
        s->filt[t] = caribou_cuda_create(&cfg);
        if (!s->filt[t]) { caribou_cuda_rxstream_destroy(s); return nullptr; }
    }
    cudaSetDevice(dev->device);
    if (!s->interm.reserve((s->mtu + 4) * 4)) { caribou_cuda_rxstream_destroy(s); return nullptr; }
    cudaMemset(s->interm.p, 0, (s->mtu + 4) * 4);
    return s;
}

void caribou_cuda_rxstream_destroy(caribou_cuda_rxstream* s)
{
    if (!s) return;
    for (int t = 0; t < 4; t++) if (s->filt[t]) caribou_cuda_destroy(s->filt[t]);
    s->interm.release(); s->user.release(); s->conv.release(); s->bounce.release();
    delete s;
}

int caribou_cuda_rxstream_set_format(caribou_cuda_rxstream* s, const char* fmt)
{
    if (!s || !fmt) return -1;                                             // CaribouliteStream.cpp:158-173
    if (!strcmp(fmt, "CS16")) s->format = CARIBOU_CUDA_FMT_CS16;
    else if (!strcmp(fmt, "CS8")) s->format = CARIBOU_CUDA_FMT_CS8;
    else if (!strcmp(fmt, "CF32")) s->format = CARIBOU_CUDA_FMT_CF32;
    else if (!strcmp(fmt, "CF64")) s->format = CARIBOU_CUDA_FMT_CF64;
    else return -1;
    return 0;
}

void caribou_cuda_rxstream_set_filter(caribou_cuda_rxstream* s, int type)
{
    if (!s) return;
    s->filter_type = (type >= 1 && type <= 3) ? type : 0;                  // CaribouliteStream.cpp:127-141 (200 kHz has no case)
}

size_t caribou_cuda_rxstream_mtu(caribou_cuda_rxstream* s) { return s ? s->mtu : 0; }

// Stream::ReadSamples(cs16*) on a device buffer: read + unpack into d_buf, then the in-place int16 filter - queued,
// not waited for.  The filter is gated on the channel context's lost-sync flag: a read that lost word sync returns 0
// elements in the reference and never reaches the filter loop (CaribouliteStream.cpp:266-276,291-298).
static int stream_read_cs16_device(caribou_cuda_rxstream* s, caribou_cuda_ctx* c, uint32_t* d_buf, size_t n, size_t* got)
{
    int res = read_to_device(s->dev, c, d_buf, nullptr, n, got);
    if (res > 0 && s->filter_type != 0) {
        if (filter_device(s->filt[s->filter_type], CARIBOU_CUDA_IN_CS16, d_buf, (size_t)res, d_buf, lost_flag(c)) != CARIBOU_CUDA_OK)
            return CARIBOU_CUDA_ERR;
    }
    return res;
}

int caribou_cuda_rxstream_read(caribou_cuda_rxstream* s, void* buffer, size_t num_elements, long timeout_us)
{
    (void)timeout_us;                                                      // unused by the reference as well
    if (!s || !buffer) { set_error("bad argument"); return 0; }
    caribou_cuda_smi* d = s->dev;
    caribou_cuda_ctx* c = smi_ctx(d, s->channel);
    if (!c) return 0;
    cudaSetDevice(d->device);
    auto finish = [&]() {      // one synchronisation per call; afterwards the stage and the bounce buffer are free
        const bool ok = mirror_counters(c) == CARIBOU_CUDA_OK && cuda_ok(cudaStreamSynchronize(c->stream), "cudaStreamSynchronize");
        d->stage_busy = false;
        return ok;
    };
    if (s->format == CARIBOU_CUDA_FMT_CS16) {
        // no MTU clamp for CS16 (CaribouliteStream.cpp:282-301).  The read decodes straight into the caller's buffer
        // in the reference, so slots a misaligned chunk leaves untouched keep the caller's bytes - and the in-place
        // filter then runs over them: the device works on an image of the caller's buffer.
        const size_t nb = num_elements * 4;
        if (!s->user.reserve(nb + 64) || !s->bounce.reserve(nb + 64)) return 0;
        memcpy(s->bounce.p, buffer, nb);
        if (!cuda_ok(cudaMemcpyAsync(s->user.p, s->bounce.p, nb, cudaMemcpyHostToDevice, c->stream), "H2D")) return 0;
        size_t got = 0;
        const int res = stream_read_cs16_device(s, c, s->user.as<uint32_t>(), num_elements, &got);
        if (got && !cuda_ok(cudaMemcpyAsync(s->bounce.p, s->user.p, got * 4, cudaMemcpyDeviceToHost, c->stream), "D2H")) return 0;
        if (!finish()) return 0;
        // chunks decoded before an error or a lost sync are in the caller's buffer in the reference as well
        if (got) memcpy(buffer, s->bounce.p, got * 4);
        return (res < 0 || lost_seen(c)) ? 0 : res;                        // CaribouliteStream.cpp:266-276
    }
    if (num_elements > s->mtu) num_elements = s->mtu;                     // CaribouliteStream.cpp:306,328,351
    size_t got = 0;
    const int res = stream_read_cs16_device(s, c, s->interm.as<uint32_t>(), num_elements, &got);
    const size_t es = s->format == CARIBOU_CUDA_FMT_CF32 ? 8 : (s->format == CARIBOU_CUDA_FMT_CF64 ? 16 : 2);
    if (res > 0) {
        if (!s->conv.reserve((size_t)res * es + 64) || !s->bounce.reserve((size_t)res * es + 64)) return 0;
        if (!launch_convert(c, s->interm.as<uint32_t>(), s->conv.p, (size_t)res, s->format)) return 0;
        if (!cuda_ok(cudaMemcpyAsync(s->bounce.p, s->conv.p, (size_t)res * es, cudaMemcpyDeviceToHost, c->stream), "D2H")) return 0;
    }
    if (!finish()) return 0;
    if (res <= 0 || lost_seen(c)) return 0;
    memcpy(buffer, s->bounce.p, (size_t)res * es);
    return res;
}

// ------------------------------------------------------------------ capture pump
namespace {
// read until `want` bytes, EOF (returns what it has) or error (-1)
ssize_t read_full(int fd, uint8_t* buf, size_t want)
{
    size_t got = 0;
    while (got < want) {
        const ssize_t r = read(fd, buf + got, want - got);
        if (r < 0) { if (errno == EINTR) continue; return -1; }
        if (r == 0) break;
        got += (size_t)r;
    }
    return (ssize_t)got;
}
bool write_full(int fd, const uint8_t* buf, size_t n)
{
    size_t put = 0;
    while (put < n) {
        const ssize_t r = write(fd, buf + put, n - put);
        if (r < 0) { if (errno == EINTR) continue; return false; }
        put += (size_t)r;
    }
    return true;
}
}  // namespace

int64_t caribou_cuda_pump(caribou_cuda_ctx* c, int in_kind, int in_fd, int out_fd, size_t block_samples, int64_t max_samples)
{
    if (!c || in_kind < 0 || in_kind > 2 || block_samples == 0) { set_error("bad argument"); return CARIBOU_CUDA_ERR; }
    if (c->cfg.n_streams != 1 || c->cfg.interleave_block_words) { set_error("pump: single-stream, single-channel contexts only"); return CARIBOU_CUDA_ERR; }
    if (!c->has_filter && in_kind != CARIBOU_CUDA_IN_WORDS) { set_error("pump: a context without filter takes raw SMI words"); return CARIBOU_CUDA_ERR; }
    const size_t D = c->has_filter ? (size_t)c->plan.D : 1;
    if (block_samples % D) { set_error("pump: block_samples must be a multiple of the decimation factor"); return CARIBOU_CUDA_ERR; }
    const size_t in_es = in_kind == CARIBOU_CUDA_IN_CF32 ? 8 : 4;
    const size_t out_es = c->has_filter ? c->out_elem : 4;
    uint8_t* hin[2] = {nullptr, nullptr};
    uint8_t* hout[2] = {nullptr, nullptr};
    auto cleanup = [&]() { for (int i = 0; i < 2; i++) { caribou_cuda_host_free(hin[i]); caribou_cuda_host_free(hout[i]); } };
    for (int i = 0; i < 2; i++) {
        hin[i] = (uint8_t*)caribou_cuda_host_alloc(block_samples * in_es + 64);
        hout[i] = (uint8_t*)caribou_cuda_host_alloc(block_samples / D * out_es + 64);
        if (!hin[i] || !hout[i]) { cleanup(); return CARIBOU_CUDA_ERR; }
    }
    int64_t consumed = 0, rc = 0;
    size_t prev_out_bytes = 0;      // output of the block before the one in flight, not yet written
    int prev_slot = -1;
    // block 0 is read up front; afterwards block k+1 is read while block k is on the device
    auto want_bytes = [&](void) -> size_t {
        size_t w = block_samples;
        if (max_samples > 0) w = (size_t)std::min<int64_t>((int64_t)w, max_samples - consumed);
        return w * in_es;
    };
    ssize_t have = read_full(in_fd, hin[0], want_bytes());
    int slot = 0;
    while (have > 0) {
        size_t n = (size_t)have / in_es;
        n -= n % D;                                     // a tail shorter than D is not processed
        if (n == 0) break;
        // ---- block `slot` goes to the device (asynchronous for pinned memory)
        int sr;
        if (!c->has_filter) {
            caribou_cuda_unpack_out o;
            memset(&o, 0, sizeof(o));
            o.cs16 = hout[slot];
            sr = caribou_cuda_unpack_submit(c, hin[slot], n * 4, 0, CARIBOU_CUDA_MEM_HOST_PINNED, &o);
        } else if (in_kind == CARIBOU_CUDA_IN_WORDS) {
            sr = caribou_cuda_rx_submit(c, hin[slot], n * 4, 0, CARIBOU_CUDA_MEM_HOST_PINNED, nullptr, hout[slot], 0);
        } else {
            sr = caribou_cuda_filter_submit(c, in_kind, hin[slot], n, 0, CARIBOU_CUDA_MEM_HOST_PINNED, hout[slot], 0);
        }
        if (sr != CARIBOU_CUDA_OK) { rc = CARIBOU_CUDA_ERR; break; }
        consumed += (int64_t)n;
        // ---- meanwhile on the host: write the previous block's result, read the next block
        if (prev_slot >= 0 && !write_full(out_fd, hout[prev_slot], prev_out_bytes)) { set_error("pump: write failed (%s)", strerror(errno)); rc = CARIBOU_CUDA_ERR; prev_slot = -1; break; }
        prev_slot = -1;
        ssize_t next_have = 0;
        if (!(max_samples > 0 && consumed >= max_samples) && (size_t)have == block_samples * in_es)
            next_have = read_full(in_fd, hin[slot ^ 1], want_bytes());
        // ---- the block in flight
        const int64_t wr = caribou_cuda_wait(c, nullptr);
        if (wr < 0) { rc = wr; break; }
        prev_slot = slot;
        prev_out_bytes = (size_t)(c->has_filter ? c->last_n_filtered : wr) * out_es;
        if (next_have < 0) { set_error("pump: read failed (%s)", strerror(errno)); rc = CARIBOU_CUDA_ERR; break; }
        have = next_have;
        slot ^= 1;
    }
    if (have < 0 && rc == 0) { set_error("pump: read failed (%s)", strerror(errno)); rc = CARIBOU_CUDA_ERR; }
    if (prev_slot >= 0 && rc >= 0 && !write_full(out_fd, hout[prev_slot], prev_out_bytes)) { set_error("pump: write failed (%s)", strerror(errno)); rc = CARIBOU_CUDA_ERR; }
    cleanup();
    return rc < 0 ? rc : consumed;
}

// ------------------------------------------------------------------ TX shims
// caribou_smi_write (caribou_smi.c:720-762): pack on the GPU, then hand native_batch_len bytes at a time to the fd
// exactly as the reference's loop does.  `format` selects the Stream::WriteSamples conversion that runs first
// (CaribouliteStream.cpp:199-244); CS16 is caribou_smi_write itself.
static int tx_write(caribou_cuda_smi* d, int channel, int format, const void* buffer, size_t n)
{
    if (!d) { set_error("null device"); return CARIBOU_CUDA_ERR; }
    caribou_cuda_ctx* c = smi_ctx(d, channel);
    if (!c) return CARIBOU_CUDA_ERR;
    if (n == 0) return 0;
    if (!buffer) { set_error("null buffer"); return CARIBOU_CUDA_ERR; }
    if (!stage_reserve(d, n * 4)) return CARIBOU_CUDA_ERR;
    if (caribou_cuda_tx_pack_submit(c, format, buffer, n, CARIBOU_CUDA_MEM_HOST, d->h_stage) != CARIBOU_CUDA_OK) return CARIBOU_CUDA_ERR;
    if (caribou_cuda_wait(c, nullptr) < 0) return CARIBOU_CUDA_ERR;
    size_t left = n * CARIBOU_CUDA_BYTES_PER_SAMPLE, written = 0;           // bytes, samples (caribou_smi.c:723-724)
    uint32_t to_ms = (uint32_t)((2 * n * 1000) / d->sample_rate);          // :725-726
    if (to_ms < 2) to_ms = 2;
    while (left) {
        size_t cur = left > d->native_batch_len ? d->native_batch_len : left;
        cur &= ~size_t(3);                                                  // :745
        if (!cur || (written * 4 + cur) > n * 4) break;                     // (the second test: a short write(2) makes the
                                                                            // reference's sample_offset run ahead of `left`)
        const int ret = timeout_write(d->fd, d->h_stage + written * 4, cur, (int)to_ms);   // :750
        if (ret < 0) { set_error("smi write failed (%s)", strerror(errno)); return CARIBOU_CUDA_ERR; }
        if (ret == 0) break;                                                // POLLOUT timeout
        written += cur / CARIBOU_CUDA_BYTES_PER_SAMPLE;                     // :757-758: samples advance by the chunk,
        left -= (size_t)ret > left ? left : (size_t)ret;                    //           bytes by what write(2) took
    }
    return (int)written;
}

int caribou_cuda_smi_write(caribou_cuda_smi* d, int channel, const caribou_cuda_sample_complex_int16* samples,
                           size_t length_samples)
{
    return tx_write(d, channel, CARIBOU_CUDA_FMT_CS16, samples, length_samples);
}

// Stream::WriteSamplesGen (CaribouliteStream.cpp:246-258): formats other than CS16 are clamped to the MTU (:199,214,229),
// negative results of the CS16 write become 0 (:183-194).
int caribou_cuda_txstream_write(caribou_cuda_smi* d, int channel, const char* fmt, const void* buffer, size_t num_elements)
{
    if (!d || !fmt) { set_error("bad argument"); return 0; }
    int format = CARIBOU_CUDA_FMT_CS16;
    if (!strcmp(fmt, "CS16")) format = CARIBOU_CUDA_FMT_CS16;
    else if (!strcmp(fmt, "CS8")) format = CARIBOU_CUDA_FMT_CS8;
    else if (!strcmp(fmt, "CF32")) format = CARIBOU_CUDA_FMT_CF32;
    else if (!strcmp(fmt, "CF64")) format = CARIBOU_CUDA_FMT_CF64;
    else { set_error("unknown format %s", fmt); return 0; }
    const size_t mtu = caribou_cuda_smi_get_native_batch_samples(d);
    if (format != CARIBOU_CUDA_FMT_CS16 && num_elements > mtu) num_elements = mtu;
    const int ret = tx_write(d, channel, format, buffer, num_elements);
    return ret < 0 ? 0 : ret;
}

}  // extern "C"
