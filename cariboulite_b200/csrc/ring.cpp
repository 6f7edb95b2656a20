// Host side of continuous streaming: the sample queue between a reader thread and the consumer,
// and the reader thread itself with the sample path on the GPU.
//
// The queue keeps the behaviour of the reference's circular_buffer<T>(size, override_write, block_read)
// (LC/datatypes/circular_buffer.h:16-92): capacity rounded up to a power of two, free-running head / tail counters,
// put() that drops the OLDEST elements when the writer outruns the reader, get() that waits up to timeout_us for
// `length` elements and takes nothing (returns 0) when they did not arrive.  One difference: a put() longer than
// the capacity overruns the reference's buffer; here it keeps the newest `capacity` elements.
//
// The reader is the reference's ReaderThread (LC/soapy_api/CaribouliteStream.cpp:16-49): while running, read one
// MTU through caribou_smi_read - here caribou_cuda_smi_read, i.e. the GPU path - and put() what came back; a
// negative read counts as nothing.  Sized as the reference sizes it: NUM_NATIVE_MTUS_PER_QUEUE = 10 MTUs
// (CaribouliteStream.cpp:8,71) unless the caller asks otherwise.
#include <string.h>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <new>
#include <thread>
#include <vector>
#include "../../include/caribou_cuda.h"

struct caribou_cuda_ring {
    std::vector<uint8_t> buf;
    size_t elem = 4, cap = 0;              // element bytes, capacity in elements (power of two)
    size_t head = 0, tail = 0;             // free-running element counters
    bool override_write = true, block_read = true;
    uint64_t dropped = 0;                  // elements overwritten before they were read
    std::mutex m;
    std::condition_variable cv;
    size_t size_locked() const { return head - tail; }
};

struct caribou_cuda_async {
    caribou_cuda_smi* dev = nullptr;
    int channel = 0;
    caribou_cuda_ring* ring = nullptr;
    std::thread th;
    std::atomic<int> running{0}, active{0}, idle{0};
    std::atomic<uint64_t> blocks{0};
    size_t mtu = 0;
    std::vector<caribou_cuda_sample_complex_int16> block;
    std::vector<caribou_cuda_sample_meta> meta;
};

extern "C" {

caribou_cuda_ring* caribou_cuda_ring_create(size_t capacity, size_t elem_bytes, int override_write, int block_read)
{
    if (capacity == 0 || elem_bytes == 0) return nullptr;
    caribou_cuda_ring* r = new (std::nothrow) caribou_cuda_ring();
    if (!r) return nullptr;
    size_t cap = 1;
    while (cap < capacity) cap <<= 1;                      // circular_buffer.h:21-24
    r->cap = cap;
    r->elem = elem_bytes;
    r->override_write = override_write != 0;
    r->block_read = block_read != 0;
    r->buf.resize(cap * elem_bytes);
    return r;
}

void caribou_cuda_ring_destroy(caribou_cuda_ring* r) { delete r; }

size_t caribou_cuda_ring_put(caribou_cuda_ring* r, const void* data, size_t length)
{
    if (!r || (!data && length)) return 0;
    const uint8_t* src = static_cast<const uint8_t*>(data);
    {
        std::lock_guard<std::mutex> lock(r->m);
        if (length > r->cap) {                              // keep the newest cap elements (see the header comment)
            const size_t skip = length - r->cap;
            r->dropped += skip;
            src += skip * r->elem;
            length = r->cap;
        }
        const size_t free_now = r->cap - r->size_locked();
        if (free_now < length && r->override_write) {       // :41-45: pop what is needed
            r->tail += length - free_now;
            r->dropped += length - free_now;
        }
        const size_t room = r->cap - r->size_locked();
        const size_t len = length < room ? length : room;   // without override: as much as fits
        const size_t at = r->head & (r->cap - 1);
        const size_t first = len < r->cap - at ? len : r->cap - at;
        memcpy(r->buf.data() + at * r->elem, src, first * r->elem);
        memcpy(r->buf.data(), src + first * r->elem, (len - first) * r->elem);
        r->head += len;
        length = len;
    }
    if (r->block_read) r->cv.notify_one();
    return length;
}

size_t caribou_cuda_ring_get(caribou_cuda_ring* r, void* data, size_t length, long timeout_us)
{
    if (!r) return 0;
    std::unique_lock<std::mutex> lock(r->m);
    if (r->block_read) {                                    // :67-82
        r->cv.wait_for(lock, std::chrono::microseconds(timeout_us < 0 ? 0 : timeout_us),
                       [&]() { return r->size_locked() >= length; });
        if (r->size_locked() < length) return 0;
    }
    const size_t have = r->size_locked();
    const size_t len = length < have ? length : have;
    const size_t at = r->tail & (r->cap - 1);
    const size_t first = len < r->cap - at ? len : r->cap - at;
    if (data) {                                             // NULL: just consume (:87-91)
        uint8_t* dst = static_cast<uint8_t*>(data);
        memcpy(dst, r->buf.data() + at * r->elem, first * r->elem);
        memcpy(dst + first * r->elem, r->buf.data(), (len - first) * r->elem);
    }
    r->tail += len;
    return len;
}

void caribou_cuda_ring_reset(caribou_cuda_ring* r)
{
    if (!r) return;
    std::lock_guard<std::mutex> lock(r->m);
    r->head = r->tail = 0;
}
size_t caribou_cuda_ring_size(caribou_cuda_ring* r)
{
    if (!r) return 0;
    std::lock_guard<std::mutex> lock(r->m);
    return r->size_locked();
}
size_t caribou_cuda_ring_capacity(caribou_cuda_ring* r) { return r ? r->cap : 0; }
uint64_t caribou_cuda_ring_dropped(caribou_cuda_ring* r)
{
    if (!r) return 0;
    std::lock_guard<std::mutex> lock(r->m);
    return r->dropped;
}

// ------------------------------------------------------------------ reader thread
static void reader_main(caribou_cuda_async* q)
{
    while (q->running.load()) {
        if (!q->active.load()) {                            // CaribouliteStream.cpp:23-27
            std::this_thread::sleep_for(std::chrono::milliseconds(10));
            continue;
        }
        int ret = caribou_cuda_smi_read(q->dev, q->channel, q->block.data(), q->meta.data(), q->mtu);
        if (ret < 0) ret = 0;                               // :33-42
        if (ret) {
            caribou_cuda_ring_put(q->ring, q->block.data(), (size_t)ret);
            q->blocks.fetch_add(1);
            q->idle.store(0);
        } else {
            q->idle.store(1);                               // a file or pipe that ran dry: do not spin on it
            std::this_thread::sleep_for(std::chrono::milliseconds(1));
        }
    }
}

caribou_cuda_async* caribou_cuda_async_open(caribou_cuda_smi* dev, int channel, size_t capacity_samples)
{
    if (!dev) return nullptr;
    caribou_cuda_async* q = new (std::nothrow) caribou_cuda_async();
    if (!q) return nullptr;
    q->dev = dev;
    q->channel = channel;
    q->mtu = caribou_cuda_smi_get_native_batch_samples(dev);
    if (capacity_samples == 0) capacity_samples = q->mtu * 10;          // NUM_NATIVE_MTUS_PER_QUEUE
    q->ring = caribou_cuda_ring_create(capacity_samples, sizeof(caribou_cuda_sample_complex_int16), 1, 1);
    if (!q->ring || q->mtu == 0) { caribou_cuda_ring_destroy(q->ring); delete q; return nullptr; }
    q->block.resize(q->mtu + 1);
    q->meta.resize(q->mtu + 1);
    q->running.store(1);
    q->active.store(0);                                     // :94-96: the thread starts, the stream is not active yet
    q->th = std::thread(reader_main, q);
    return q;
}

void caribou_cuda_async_activate(caribou_cuda_async* q, int active) { if (q) q->active.store(active ? 1 : 0); }

size_t caribou_cuda_async_read(caribou_cuda_async* q, caribou_cuda_sample_complex_int16* buffer, size_t num_samples,
                               long timeout_us)
{
    return q ? caribou_cuda_ring_get(q->ring, buffer, num_samples, timeout_us) : 0;    // Stream::Read, :263
}

int caribou_cuda_async_idle(caribou_cuda_async* q) { return q ? q->idle.load() : 1; }
caribou_cuda_ring* caribou_cuda_async_ring(caribou_cuda_async* q) { return q ? q->ring : nullptr; }

void caribou_cuda_async_close(caribou_cuda_async* q)
{
    if (!q) return;
    q->active.store(0);                                     // :108-113
    q->running.store(0);
    if (q->th.joinable()) q->th.join();
    caribou_cuda_ring_destroy(q->ring);
    delete q;
}

}  // extern "C"
