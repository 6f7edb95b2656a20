// oracle/_ref glue - TEST INFRASTRUCTURE ONLY.
//
// Builds a shared library out of the reference's UNMODIFIED sources (compiled
// where they lie under /root/reference by oracle/Makefile) plus this file, which
// contains no algorithm of its own: it only
//   * provides the three thin cariboulite_radio_* entry points the reference's
//     Stream class links against (they forward to caribou_smi_*, exactly like
//     software/libcariboulite/src/cariboulite_radio.c:1258-1315 do),
//   * stubs io_utils_set_gpio_mode (hardware, init-only), and
//   * exposes flat C entry points so tests/bench can drive the reference code on
//     synthetic buffers held in a memfd (read(2) is the only "hardware" the RX
//     path touches: caribou_smi.c:466-492).
//
// Nothing in the product path (cariboulite_b200/, include/) may link or load this.
#define _GNU_SOURCE 1
#include <sys/mman.h>
#include <unistd.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <new>

#include "Cariboulite.hpp"   // reference header: pulls Stream, radio and sys_st types
#include <Iir.h>             // reference's vendored iir1

extern "C" {
#include "caribou_smi/caribou_smi.h"
}

// ---------------------------------------------------------------- hardware stubs
extern "C" void io_utils_set_gpio_mode(int gpio, io_utils_alt_en mode) { (void)gpio; (void)mode; }
// The TX path announces its direction to the kernel driver with an ioctl (caribou_smi.c:30-40) a memfd does not
// implement; inside this library (linked -Bsymbolic-functions) the driver says yes.  The RX path makes no ioctl.
extern "C" int ioctl(int fd, unsigned long req, ...) { (void)fd; (void)req; return 0; }

// ---------------------------------------------------------------- radio shims
// Same forwarding the reference does in cariboulite_radio.c:1258-1315.
extern "C" int cariboulite_radio_read_samples(cariboulite_radio_state_st* radio,
                                              cariboulite_sample_complex_int16* buffer,
                                              cariboulite_sample_meta* metadata, size_t length)
{
    return caribou_smi_read(&radio->sys->smi, (caribou_smi_channel_en)radio->smi_channel_id,
                            (caribou_smi_sample_complex_int16*)buffer,
                            (caribou_smi_sample_meta*)metadata, length);
}
extern "C" int cariboulite_radio_write_samples(cariboulite_radio_state_st* radio,
                                               cariboulite_sample_complex_int16* buffer, size_t length)
{
    (void)radio; (void)buffer; (void)length;
    return -1;  // TX is out of scope
}
extern "C" size_t cariboulite_radio_get_native_mtu_size_samples(cariboulite_radio_state_st* radio)
{
    return caribou_smi_get_native_batch_samples(&radio->sys->smi);
}

// ---------------------------------------------------------------- device handle
struct RefDev {
    sys_st* sys;
    cariboulite_radio_state_st* radio;   // points into sys
    SoapySDR::Stream* stream;
};

extern "C" {

// Open a fake SMI device whose fd is a memfd pre-loaded with `nbytes` of raw SMI bytes.
void* ref_dev_open(const uint8_t* bytes, size_t nbytes, size_t native_batch_len, int channel /*0=S1G,1=HiF*/)
{
    RefDev* d = new (std::nothrow) RefDev();
    if (!d) return nullptr;
    d->sys = (sys_st*)calloc(1, sizeof(sys_st));
    int fd = memfd_create("ref_smi", 0);
    if (fd < 0) { free(d->sys); delete d; return nullptr; }
    size_t done = 0;
    while (done < nbytes) {
        ssize_t w = write(fd, bytes + done, nbytes - done);
        if (w <= 0) { close(fd); free(d->sys); delete d; return nullptr; }
        done += (size_t)w;
    }
    lseek(fd, 0, SEEK_SET);
    caribou_smi_st* smi = &d->sys->smi;
    smi->initialized = 1;
    smi->filedesc = fd;
    smi->native_batch_len = native_batch_len;
    smi->sample_rate = CARIBOU_SMI_SAMPLE_RATE;
    smi->debug_mode = caribou_smi_none;
    smi->read_temp_buffer = (uint8_t*)malloc(native_batch_len + 1024);   // caribou_smi.c:564
    smi->write_temp_buffer = nullptr;
    d->radio = &d->sys->radio_low;
    d->radio->sys = d->sys;
    d->radio->smi_channel_id = channel ? caribou_smi_channel_2400 : caribou_smi_channel_900;
    d->stream = nullptr;
    return d;
}

void ref_dev_close(void* h)
{
    RefDev* d = (RefDev*)h;
    if (!d) return;
    if (d->stream) delete d->stream;
    close(d->sys->smi.filedesc);
    free(d->sys->smi.read_temp_buffer);
    free(d->sys->smi.write_temp_buffer);
    free(d->sys);
    delete d;
}

// Re-arm the memfd so the same capture can be read again (bench repetitions).
void ref_dev_rewind(void* h) { lseek(((RefDev*)h)->sys->smi.filedesc, 0, SEEK_SET); }

// caribou_smi_read (caribou_smi.c:632-682), unmodified, on the memfd.
int ref_dev_smi_read(void* h, int16_t* cs16, uint8_t* meta, size_t length_samples)
{
    RefDev* d = (RefDev*)h;
    return caribou_smi_read(&d->sys->smi, (caribou_smi_channel_en)d->radio->smi_channel_id,
                            (caribou_smi_sample_complex_int16*)cs16, (caribou_smi_sample_meta*)meta,
                            length_samples);
}

// caribou_smi_write (caribou_smi.c:720-762), unmodified: the memfd receives what the reference would hand to the
// driver.  write_temp_buffer is what caribou_smi_init allocates (caribou_smi.c:565-571).  The fd is rewound
// first; returns the reference's return value, the bytes it wrote are read back with ref_dev_read_back.
int ref_dev_smi_write(void* h, const int16_t* cs16, size_t length_samples)
{
    RefDev* d = (RefDev*)h;
    caribou_smi_st* smi = &d->sys->smi;
    if (!smi->write_temp_buffer) smi->write_temp_buffer = (uint8_t*)malloc(smi->native_batch_len + 1024);
    if (ftruncate(smi->filedesc, 0) != 0) return -100;
    lseek(smi->filedesc, 0, SEEK_SET);
    return caribou_smi_write(smi, (caribou_smi_channel_en)d->radio->smi_channel_id,
                             (caribou_smi_sample_complex_int16*)cs16, length_samples);
}
long ref_dev_read_back(void* h, uint8_t* out, size_t cap)
{
    RefDev* d = (RefDev*)h;
    const int fd = d->sys->smi.filedesc;
    const off_t size = lseek(fd, 0, SEEK_END);
    lseek(fd, 0, SEEK_SET);
    size_t done = 0;
    while (done < cap && done < (size_t)size) {
        ssize_t r = read(fd, out + done, ((size_t)size < cap ? (size_t)size : cap) - done);
        if (r <= 0) break;
        done += (size_t)r;
    }
    return (long)done;
}

// The read(2) loop of caribou_smi_read alone (same chunking into the device's temp buffer, no decode): what the
// memfd copy costs, so that the CPU baseline can be quoted with and without it (BASELINE.md section 3 run 1).
long ref_dev_read_only(void* h, size_t length_samples)
{
    RefDev* d = (RefDev*)h;
    caribou_smi_st* smi = &d->sys->smi;
    size_t left = length_samples * CARIBOU_SMI_BYTES_PER_SAMPLE, done = 0;
    while (left) {
        const size_t cur = left > smi->native_batch_len ? smi->native_batch_len : left;
        const ssize_t r = read(smi->filedesc, smi->read_temp_buffer, cur);
        if (r <= 0) break;
        done += (size_t)r; left -= (size_t)r;
    }
    return (long)(done / CARIBOU_SMI_BYTES_PER_SAMPLE);
}

// cariboulite_radio_read_samples-shaped call (through the shim above).
int ref_dev_radio_read(void* h, int16_t* cs16, uint8_t* meta, size_t length_samples)
{
    RefDev* d = (RefDev*)h;
    return cariboulite_radio_read_samples(d->radio, (cariboulite_sample_complex_int16*)cs16,
                                          (cariboulite_sample_meta*)meta, length_samples);
}

// SoapySDR::Stream (CaribouliteStream.cpp), unmodified.
int ref_stream_open(void* h)
{
    RefDev* d = (RefDev*)h;
    if (d->stream) return 0;
    d->stream = new SoapySDR::Stream(d->radio);
    return 0;
}
int ref_stream_set_format(void* h, const char* fmt) { return ((RefDev*)h)->stream->setFormat(fmt); }
void ref_stream_set_filter(void* h, int type)
{
    ((RefDev*)h)->stream->setDigitalFilter((SoapySDR::Stream::DigitalFilterType)type);
}
size_t ref_stream_mtu(void* h) { return ((RefDev*)h)->stream->mtu_size; }
int ref_stream_read(void* h, void* buffer, size_t num_elements)
{
    return ((RefDev*)h)->stream->ReadSamplesGen(buffer, num_elements, 1000000);
}

}  // extern "C"

// ---------------------------------------------------------------- iir1, unmodified
struct RefFilt {
    virtual ~RefFilt() {}
    virtual float  f32(float x) = 0;
    virtual double f64(double x) = 0;
    virtual int    sos(double* out) = 0;   // rows of (b0,b1,b2,a0,a1,a2)
    virtual void   reset() = 0;
    // one rail of a CS16 block: (float)v / 4096.0f (LC/CaribouLiteRadioCpp.cpp:91) -> filter<float>
    // (LC/iir/iir/Cascade.h:161-169) at full rate; every D-th output (global phase *pos) is kept.
    virtual size_t rail_f32(const int16_t* iq, int rail, size_t n, size_t pos, int D, float* out2) = 0;
};
template <class F>
static size_t rail_loop(F& f, const int16_t* iq, int rail, size_t n, size_t pos, int D, float* out2)
{
    size_t m = 0;
    size_t ph = pos % (size_t)D;                      // decimation phase carried, not divided per sample
    for (size_t k = 0; k < n; k++) {
        const float y = f.template filter<float>((float)iq[2 * k + rail] / 4096.0f);
        if (ph == 0) { out2[2 * m + rail] = y; m++; }
        if (++ph == (size_t)D) ph = 0;
    }
    return m;
}
template <class F>
static int dump_sos(F& f, double* out)
{
    const Iir::Cascade::Storage st = f.getCascadeStorage();
    for (int i = 0; i < st.maxStages; i++) {
        const Iir::Biquad& b = st.stageArray[i];
        out[6 * i + 0] = b.getB0(); out[6 * i + 1] = b.getB1(); out[6 * i + 2] = b.getB2();
        out[6 * i + 3] = b.getA0(); out[6 * i + 4] = b.getA1(); out[6 * i + 5] = b.getA2();
    }
    return st.maxStages;
}
template <int N>
struct RefButter : RefFilt {
    Iir::Butterworth::LowPass<N> f;   // default state = DirectFormII, as in CaribouliteStream.hpp:124-131
    RefButter(double fs, double fc) { f.setup(fs, fc); }
    float  f32(float x) override { return f.template filter<float>(x); }
    double f64(double x) override { return f.template filter<double>(x); }
    int    sos(double* out) override { return dump_sos(f, out); }
    void   reset() override { f.reset(); }
    size_t rail_f32(const int16_t* iq, int rail, size_t n, size_t pos, int D, float* out2) override
    { return rail_loop(f, iq, rail, n, pos, D, out2); }
};
template <int K>
struct RefSos : RefFilt {
    Iir::Custom::SOSCascade<K> f;
    RefSos(const double* c) {
        double a[K][6];
        memcpy(a, c, sizeof(a));
        f.setup(a);
    }
    float  f32(float x) override { return f.template filter<float>(x); }
    double f64(double x) override { return f.template filter<double>(x); }
    int    sos(double* out) override { return dump_sos(f, out); }
    void   reset() override { f.reset(); }
    size_t rail_f32(const int16_t* iq, int rail, size_t n, size_t pos, int D, float* out2) override
    { return rail_loop(f, iq, rail, n, pos, D, out2); }
};

extern "C" {

void* ref_butter_create(int order, double fs, double fc)
{
    try {
        switch (order) {
            case 1: return new RefButter<1>(fs, fc);
            case 2: return new RefButter<2>(fs, fc);
            case 3: return new RefButter<3>(fs, fc);
            case 4: return new RefButter<4>(fs, fc);
            case 5: return new RefButter<5>(fs, fc);
            case 6: return new RefButter<6>(fs, fc);
            case 7: return new RefButter<7>(fs, fc);
            case 8: return new RefButter<8>(fs, fc);
            default: return nullptr;
        }
    } catch (...) { return nullptr; }
}
// sos rows are scipy order (b0,b1,b2,a0,a1,a2), as Iir::Custom::SOSCascade takes them.
void* ref_sos_create(int nsos, const double* sos)
{
    try {
        switch (nsos) {
            case 1: return new RefSos<1>(sos);
            case 2: return new RefSos<2>(sos);
            case 3: return new RefSos<3>(sos);
            case 4: return new RefSos<4>(sos);
            default: return nullptr;
        }
    } catch (...) { return nullptr; }
}
void ref_filt_destroy(void* f) { delete (RefFilt*)f; }
void ref_filt_reset(void* f) { ((RefFilt*)f)->reset(); }
int  ref_filt_sos(void* f, double* out) { return ((RefFilt*)f)->sos(out); }
void ref_filt_run_f32(void* f, const float* in, float* out, size_t n)
{
    RefFilt* p = (RefFilt*)f;
    for (size_t i = 0; i < n; i++) out[i] = p->f32(in[i]);
}
void ref_filt_run_f64(void* f, const double* in, double* out, size_t n)
{
    RefFilt* p = (RefFilt*)f;
    for (size_t i = 0; i < n; i++) out[i] = p->f64(in[i]);
}
// The reference's in-place CS16 filter loop (CaribouliteStream.cpp:291-298) on one rail pair.
void ref_filt_run_cs16_inplace(void* fi, void* fq, int16_t* iq, size_t n)
{
    RefFilt* pi = (RefFilt*)fi; RefFilt* pq = (RefFilt*)fq;
    for (size_t k = 0; k < n; k++) {
        iq[2 * k + 0] = (int16_t)pi->f32((float)iq[2 * k + 0]);
        iq[2 * k + 1] = (int16_t)pq->f32((float)iq[2 * k + 1]);
    }
}

}  // extern "C"

// The whole RX path for one capture, composed of the reference's own functions in the order the
// north star names: caribou_smi_read (unmodified, MTU-sized requests as Stream::ReadSamples issues
// them, CaribouliteStream.cpp:304-323) -> CS16 -> CF32 (/4096.0f) -> iir1 cascade per rail ->
// keep every D-th sample.  Returns filtered samples written to out (interleaved CF32), or < 0.
extern "C" long ref_rx_path_f32(void* h, void* fi, void* fq, int D, float* out, size_t n_samples)
{
    RefDev* d = (RefDev*)h;
    RefFilt* pi = (RefFilt*)fi; RefFilt* pq = (RefFilt*)fq;
    const size_t mtu = caribou_smi_get_native_batch_samples(&d->sys->smi);
    int16_t* cs16 = (int16_t*)malloc((mtu + 1) * 4);
    uint8_t* meta = (uint8_t*)malloc(mtu + 1);
    size_t pos = 0, m = 0;
    long rc = 0;
    while (pos < n_samples) {
        const size_t want = n_samples - pos < mtu ? n_samples - pos : mtu;
        const int got = ref_dev_radio_read(h, cs16, meta, want);
        if (got < 0) { rc = got; break; }
        if (got == 0) break;
        const size_t mi = pi->rail_f32(cs16, 0, (size_t)got, pos, D, out + 2 * m);
        pq->rail_f32(cs16, 1, (size_t)got, pos, D, out + 2 * m);
        m += mi;
        pos += (size_t)got;
    }
    free(cs16); free(meta);
    return rc < 0 ? rc : (long)m;
}

// ---------------------------------------------------------------- circular_buffer<T>, unmodified
// (LC/datatypes/circular_buffer.h, header-only; reaches this file through Cariboulite.hpp.)  Elements are the
// 4-byte native samples the reference queues (CaribouliteStream.cpp:71).
extern "C" {
typedef circular_buffer<cariboulite_sample_complex_int16> RefRing;
void* ref_ring_create(size_t size, int override_write, int block_read)
{
    return new (std::nothrow) RefRing(size, override_write != 0, block_read != 0);
}
void ref_ring_destroy(void* r) { delete (RefRing*)r; }
size_t ref_ring_put(void* r, const void* data, size_t length)
{
    return ((RefRing*)r)->put((const cariboulite_sample_complex_int16*)data, length);
}
size_t ref_ring_get(void* r, void* data, size_t length, int timeout_us)
{
    return ((RefRing*)r)->get((cariboulite_sample_complex_int16*)data, length, timeout_us);
}
size_t ref_ring_size(void* r) { return ((RefRing*)r)->size(); }
size_t ref_ring_capacity(void* r) { return ((RefRing*)r)->capacity(); }
void ref_ring_reset(void* r) { ((RefRing*)r)->reset(); }
}  // extern "C"
