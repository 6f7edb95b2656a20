// tests/dropin glue - TEST INFRASTRUCTURE ONLY.  Flat C entry points so the python test can drive the reference's
// own (unmodified) SoapySDR::Stream and cariboulite_radio_* code, linked against the GPU-backed caribou_smi_read of
// dropin_caribou_smi.c, on a synthetic capture held in a memfd.  No algorithm lives here.
#define _GNU_SOURCE 1
#include <sys/mman.h>
#include <unistd.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <new>

#include "Cariboulite.hpp"   // reference header: Stream, radio and sys_st types

extern "C" {
#include "caribou_smi/caribou_smi.h"
int  dropin_smi_attach(caribou_smi_st* dev, int cuda_device);
void dropin_smi_detach(caribou_smi_st* dev);
// hardware (init-only GPIO set-up of caribou_smi_setup_ios)
void io_utils_set_gpio_mode(int gpio, io_utils_alt_en mode) { (void)gpio; (void)mode; }
}

struct DropDev {
    sys_st* sys;
    cariboulite_radio_state_st* radio;
    SoapySDR::Stream* stream;
};

extern "C" {

void* dropin_dev_open(const uint8_t* bytes, size_t nbytes, size_t native_batch_len, int channel, int cuda_device)
{
    DropDev* d = new (std::nothrow) DropDev();
    if (!d) return nullptr;
    d->sys = (sys_st*)calloc(1, sizeof(sys_st));
    int fd = memfd_create("dropin_smi", 0);
    if (fd < 0) { free(d->sys); delete d; return nullptr; }
    size_t done = 0;
    while (done < nbytes) {
        ssize_t w = write(fd, bytes + done, nbytes - done);
        if (w <= 0) { close(fd); free(d->sys); delete d; return nullptr; }
        done += (size_t)w;
    }
    lseek(fd, 0, SEEK_SET);
    caribou_smi_st* smi = &d->sys->smi;       // what caribou_smi_init fills in (caribou_smi.c:513-581) minus the hardware
    smi->initialized = 1;
    smi->filedesc = fd;
    smi->native_batch_len = native_batch_len;
    smi->sample_rate = CARIBOU_SMI_SAMPLE_RATE;
    smi->debug_mode = caribou_smi_none;
    d->radio = &d->sys->radio_low;
    d->radio->sys = d->sys;
    d->radio->smi_channel_id = channel ? caribou_smi_channel_2400 : caribou_smi_channel_900;
    d->stream = nullptr;
    if (dropin_smi_attach(smi, cuda_device) != 0) { close(fd); free(d->sys); delete d; return nullptr; }
    return d;
}

void dropin_dev_close(void* h)
{
    DropDev* d = (DropDev*)h;
    if (!d) return;
    if (d->stream) delete d->stream;
    dropin_smi_detach(&d->sys->smi);
    close(d->sys->smi.filedesc);
    free(d->sys);
    delete d;
}

// the reference's cariboulite_radio_read_samples (LC/cariboulite_radio.c:1258-1285), unmodified
int dropin_radio_read(void* h, int16_t* cs16, uint8_t* meta, size_t length_samples)
{
    DropDev* d = (DropDev*)h;
    return cariboulite_radio_read_samples(d->radio, (cariboulite_sample_complex_int16*)cs16,
                                          (cariboulite_sample_meta*)meta, length_samples);
}
size_t dropin_radio_mtu(void* h) { return cariboulite_radio_get_native_mtu_size_samples(((DropDev*)h)->radio); }

// the reference's SoapySDR::Stream (LC/soapy_api/CaribouliteStream.cpp), unmodified
int dropin_stream_open(void* h)
{
    DropDev* d = (DropDev*)h;
    if (!d->stream) d->stream = new SoapySDR::Stream(d->radio);
    return 0;
}
int dropin_stream_set_format(void* h, const char* fmt) { return ((DropDev*)h)->stream->setFormat(fmt); }
void dropin_stream_set_filter(void* h, int type)
{
    ((DropDev*)h)->stream->setDigitalFilter((SoapySDR::Stream::DigitalFilterType)type);
}
size_t dropin_stream_mtu(void* h) { return ((DropDev*)h)->stream->mtu_size; }
int dropin_stream_read(void* h, void* buffer, size_t num_elements)
{
    return ((DropDev*)h)->stream->ReadSamplesGen(buffer, num_elements, 1000000);
}

}  // extern "C"
