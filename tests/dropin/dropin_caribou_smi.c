/* tests/dropin - TEST INFRASTRUCTURE: the one translation unit a maintainer replaces.
 *
 * caribou_smi_read / caribou_smi_write with the reference's signatures (LC/caribou_smi/caribou_smi.h:96-100) and the
 * bodies INTEGRATION.md section 1 shows: they forward to libcaribou_cuda.  Everything else of caribou_smi.c is the
 * reference's own code (tests/dropin/Makefile compiles the unmodified file with these two functions renamed away).
 * caribou_smi_st stays byte-for-byte the reference's (objects compiled against the unmodified header link
 * unchanged): the GPU handle of a device lives in a side table keyed by the caribou_smi_st pointer instead of in a
 * new struct member (INTEGRATION.md shows both variants). */
#include <stddef.h>
#include "caribou_smi/caribou_smi.h"
#include "caribou_cuda.h"

#define DROPIN_MAX_DEV 16
static struct { caribou_smi_st* dev; caribou_cuda_smi* cuda; } g_tab[DROPIN_MAX_DEV];

static caribou_cuda_smi* cuda_of(caribou_smi_st* dev)
{
    for (int i = 0; i < DROPIN_MAX_DEV; i++)
        if (g_tab[i].dev == dev) return g_tab[i].cuda;
    return NULL;
}

/* what the tail of caribou_smi_init (caribou_smi.c:513-581) does once dev->filedesc is open */
int dropin_smi_attach(caribou_smi_st* dev, int cuda_device)
{
    for (int i = 0; i < DROPIN_MAX_DEV; i++) {
        if (g_tab[i].dev == NULL) {
            caribou_cuda_smi* c = caribou_cuda_smi_open(dev->filedesc, dev->native_batch_len, cuda_device);
            if (c == NULL) return -1;                      /* no CPU fallback */
            g_tab[i].dev = dev; g_tab[i].cuda = c;
            return 0;
        }
    }
    return -1;
}

/* caribou_smi_close */
void dropin_smi_detach(caribou_smi_st* dev)
{
    for (int i = 0; i < DROPIN_MAX_DEV; i++) {
        if (g_tab[i].dev == dev) {
            caribou_cuda_smi_close(g_tab[i].cuda);
            g_tab[i].dev = NULL; g_tab[i].cuda = NULL;
        }
    }
}

int caribou_smi_read(caribou_smi_st* dev, caribou_smi_channel_en channel,
                     caribou_smi_sample_complex_int16* buffer, caribou_smi_sample_meta* metadata,
                     size_t length_samples)
{
    if (dev->debug_mode != caribou_smi_none) return -2;                       /* caribou_smi.c:653, unchanged */
    return caribou_cuda_smi_read(cuda_of(dev),
                                 channel == caribou_smi_channel_2400 ? CARIBOU_CUDA_CH_HIF : CARIBOU_CUDA_CH_S1G,
                                 (caribou_cuda_sample_complex_int16*)buffer,   /* same packed layout, caribou_smi.h:53-64 */
                                 (caribou_cuda_sample_meta*)metadata, length_samples);
}

int caribou_smi_write(caribou_smi_st* dev, caribou_smi_channel_en channel,
                      caribou_smi_sample_complex_int16* samples, size_t length_samples)
{
    if (caribou_smi_set_driver_streaming_state(dev, smi_stream_tx_channel) != 0) return -1;   /* :728-735, unchanged */
    return caribou_cuda_smi_write(cuda_of(dev),
                                  channel == caribou_smi_channel_2400 ? CARIBOU_CUDA_CH_HIF : CARIBOU_CUDA_CH_S1G,
                                  (const caribou_cuda_sample_complex_int16*)samples, length_samples);
}
