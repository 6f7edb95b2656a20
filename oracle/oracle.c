/* oracle/oracle.c - integer half of the CPU oracle (see oracle.h header note).
 * TEST INFRASTRUCTURE ONLY - never part of the product path. */
#include "oracle.h"
#include <string.h>

static uint32_t ld_u32(const uint8_t* p)
{
    uint32_t v;
    memcpy(&v, p, 4);   /* reference does an unaligned *(uint32_t*) on a little-endian host */
    return v;
}

static int hdr_ok(uint32_t w) { return (w & 0xC001C000u) == 0x80004000u; }

/* LC/caribou_smi/caribou_smi.c:235-292 */
int orc_find_buffer_offset(const uint8_t* buffer, size_t len)
{
    if (len <= (size_t)(ORC_BYTES_PER_SAMPLE * 4)) return 0;            /* :240-243 */
    for (size_t offs = 0; offs < len - (size_t)(ORC_BYTES_PER_SAMPLE * 4); offs++) {  /* :248 */
        if (hdr_ok(ld_u32(buffer + offs)) && hdr_ok(ld_u32(buffer + offs + 4)) &&
            hdr_ok(ld_u32(buffer + offs + 8)) && hdr_ok(ld_u32(buffer + offs + 12)))
            return (int)offs;
    }
    return -1;                                                          /* :285-289 */
}

static int16_t sx13(uint32_t v)
{
    int16_t s = (int16_t)(v & 0x1FFFu);
    if (s >= (int16_t)0x1000) s -= (int16_t)0x2000;                     /* :356-357 */
    return s;
}

/* LC/caribou_smi/caribou_smi.c:295-393 */
int orc_rx_data_analyze(int channel, const uint8_t* data, size_t data_length,
                        int16_t* cs16, uint8_t* meta)
{
    int offs = orc_find_buffer_offset(data, data_length);
    if (offs < 0) return -1;
    size_t actual_length = data_length;
    int shortening = (offs > 0) ? (offs / ORC_BYTES_PER_SAMPLE + 1) : 0;    /* :323 */
    actual_length -= (size_t)shortening * ORC_BYTES_PER_SAMPLE;
    const uint8_t* p = data + offs;
    size_t n = actual_length / ORC_BYTES_PER_SAMPLE;
    size_t i;
    for (i = 0; i < n; i++) {
        uint32_t s = ld_u32(p + 4 * i);
        if (meta) meta[i] = (uint8_t)(s & 1u);                          /* :348 */
        if (cs16) {
            uint32_t lo = (s >> 1) & 0x1FFFu;                            /* bits 13:1  */
            uint32_t hi = (s >> 17) & 0x1FFFu;                           /* bits 29:17 */
            if (channel != ORC_CH_HIF) {                                 /* S1G: q=lo, i=hi (:351-354) */
                cs16[2 * i + 0] = sx13(hi);
                cs16[2 * i + 1] = sx13(lo);
            } else {                                                     /* HiF: i=lo, q=hi (:371-373) */
                cs16[2 * i + 0] = sx13(lo);
                cs16[2 * i + 1] = sx13(hi);
            }
        }
    }
    /* :382-389 - one extrapolated sample after a misaligned chunk; its meta is not written.
     * (The reference dereferences cmplx_vec unconditionally here; with cs16 == NULL that is
     * undefined behaviour upstream, so the oracle just skips it.) */
    if (shortening > 0 && cs16 && i >= 2) {
        cs16[2 * i + 0] = (int16_t)(110 * cs16[2 * (i - 1) + 0] / 100 - cs16[2 * (i - 2) + 0] / 10);
        cs16[2 * i + 1] = (int16_t)(110 * cs16[2 * (i - 1) + 1] / 100 - cs16[2 * (i - 2) + 1] / 10);
    }
    return offs;
}

/* LC/caribou_smi/caribou_smi.c:632-682 */
long orc_smi_read(const uint8_t* bytes, size_t nbytes, size_t* cursor, size_t native_batch_len,
                  int channel, int16_t* cs16, uint8_t* meta, size_t length_samples)
{
    size_t left_to_read = length_samples * ORC_BYTES_PER_SAMPLE;
    size_t read_so_far = 0;
    while (left_to_read) {
        int16_t* so = cs16 ? cs16 + 2 * read_so_far : NULL;
        uint8_t* mo = meta ? meta + read_so_far : NULL;
        size_t cur = left_to_read > native_batch_len ? native_batch_len : left_to_read;   /* :649 */
        size_t avail = nbytes - *cursor;
        size_t ret = cur < avail ? cur : avail;                       /* read(2) on the stream */
        if (ret == 0) break;                                            /* :657-661 timeout */
        int r = orc_rx_data_analyze(channel, bytes + *cursor, ret, so, mo);
        *cursor += ret;
        if (r < 0) return -3;                                           /* :665-668 */
        read_so_far += ret / ORC_BYTES_PER_SAMPLE;                      /* :677 */
        left_to_read -= ret;
    }
    return (long)read_so_far;
}

void orc_flag_scan(const uint32_t* words, size_t n,
                   uint32_t* invalid_idx, size_t cap_invalid, uint64_t* n_invalid,
                   uint32_t* sync_idx, size_t cap_sync, uint64_t* n_sync)
{
    uint64_t ni = 0, ns = 0;
    for (size_t k = 0; k < n; k++) {
        uint32_t w = words[k];
        if (!hdr_ok(w)) { if (invalid_idx && ni < cap_invalid) invalid_idx[ni] = (uint32_t)k; ni++; }
        if (w & 1u)     { if (sync_idx && ns < cap_sync) sync_idx[ns] = (uint32_t)k; ns++; }
    }
    if (n_invalid) *n_invalid = ni;
    if (n_sync) *n_sync = ns;
}

size_t orc_demux_gather(const uint32_t* words, size_t n, size_t B, int sub, uint32_t* out)
{
    size_t m = 0;
    for (size_t k = 0; k < n; k++)
        if ((int)((k / B) & 1u) == sub) out[m++] = words[k];
    return m;
}

static uint64_t splitmix64(uint64_t x)
{
    uint64_t z = x + 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

void orc_gen_words(uint64_t n0, size_t n, int hif_layout, uint64_t seed,
                   const int16_t* tone_i, const int16_t* tone_q, uint32_t* out)
{
    for (size_t k = 0; k < n; k++) {
        uint64_t g = n0 + k;
        int ph = (int)(g % 800u);
        int ui = (int)(splitmix64(seed ^ (g * 2 + 0)) & 127u) - 64;
        int uq = (int)(splitmix64(seed ^ (g * 2 + 1)) & 127u) - 64;
        int I = tone_i[ph] + ui, Q = tone_q[ph] + uq;
        if (I < -4096) I = -4096; if (I > 4095) I = 4095;
        if (Q < -4096) Q = -4096; if (Q > 4095) Q = 4095;
        uint32_t S = (g % 4000000u) == 0;
        uint32_t hi = (uint32_t)(hif_layout ? Q : I) & 0x1FFFu;
        uint32_t lo = (uint32_t)(hif_layout ? I : Q) & 0x1FFFu;
        out[k] = 0x80004000u | (hi << 17) | (lo << 1) | S;
    }
}

void orc_cs16_to_cf32(const int16_t* iq, float* out, size_t n)
{
    const float max_val = 4096.0f;                                      /* CaribouliteStream.cpp:313 */
    for (size_t k = 0; k < 2 * n; k++) out[k] = (float)iq[k] / max_val;
}
void orc_cs16_to_cf64(const int16_t* iq, double* out, size_t n)
{
    const double max_val = 4096.0;                                      /* :335 */
    for (size_t k = 0; k < 2 * n; k++) out[k] = (double)iq[k] / max_val;
}
void orc_cs16_to_cs8(const int16_t* iq, int8_t* out, size_t n)
{
    for (size_t k = 0; k < 2 * n; k++) out[k] = (int8_t)((iq[k] >> 5) & 0x00FF);   /* :362-363 */
}

/* ---- TX direction.  caribou_smi_generate_data, caribou_smi.c:688-717: the reference's loop with
 * the sample read it has commented out (`cmplx_vec[i].i` / `.q`) put back - as committed it packs the constant
 * ii = 0xFFFF, qq = 0, which orc_tx_pack reproduces for the sample (i = -1, q = 0); that word is the one vector
 * the reference pins.  Layout comment caribou_smi.c:693-696:
 *   [SOF TXC CTX I12..I8] [0 I7..I1] [0 I0 Q12..Q7] [0 Q6..Q0], byte 0 first on the wire (bswap32 on a
 *   little-endian host). */
void orc_tx_pack(const int16_t* iq, size_t n, uint32_t* out)
{
    for (size_t k = 0; k < n; k++) {
        int32_t ii = iq[2 * k], qq = iq[2 * k + 1];
        ii &= 0x1FFF;
        qq &= 0x1FFF;
        uint32_t s = (1u << 2) | (1u << 1) | (1u << 0);   /* SMI_TX_SAMPLE_SOF | MODEM_TX_CTRL | COND_TX_CTRL */
        s <<= 5; s |= (uint32_t)(ii >> 8) & 0x1Fu;
        s <<= 8; s |= (uint32_t)(ii >> 1) & 0x7Fu;
        s <<= 2; s |= (uint32_t)(ii & 0x1);
        s <<= 6; s |= (uint32_t)(qq >> 7) & 0x3Fu;
        s <<= 8; s |= (uint32_t)(qq & 0x7F);
        out[k] = __builtin_bswap32(s);
    }
}
/* Stream::WriteSamples conversions to the native int16 pair, CaribouliteStream.cpp:199-244.  The reference's
 * (int16_t)(x * 4096) is undefined outside the int range; here, as on x86-64 and on the GPU within +-2^31/4096,
 * it is truncation toward zero to int32 followed by keeping the low 16 bits. */
static int16_t tx_i16_from_f64(double v)
{
    if (!(v > -2147483648.0 && v < 2147483648.0)) return 0;   /* cvttsd2si's "integer indefinite" 0x80000000: low half 0 */
    return (int16_t)(int32_t)v;
}
void orc_tx_from_cf32(const float* x, size_t n, int16_t* iq)
{
    const float max_val = 4096.0f;                                      /* :202 */
    for (size_t k = 0; k < 2 * n; k++) iq[k] = tx_i16_from_f64((double)(x[k] * max_val));
}
void orc_tx_from_cf64(const double* x, size_t n, int16_t* iq)
{
    const double max_val = 4096.0;                                      /* :217 */
    for (size_t k = 0; k < 2 * n; k++) iq[k] = tx_i16_from_f64(x[k] * max_val);
}
void orc_tx_from_cs8(const int8_t* x, size_t n, int16_t* iq)
{
    for (size_t k = 0; k < 2 * n; k++) iq[k] = (int16_t)(((int16_t)x[k]) << 5);     /* :236-237 */
}
