/* libcaribou_cuda - B200 (sm_100a) implementation of CaribouLite's receive-side sample path.
 *
 * Plain C ABI: pointers, sizes and PODs only.  Every entry point names the reference
 * interface it stands in for ("LC/" = software/libcariboulite/src/ of Habraken/cariboulite).
 *
 *   raw SMI words -> (alignment search per chunk) -> demux -> 13-bit sign-extend -> CS16 / meta
 *                 -> CF32 -> Butterworth biquad cascade (iir1 DF-II semantics, FP64) -> decimate
 *
 * There is no CPU fallback: every compute entry point launches CUDA kernels and fails with
 * CARIBOU_CUDA_ERR (-1) if the device or the kernels are unavailable.
 *
 * Return convention (same as LC/caribou_smi/caribou_smi.c:632-682 and
 * LC/cariboulite_radio.c:1258-1285): >= 0 number of samples delivered, 0 = nothing to
 * read ("timeout"), -1 = I/O / CUDA failure, -2 = reserved (debug mode, never produced),
 * -3 = a chunk had no word alignment ("sync lost").
 */
#ifndef CARIBOU_CUDA_H
#define CARIBOU_CUDA_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define CARIBOU_CUDA_OK            0
#define CARIBOU_CUDA_ERR          (-1)
#define CARIBOU_CUDA_ERR_DEBUG    (-2)
#define CARIBOU_CUDA_ERR_SYNC     (-3)

#define CARIBOU_CUDA_BYTES_PER_SAMPLE  4        /* LC/caribou_smi/caribou_smi.h:42 */
#define CARIBOU_CUDA_NATIVE_BATCH_LEN  524288   /* LC/caribou_smi/caribou_smi.c:74-79 */
#define CARIBOU_CUDA_MAX_SOS           4

/* caribou_smi_channel_en, LC/caribou_smi/caribou_smi.h:45-49: the channel only selects
 * which 13-bit field is I and which is Q (LC/caribou_smi/caribou_smi.c:342-379). */
typedef enum { CARIBOU_CUDA_CH_S1G = 0, CARIBOU_CUDA_CH_HIF = 1 } caribou_cuda_channel;

typedef enum { CARIBOU_CUDA_MEM_DEVICE = 0,        /* pointers are device memory */
               CARIBOU_CUDA_MEM_HOST_PINNED = 1,   /* page-locked host memory: async copies */
               CARIBOU_CUDA_MEM_HOST = 2           /* pageable host memory: staged, synchronous */
} caribou_cuda_mem;
/* caribou_cuda_unpack_submit only: input bytes of kind `in`, output arrays of kind `out` (the shims read(2) into a
 * page-locked buffer and deliver into the caller's pageable arrays) */
#define CARIBOU_CUDA_MEM_IO(in, out)  (0x100 | ((in) & 0xF) | (((out) & 0xF) << 4))

/* Sample formats of SoapyCariboulite's setupStream (LC/soapy_api/CaribouliteStreamFunctions.cpp:11-19). */
typedef enum { CARIBOU_CUDA_FMT_CF32 = 0, CARIBOU_CUDA_FMT_CS16 = 1,
               CARIBOU_CUDA_FMT_CS8 = 2,  CARIBOU_CUDA_FMT_CF64 = 3 } caribou_cuda_format;

/* Where the int16 truncation sits relative to the filter.
 * Q16: reference ordering, LC/soapy_api/CaribouliteStream.cpp:291-298 then :315-321:
 *      y = (int16)(float)cascade((float)x_int16); out = y / 4096.
 * F32: CS16 -> CF32 first, cascade on the float, no integer truncation:
 *      out = (float)cascade((float)x_int16 / 4096). */
typedef enum { CARIBOU_CUDA_PARITY_Q16 = 0, CARIBOU_CUDA_PARITY_F32 = 1 } caribou_cuda_parity;

/* Input element type of caribou_cuda_filter_submit / rx path. */
typedef enum { CARIBOU_CUDA_IN_WORDS = 0,   /* raw 32-bit SMI words (fused unpack) */
               CARIBOU_CUDA_IN_CS16 = 1,    /* packed (int16 i, int16 q)           */
               CARIBOU_CUDA_IN_CF32 = 2     /* packed (float i, float q)            */
} caribou_cuda_input;

/* Realisation of the cascade inside the scan kernels (caribou_cuda_config.kernel_form). */
typedef enum { CARIBOU_CUDA_FORM_AUTO = 0, CARIBOU_CUDA_FORM_CASCADE = 1, CARIBOU_CUDA_FORM_PARALLEL = 2 } caribou_cuda_form;

typedef struct caribou_cuda_ctx caribou_cuda_ctx;

typedef struct {
    int      device;                  /* CUDA device ordinal */
    int      channel;                 /* caribou_cuda_channel (single-channel captures) */
    uint32_t interleave_block_words;  /* 0: single channel.  B > 0: dual-channel capture, word k belongs to
                                         S1G if (k / B) % 2 == 0 else HiF */
    size_t   chunk_bytes;             /* read(2) granularity = caribou_smi_st.native_batch_len; 0 -> 524288.
                                         Must be a multiple of 16. */
    int      n_streams;               /* independent captures per submit (equal length); 0 -> 1 */
    /* digital filter: cascade of n_sos biquads, rows (b0,b1,b2,a0,a1,a2) as Iir::Custom::SOSCascade
     * takes them (LC/iir/iir/Custom.h:90-112).  n_sos == 0: no filter. */
    int      n_sos;
    double   sos[CARIBOU_CUDA_MAX_SOS][6];
    int      decimation;              /* keep every D-th filtered sample (phase 0); 0 -> 1.  D in {1,2,4,8,16,32} */
    int      parity;                  /* caribou_cuda_parity */
    int      out_format;              /* caribou_cuda_format of the filtered output */
    void*    cuda_stream;             /* cudaStream_t to run on; NULL -> the library creates one */
    int      kernel_form;             /* caribou_cuda_form: AUTO picks the parallel (partial-fraction) realisation of the
                                         cascade when the poles allow it; CASCADE keeps the reference's stage-by-stage
                                         arithmetic order inside every run */
} caribou_cuda_config;

/* What the scan kernel does with the configured cascade (for tests, DESIGN.md numbers and the curious). */
typedef struct {
    int      n_sos, decimation;
    int      form;                 /* CASCADE or PARALLEL: the realisation the kernels run */
    int      run_len;              /* samples one thread filters sequentially */
    int      tile_samples;         /* samples per warp tile (32 runs) */
    int      lookback_window;      /* tiles of look-back (aggregates only); meaningless when chained */
    int      chained;              /* 1: no useful contraction, tiles wait for the predecessor's inclusive state */
    uint64_t decay_samples;        /* samples after which an initial state has decayed below 2^-60 (0 if chained) */
    double   parallel_form_error;  /* parallel vs cascade realisation, rel. rms in extended precision */
    char     note[96];             /* why the cascade form was kept, if it was; full-rate plans that apply the numerator
                                      as integer two-tap sums say so here (with the form's measured error) */
} caribou_cuda_filter_info_t;

/* Output buffers of the unpack stage.  Any pointer may be NULL (= not wanted), as in
 * LC/caribou_smi/caribou_smi.c:348-349.  Slots the reference leaves unwritten (tail of a
 * misaligned chunk, everything from the first un-alignable chunk on) are left untouched.
 * In dual-channel mode the *_b members receive the HiF sub-stream and the plain members
 * the S1G sub-stream; each sub-stream is chunked on its own. */
typedef struct {
    void*     cs16;          /* [n] (int16 i, int16 q): caribou_smi_sample_complex_int16 */
    void*     meta;          /* [n] uint8 sync bit:     caribou_smi_sample_meta          */
    void*     cf32;          /* [n] (float i, float q) = cs16 / 4096 (LC/CaribouLiteRadioCpp.cpp:91) */
    void*     cs16_b;
    void*     meta_b;
    void*     cf32_b;
    uint32_t* invalid_idx;   /* ascending indices of words failing (w & 0xC001C000) == 0x80004000 */
    size_t    invalid_cap;
    uint32_t* sync_idx;      /* ascending indices of words with S == 1 */
    size_t    sync_cap;
} caribou_cuda_unpack_out;

typedef struct {
    int64_t  n_samples;      /* samples delivered (per sub-stream in dual-channel mode: S1G) */
    int64_t  n_samples_b;    /* HiF sub-stream */
    int64_t  n_filtered;     /* filtered/decimated samples written per stream (batch: the minimum over the streams) */
    uint64_t n_invalid;      /* exact, even when the index list is capped */
    uint64_t n_sync;
    int32_t  first_bad_chunk;    /* first chunk with no alignment, -1 if none */
    int32_t  n_misaligned_chunks;/* chunks with offs > 0 */
} caribou_cuda_result;

/* One capture of a batch (caribou_cuda_rx_submit with n_streams > 1): the captures are independent reads, so each
 * has its own return value exactly as caribou_smi_read would give it (LC/caribou_smi/caribou_smi.c:632-682). */
typedef struct {
    int64_t  n_samples;          /* samples consumed by the filter, or CARIBOU_CUDA_ERR_SYNC (-3) when a chunk of THIS
                                    capture had no alignment (caribou_smi.c:665-668) */
    int64_t  n_filtered;         /* filtered/decimated samples written for this capture (on -3: those of the chunks
                                    before the un-alignable one) */
    int32_t  first_bad_chunk;    /* -1 if none */
    int32_t  n_misaligned_chunks;/* chunks of this capture with offs > 0 */
} caribou_cuda_stream_result;

/* ---------------------------------------------------------------- lifecycle */
/* Replaces caribou_smi_init's buffer/stream set-up (LC/caribou_smi/caribou_smi.c:513-581) and the
 * filter objects of SoapySDR::Stream (LC/soapy_api/CaribouliteStream.cpp:51-97). */
caribou_cuda_ctx* caribou_cuda_create(const caribou_cuda_config* cfg);
void        caribou_cuda_destroy(caribou_cuda_ctx* ctx);
const char* caribou_cuda_last_error(void);                 /* thread-local message of the last failure */
void*       caribou_cuda_get_stream(caribou_cuda_ctx* ctx); /* cudaStream_t the context launches on */
/* Block until everything submitted on the context has finished; fills *res (may be NULL) with the
 * counters of the last submit.  Returns the last submit's sample count or error code. */
int64_t     caribou_cuda_wait(caribou_cuda_ctx* ctx, caribou_cuda_result* res);
/* Per-capture results of the last caribou_cuda_rx_submit (call after caribou_cuda_wait): fills out[0..n_streams).
 * A capture that lost word alignment does not disturb the others: they stay on the fused kernel, the affected
 * capture alone is redone through unpack -> CS16 -> filter with the reference's chunk semantics.  Several
 * submits may be queued before one wait; a misaligned capture is then redone for all of them, in order. */
int         caribou_cuda_stream_results(caribou_cuda_ctx* ctx, caribou_cuda_stream_result* out);
/* Number of kernels this context has launched so far (bench.py's gpu_launches). */
uint64_t    caribou_cuda_launch_count(const caribou_cuda_ctx* ctx);

/* Profiling aid for bench.py's roofline leg: when enabled, every submit brackets its dominant kernel
 * (the IIR scan kernel; the unpack kernel for a context without filter) with CUDA events on the
 * launching stream.  caribou_cuda_read_profile waits for the context, writes the device time in
 * milliseconds of each bracketed launch since the previous read (oldest first, at most the last 256)
 * to ms[0..cap) and returns how many it wrote (-1 on error). */
int caribou_cuda_set_profiling(caribou_cuda_ctx* ctx, int enable);
int caribou_cuda_read_profile(caribou_cuda_ctx* ctx, float* ms, int cap);

/* page-locked host buffers for CARIBOU_CUDA_MEM_HOST_PINNED */
void* caribou_cuda_host_alloc(size_t nbytes);
void  caribou_cuda_host_free(void* p);

/* ---------------------------------------------------------------- kernel (a): unpack
 * caribou_smi_rx_data_analyze + caribou_smi_find_buffer_offset for every chunk of the byte
 * stream (LC/caribou_smi/caribou_smi.c:235-393) driven like caribou_smi_read's loop (:632-682),
 * with the byte stream handed over in one piece instead of read(2) by read(2).
 * `bytes` holds nbytes of raw SMI data per stream (streams `stream_stride_bytes` apart).
 * Asynchronous for DEVICE / HOST_PINNED memory; call caribou_cuda_wait for the result. */
int caribou_cuda_unpack_submit(caribou_cuda_ctx* ctx, const void* bytes, size_t nbytes,
                               size_t stream_stride_bytes, int mem,
                               const caribou_cuda_unpack_out* out);

/* Kernel (a) without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same kernels and same outputs as caribou_cuda_unpack_submit on DEVICE memory for one single-channel capture,
 * but no context: the per-chunk offsets, per-tile flag counts, their scan and the work list live in ONE fixed
 * device array inside the library (CARIBOU_CUDA_NOALLOC_SCRATCH_BYTES per GPU that loads the library, held until the
 * process exits).  The call makes no cudaMalloc / cudaFree / host allocation and creates no stream or event: it only
 * queues memsets and kernels on `cuda_stream` (a cudaStream_t; NULL = the default stream) and returns without
 * synchronising.
 *   device        CUDA device index (made current)
 *   channel       CARIBOU_CUDA_CH_S1G / CARIBOU_CUDA_CH_HIF (which 13-bit field is I)
 *   chunk_bytes   read(2) granularity, 0 -> 524288; multiple of 16
 *   d_bytes       nbytes of raw SMI data in DEVICE memory, 16-byte aligned
 *   out           DEVICE pointers (any may be NULL; the *_b members must be NULL; index lists allowed)
 *   d_result      DEVICE memory for one caribou_cuda_result (may be NULL): what caribou_cuda_wait would report
 * Limits: nbytes <= CARIBOU_CUDA_NOALLOC_MAX_BYTES and ceil(nbytes / chunk_bytes) <= CARIBOU_CUDA_NOALLOC_MAX_CHUNKS;
 * anything larger is rejected with CARIBOU_CUDA_ERR and a message (caribou_cuda_last_error), nothing is launched.
 * Every call on a GPU shares the one array: calls are safe back to back on one stream (each clears what it needs),
 * calls that overlap on the GPU (different streams without an event between them, or different threads) RACE. */
#define CARIBOU_CUDA_NOALLOC_MAX_BYTES     (64u << 20)   /* 2^24 SMI words per call */
#define CARIBOU_CUDA_NOALLOC_MAX_CHUNKS    16384
#define CARIBOU_CUDA_NOALLOC_SCRATCH_BYTES 147520
int caribou_cuda_unpack_noalloc(int device, void* cuda_stream, int channel, size_t chunk_bytes,
                                const void* d_bytes, size_t nbytes, const caribou_cuda_unpack_out* out,
                                caribou_cuda_result* d_result);

/* ---------------------------------------------------------------- kernel (b): IIR + decimate
 * SoapySDR::Stream::ReadSamples' filter loop and format conversion
 * (LC/soapy_api/CaribouliteStream.cpp:282-367) on `n_samples` per stream, as a block-parallel
 * scan; persistent per-stream filter state like the never-reset filter objects of the Stream.
 * in_kind = WORDS fuses kernel (a)'s decode (aligned chunks only: offs == 0 everywhere; use
 * caribou_cuda_rx_submit for arbitrary input).  With decimation D, n_samples must be a
 * multiple of D.  out receives n_samples / D elements of cfg.out_format per stream
 * (streams `out_stride_elems` elements apart). */
int caribou_cuda_filter_submit(caribou_cuda_ctx* ctx, int in_kind, const void* in, size_t n_samples,
                               size_t in_stride_elems, int mem, void* out, size_t out_stride_elems);

/* ---------------------------------------------------------------- fused path
 * raw SMI bytes -> filtered, decimated output (the north-star path): alignment search, then the
 * fused unpack+IIR kernel when every chunk is aligned, else unpack -> CS16 scratch -> IIR.
 * `unpack_out` may be NULL (filtered output only). */
int caribou_cuda_rx_submit(caribou_cuda_ctx* ctx, const void* bytes, size_t nbytes,
                           size_t stream_stride_bytes, int mem,
                           const caribou_cuda_unpack_out* unpack_out,
                           void* filt_out, size_t filt_stride_elems);

/* ---------------------------------------------------------------- filter state
 * The reference's only persistent state: DirectFormII (v1, v2) per biquad per rail
 * (LC/iir/iir/State.h:102-131).  Layout per stream: [rail I: v1_0,v2_0,v1_1,v2_1,...][rail Q: ...],
 * 4 * n_sos doubles.  These calls synchronise the context. */
int caribou_cuda_get_state(caribou_cuda_ctx* ctx, int stream, double* state);
int caribou_cuda_set_state(caribou_cuda_ctx* ctx, int stream, const double* state);
int caribou_cuda_reset(caribou_cuda_ctx* ctx);   /* zero state, decimation phase and sample position */
int caribou_cuda_filter_info(caribou_cuda_ctx* ctx, caribou_cuda_filter_info_t* info);

/* Multi-GPU time sharding: a shard is filtered from zero state; its end state f
 * (caribou_cuda_get_state) travels to the next rank; the receiving rank then re-runs the first
 * few tiles of its shard from the true incoming state with caribou_cuda_filter_fixup.
 * caribou_cuda_shard_matrix returns A^n_samples (row-major, (2*n_sos)^2 doubles, reference
 * state coordinates) so that s_end = M * s_in + f.  caribou_cuda_fixup_samples returns how many
 * leading samples a non-zero incoming state can influence above 2^-60 (bounded by n_samples). */
int    caribou_cuda_shard_matrix(caribou_cuda_ctx* ctx, uint64_t n_samples, double* m);
size_t caribou_cuda_fixup_samples(caribou_cuda_ctx* ctx, size_t n_samples);
int    caribou_cuda_filter_fixup(caribou_cuda_ctx* ctx, int in_kind, const void* in, size_t n_samples,
                                 size_t in_stride_elems, int mem, void* out, size_t out_stride_elems,
                                 const double* state_in /* 4*n_sos doubles per stream, host */);

/* Device-side hand-off (BASELINE.json configs[4]): one long stream, time-chunked over the GPUs of
 * a box.  Rank g of `world` owns samples [g n, (g+1) n) of every call; state and decimation phase carry over from
 * call to call exactly as on one GPU (the reference's behaviour to match: filter objects that are never reset,
 * LC/soapy_api/CaribouliteStream.hpp:124-131, fed MTU after MTU by LC/caribou_smi/caribou_smi.c:643-679).
 *
 *   caribou_cuda_shard_submit =  pass 1   fused kernel over the shard from ZERO state -> outputs + end vector f_g
 *                                hand-off ncclAllGather of the 4K doubles per stream on the context's stream (or the
 *                                         ncclSend/ncclRecv chain, mode 1), then a one-CTA kernel composes
 *                                         s_in(g) = sum_{j<g} M^(g-1-j) f_j + M^g carry   and the new carry, M = A^n
 *                                fix-up   the leading tiles (caribou_cuda_fixup_samples) re-run from s_in(g), read
 *                                         from device memory
 * Nothing on that path touches the host: no stream synchronisation, no host copy.  caribou_cuda_wait returns
 * n_samples (or -3 if a chunk of the shard lost word alignment: shards must consist of aligned chunks).
 * With the peer exchange (mode 2) the composition and the fix-up of a call are queued one call late - after the pass
 * of the next call, when every peer's vector has long arrived - so `in` and `out` of a sharded call must stay valid
 * until the next caribou_cuda_shard_submit has returned or caribou_cuda_wait (or any other call on the context) has
 * flushed the pending hand-off; the outputs are final after caribou_cuda_wait, as for every asynchronous submit.
 *
 * NCCL is bound at run time (dlopen of libnccl.so.2, so that a process that already carries an NCCL - PyTorch's -
 * shares it): caribou_cuda_shard_unique_id wraps ncclGetUniqueId (rank 0 calls it and ships the 128 bytes to the
 * other ranks by any means), caribou_cuda_shard_init wraps ncclCommInitRank; caribou_cuda_shard_attach_comm borrows an
 * ncclComm_t the caller owns.  With id == NULL no communicator is created and the caller moves the end vectors
 * itself between caribou_cuda_shard_begin and caribou_cuda_shard_finish (caribou_cuda_shard_gather_buffer: device
 * memory, [world][n_streams * 4 * n_sos] doubles, slot `rank` written by begin, all slots read by finish), or lets
 * caribou_cuda_shard_exchange_local do it for `world` contexts of ONE process (one per GPU, peer copies). */
#define CARIBOU_CUDA_SHARD_ID_BYTES 128
int    caribou_cuda_shard_unique_id(void* id /* CARIBOU_CUDA_SHARD_ID_BYTES */);
int    caribou_cuda_shard_init(caribou_cuda_ctx* ctx, int rank, int world, const void* id);
int    caribou_cuda_shard_attach_comm(caribou_cuda_ctx* ctx, int rank, int world, void* nccl_comm);
/* 0: ncclAllGather, 1: ncclSend/ncclRecv chain, 2: peer stores - a kernel writes the end vector straight into every
 * rank's exchange buffer over NVLink (buffers mapped with CUDA IPC at init; the default whenever every pair of GPUs of
 * the communicator can map each other, else 0).  All ranks must use the same mode.  caribou_cuda_shard_mode returns it. */
int    caribou_cuda_shard_set_mode(caribou_cuda_ctx* ctx, int mode);
int    caribou_cuda_shard_mode(caribou_cuda_ctx* ctx);
int    caribou_cuda_shard_submit(caribou_cuda_ctx* ctx, int in_kind, const void* in, size_t n_samples,
                                 size_t in_stride_elems, void* out, size_t out_stride_elems);
int    caribou_cuda_shard_begin(caribou_cuda_ctx* ctx, int in_kind, const void* in, size_t n_samples,
                                size_t in_stride_elems, void* out, size_t out_stride_elems);
void*  caribou_cuda_shard_gather_buffer(caribou_cuda_ctx* ctx, size_t* bytes_per_rank);
int    caribou_cuda_shard_finish(caribou_cuda_ctx* ctx);
int    caribou_cuda_shard_exchange_local(caribou_cuda_ctx* const* ctxs, int world);
/* Device times of the last sharded call when profiling is enabled (caribou_cuda_set_profiling): pass 1, and
 * hand-off + fix-up, in milliseconds.  Synchronises the context. */
int    caribou_cuda_shard_profile(caribou_cuda_ctx* ctx, float* pass_ms, float* handoff_ms);

/* ---------------------------------------------------------------- host-side filter design
 * Iir::Butterworth::LowPass<order>::setup(fs, fc) (LC/iir/iir/Butterworth.cpp:49-68,107-117,
 * PoleFilter.cpp:47-90, Cascade.cpp:94-118).  Writes (order+1)/2 rows of (b0,b1,b2,a0,a1,a2);
 * returns the row count or -1. */
int caribou_cuda_butter_lowpass(int order, double fs, double fc, double* sos);

/* ---------------------------------------------------------------- synthetic input (bench/test utility)
 * Device-side generator of the deterministic SMI word pattern of SURVEY 8(d); bit-identical to
 * oracle/oracle.c::orc_gen_words given the same 800-entry tone tables. */
int caribou_cuda_synth_words(caribou_cuda_ctx* ctx, void* dev_words, uint64_t n0, size_t n,
                             int hif_layout, uint64_t seed,
                             const int16_t* tone_i_host, const int16_t* tone_q_host);

/* ---------------------------------------------------------------- drop-in shims
 * Same signatures and struct layouts as the reference, fed from a file descriptor
 * (/dev/smi, a capture file or a memfd).  See INTEGRATION.md. */
#pragma pack(push, 1)
typedef struct { int16_t i; int16_t q; } caribou_cuda_sample_complex_int16;   /* caribou_smi.h:53-59 */
typedef struct { uint8_t sync; } caribou_cuda_sample_meta;                      /* caribou_smi.h:61-64 */
#pragma pack(pop)

typedef struct caribou_cuda_smi caribou_cuda_smi;   /* stands in for caribou_smi_st */
caribou_cuda_smi* caribou_cuda_smi_open(int fd, size_t native_batch_len, int device);
void              caribou_cuda_smi_close(caribou_cuda_smi* dev);
/* int caribou_smi_read(caribou_smi_st*, caribou_smi_channel_en, caribou_smi_sample_complex_int16*,
 *                      caribou_smi_sample_meta*, size_t)            LC/caribou_smi/caribou_smi.h:96-97 */
int caribou_cuda_smi_read(caribou_cuda_smi* dev, int channel,
                          caribou_cuda_sample_complex_int16* buffer,
                          caribou_cuda_sample_meta* metadata, size_t length_samples);
/* size_t caribou_smi_get_native_batch_samples(caribou_smi_st*)     LC/caribou_smi/caribou_smi.c:765-769 */
size_t caribou_cuda_smi_get_native_batch_samples(caribou_cuda_smi* dev);

/* int CaribouLiteRadio::ReadSamples(std::complex<float>* samples, size_t num_to_read, uint8_t* meta)
 *                                                LC/CaribouLite.hpp:188-189, LC/CaribouLiteRadioCpp.cpp:72-95
 * CF32 (= native / 4096.0f) and the sync metadata of one read in one call and one device synchronisation; `samples`
 * is num_to_read (float re, float im) pairs, `meta` may be NULL.  The radio's persistent _read_samples /
 * _read_metadata buffers live on the device.  Returns what cariboulite_radio_read_samples returns; nothing is
 * written for a result <= 0. */
int caribou_cuda_radio_read_cf32(caribou_cuda_smi* dev, int channel, float* samples, uint8_t* meta, size_t num_to_read);

/* SoapySDR::Stream's RX half (LC/soapy_api/CaribouliteStream.hpp:64-137): format switch, MTU clamp,
 * three never-reset order-6 Butterworth filter pairs, Q16 ordering. */
typedef struct caribou_cuda_rxstream caribou_cuda_rxstream;
caribou_cuda_rxstream* caribou_cuda_rxstream_create(caribou_cuda_smi* dev, int channel);
void   caribou_cuda_rxstream_destroy(caribou_cuda_rxstream* s);
int    caribou_cuda_rxstream_set_format(caribou_cuda_rxstream* s, const char* fmt);  /* Stream::setFormat */
void   caribou_cuda_rxstream_set_filter(caribou_cuda_rxstream* s, int type);         /* Stream::setDigitalFilter: 0 none,1 20k,2 50k,3 100k */
size_t caribou_cuda_rxstream_mtu(caribou_cuda_rxstream* s);                          /* Stream::getMTUSizeElements */
int    caribou_cuda_rxstream_read(caribou_cuda_rxstream* s, void* buffer, size_t num_elements,
                                  long timeout_us);                                  /* Stream::ReadSamplesGen */

/* ---------------------------------------------------------------- TX direction
 * Samples -> 32-bit SMI TX words: [SOF TXC CTX I12..I8][0 I7..I1][0 I0 Q12..Q7][0 Q6..Q0], byte 0 first
 * (caribou_smi_generate_data, LC/caribou_smi/caribou_smi.c:688-717 - as committed the reference packs a constant
 * test pattern, ii = 0xFFFF / qq = 0, with the sample read commented out; this is that loop reading the samples).
 * in_format is a caribou_cuda_format: CS16 is the native pair; CS8 (<< 5), CF32 and CF64 ((int16_t)(x * 4096))
 * are converted first as Stream::WriteSamples does (LC/soapy_api/CaribouliteStream.cpp:199-244).  `words`
 * receives n_samples uint32.  Device buffers must be 16-byte aligned.  caribou_cuda_wait returns n_samples. */
int caribou_cuda_tx_pack_submit(caribou_cuda_ctx* ctx, int in_format, const void* samples, size_t n_samples,
                                int mem, void* words);
/* int caribou_smi_write(caribou_smi_st*, caribou_smi_channel_en, caribou_smi_sample_complex_int16*, size_t)
 *                                                                   LC/caribou_smi/caribou_smi.c:720-762 */
int caribou_cuda_smi_write(caribou_cuda_smi* dev, int channel,
                           const caribou_cuda_sample_complex_int16* samples, size_t length_samples);
/* Stream::WriteSamplesGen (CaribouliteStream.cpp:246-258) for the format named like Stream::setFormat names it:
 * MTU clamp for the converted formats, failures reported as 0 elements. */
int caribou_cuda_txstream_write(caribou_cuda_smi* dev, int channel, const char* fmt, const void* buffer,
                                size_t num_elements);

/* ---------------------------------------------------------------- capture pump (file / pipe to file / pipe)
 * The capture loop of cariboulite_util (LC/cariboulite_util.c:340-361: read a block, fwrite ret*4 bytes of raw
 * little-endian CS16) with the sample path on the GPU and the wire format kept.  Reads blocks of `block_samples`
 * elements of `in_kind` (raw SMI words, or CS16 / CF32 samples of an earlier capture) from in_fd until EOF or
 * max_samples (<= 0: no limit), runs the context's path and writes the result to out_fd:
 *   context without filter, IN_WORDS: unpack -> raw CS16 (4 bytes per sample; metadata is dropped exactly as the
 *                                      reference drops it, cariboulite_util.c:349);
 *   context with filter:               filtered, decimated elements of cfg.out_format.
 * Host I/O overlaps the GPU: two page-locked input and output buffers, block k+1 is read and block k-1 written
 * while block k is on the device.  block_samples must be a multiple of the decimation factor (and should be a
 * multiple of chunk_bytes/4 for IN_WORDS so that chunk boundaries fall where the reference's reads put them);
 * a tail shorter than the decimation factor is not processed.  Single-stream contexts only.
 * Returns the number of input samples consumed, -1 on I/O / CUDA failure, -3 when a block lost word sync. */
int64_t caribou_cuda_pump(caribou_cuda_ctx* ctx, int in_kind, int in_fd, int out_fd, size_t block_samples,
                          int64_t max_samples);

/* ---------------------------------------------------------------- continuous streaming
 * The sample queue of the reference's async RX path and its reader thread.
 * caribou_cuda_ring = circular_buffer<T>(size, override_write, block_read), LC/datatypes/circular_buffer.h:16-92:
 * capacity rounded up to a power of two; put() drops the oldest elements when override_write is set and the
 * writer outruns the reader, otherwise takes what fits; get() waits up to timeout_us for `length` elements and
 * returns 0 without taking any when they did not arrive (block_read), or returns what is there (non-blocking);
 * data == NULL consumes without copying.  (A put longer than the capacity overruns the reference's buffer; here
 * it keeps the newest `capacity` elements.)  Elements are elem_bytes wide. */
typedef struct caribou_cuda_ring caribou_cuda_ring;
caribou_cuda_ring* caribou_cuda_ring_create(size_t capacity, size_t elem_bytes, int override_write, int block_read);
void     caribou_cuda_ring_destroy(caribou_cuda_ring* r);
size_t   caribou_cuda_ring_put(caribou_cuda_ring* r, const void* data, size_t length);
size_t   caribou_cuda_ring_get(caribou_cuda_ring* r, void* data, size_t length, long timeout_us);
void     caribou_cuda_ring_reset(caribou_cuda_ring* r);
size_t   caribou_cuda_ring_size(caribou_cuda_ring* r);
size_t   caribou_cuda_ring_capacity(caribou_cuda_ring* r);
uint64_t caribou_cuda_ring_dropped(caribou_cuda_ring* r);      /* elements overwritten before they were read */
/* ReaderThread + rx_queue of SoapySDR::Stream (LC/soapy_api/CaribouliteStream.cpp:8,16-49,71,94-113,263): a thread
 * reads one MTU at a time through caribou_cuda_smi_read (the GPU path) and puts the samples into a ring of
 * capacity_samples (0: 10 MTUs as the reference); caribou_cuda_async_read is rx_queue->get.  The stream starts
 * inactive (activateStream).  caribou_cuda_async_idle is 1 after a read that returned nothing (a capture file or
 * pipe that ran dry). */
typedef struct caribou_cuda_async caribou_cuda_async;
caribou_cuda_async* caribou_cuda_async_open(caribou_cuda_smi* dev, int channel, size_t capacity_samples);
void   caribou_cuda_async_activate(caribou_cuda_async* q, int active);
size_t caribou_cuda_async_read(caribou_cuda_async* q, caribou_cuda_sample_complex_int16* buffer, size_t num_samples,
                               long timeout_us);
int    caribou_cuda_async_idle(caribou_cuda_async* q);
caribou_cuda_ring* caribou_cuda_async_ring(caribou_cuda_async* q);
void   caribou_cuda_async_close(caribou_cuda_async* q);

#ifdef __cplusplus
}
#endif
#endif /* CARIBOU_CUDA_H */
