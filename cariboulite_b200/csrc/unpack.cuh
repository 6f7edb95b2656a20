// Kernel (a): alignment search + fused unpack / demux / sync-check / sign-extend / CS16 / CF32.
#pragma once
#include "common.cuh"

namespace cbl {

struct OffsetsParams {
    SubView  sv;                  // sub-stream view of stream 0; stream s starts stream_stride_bytes*s later
    uint64_t stream_stride_bytes;
    uint32_t n_streams;           // gridDim.y
    uint32_t chunk_bytes;
    uint32_t n_chunks;            // gridDim.x
    int32_t* offs;          // [n_streams][offs_stride] out: byte offset, or -1
    uint32_t offs_stride;
    int32_t* first_bad;     // [n_streams] (stride first_bad_stride) out: min chunk index with offs == -1 (pre-set to INT_MAX)
    uint32_t first_bad_stride;
    int32_t* n_misaligned;  // out: += chunks with offs > 0
    int32_t* any_misaligned;// out: set to 1 if any chunk has offs != 0 (shared by all subs/streams)
    // per-stream verdicts for the fused path (all optional, indexed by the stream of the launch):
    int32_t* gate;          // out: set to 1 when a chunk of the stream has offs != 0 - sticky, the fused kernels skip the stream
    int32_t* gate_seq;      // out: min over the submits that set the gate of `seq`
    int32_t* nmis;          // out: += chunks of the stream with offs > 0
    int32_t  seq;           // sequence number of this submit since the context last resolved its queue
    int32_t* lost;          // optional out: set to 1 when a chunk has no alignment at all (sticky: the caller clears it)
};

struct UnpackParams {
    const uint32_t* words;        // raw capture (device)
    uint64_t nbytes;              // raw bytes
    uint32_t B;                   // interleave block (0, 1 or a multiple of 8)
    uint32_t chunk_bytes;
    uint32_t spc;                 // samples per chunk
    int      n_subs;
    SubView  sv[2];
    int      layout[2];           // decode layout per sub: 0 = S1G (i = hi13), 1 = HiF (i = lo13)
    const int32_t* offs[2];
    const int32_t* first_bad[2];
    uint32_t* cs16[2];
    uint8_t*  meta[2];
    float2*   cf32[2];
    int       vec_ok;             // all output pointers allow 16-byte (meta: 4-byte) vector stores
    // flag compaction (A10)
    int       want_flags;
    uint32_t* inv_idx;  uint64_t inv_cap;
    uint32_t* syn_idx;  uint64_t syn_cap;
    uint64_t  index_base;         // added to the word index (stream offset in a batch)
    uint64_t* tile_counts;        // [n_tiles] flagged words per tile (invalid<<32 | sync), zeroed
    uint64_t* tile_offsets;       // [n_tiles] exclusive scan of tile_counts
    uint32_t* work;               // [n_tiles] tiles with flags
    uint32_t* n_work;             // entries in work
    const uint64_t* totals_in;    // [2] (n_invalid, n_sync) of the earlier launches of this submit: read by every tile
    uint64_t* totals_out;         // [2] totals_in + this launch, written by the last tile (a different slot)
    uint32_t  n_tiles;
};

constexpr int kUnpackThreads = 256;
constexpr int kUnpackMaxIters  = 4;                                        // 4-word groups per thread and tile
constexpr int kUnpackTileWords = kUnpackThreads * kUnpackMaxIters * 4;    // raw words per tile (4096)

void launch_find_offsets(const OffsetsParams& p, cudaStream_t st);
// returns the number of kernels launched (decode, flag scan, flag index emission)
int launch_unpack(const UnpackParams& p, int num_sms, cudaStream_t st);
void launch_synth(uint32_t* out, uint64_t n0, uint64_t n, int hif_layout, uint64_t seed,
                  const int16_t* tone_i, const int16_t* tone_q, int num_sms, cudaStream_t st);

}  // namespace cbl
