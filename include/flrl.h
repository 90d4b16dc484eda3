/*
 * flrl.h — C ABI of the B200-native fixed-length (FL) / run-length (RL) codec.
 *
 * This is the drop-in boundary for the hot path of Polyphemus980/fl-rl-compression-MPI.
 * The reference has no FFI layer; its boundary is the `compress` CLI plus the C++ functions
 * of namespace FixedLength (/root/reference/src/fl/fl_gpu.cuh:13-17, fl_cpu.cuh:9-10).
 * Every entry point below names the reference interface it replaces.  All signatures are
 * plain C: pointers, sizes, an opaque stream handle (a cudaStream_t).  No torch types.
 *
 * Two layers:
 *   1. DEVICE layer  (flrl_fl_encode, ...): device pointers in, device pointers out, no
 *      allocation, no synchronisation, re-entrant per stream/device.  This is what the
 *      host drivers, the multi-GPU orchestrator, bench.py and the parity tests call.
 *   2. HOST layer    (flrl_fl_gpu_compress, ...): host buffers in, malloc-owned host buffers
 *      out, same argument meaning / ownership / empty-input behaviour as the reference's
 *      gpuCompress / gpuDecompress, with pinned double-buffered staging inside.
 *
 * Return value: 0 on success; > 0 a cudaError_t; < 0 one of FLRL_E_*.  flrl_last_error()
 * gives the message of the last failure on the calling thread.  There is NO CPU fallback:
 * without a CUDA device every GPU entry point fails with the CUDA error.
 *
 * On-disk formats (shared with the CLI, bit-exact with the reference):
 *   FL  u64 inputSize, u64 bitsSize, u64 valuesSize, bits[bitsSize], values[valuesSize]
 *       (/root/reference/src/file_io.cu:222-280)
 *   RL  u64 inputSize, u64 pairCount, counts[pairCount], values[pairCount]
 *       (build decision: the reference documents RL in IMPLEMENTATION-PLAN.md:81-179 but
 *       ships no RL code or container)
 */
#ifndef FLRL_H
#define FLRL_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FLRL_FRAME_LENGTH 128 /* src/fl/fl_common.cuh:9 */

#define FLRL_OK 0
#define FLRL_E_ALIGN (-1)     /* a device pointer is not 16-byte aligned */
#define FLRL_E_WORKSPACE (-2) /* workspace too small */
#define FLRL_E_CAPACITY (-3)  /* output buffer too small */
#define FLRL_E_FORMAT (-4)    /* malformed compressed input (sizes / widths inconsistent) */
#define FLRL_E_ARG (-5)       /* null pointer / inconsistent sizes */
#define FLRL_E_NOMEM (-6)     /* host allocation failed */

typedef void* flrl_stream_t; /* cudaStream_t; NULL = default stream */

const char* flrl_last_error(void);
int flrl_version(void);

/* ------------------------------------------------------------------ DEVICE layer: FL ---- */

/* ceil(n / 128): number of frames == bytes of the widths array (fl_cpu.cu:21). */
size_t flrl_fl_frames(size_t n);
/* Scratch for one encode or decode call over n uncompressed bytes (per 16 KiB tile: the scanner's 12 bytes for the
 * encoder, a look-back descriptor for the decoder).  One workspace per call in flight: calls that overlap on different
 * streams need their own. */
size_t flrl_fl_workspace_bytes(size_t n);

/*
 * Replaces compressCalculateOutputBits + compressInitializeFrameStartIndiciesBits +
 * thrust::exclusive_scan + compressCalculateOutput (fl_gpu.cu:648-726, 805-808), i.e. the
 * device part of gpuCompressDevice (fl_gpu.cu:425-535), in one pass.
 *   d_in[n]            input bytes, 16-byte aligned
 *   d_bits[frames]     out: width of each frame, 1..8
 *   d_values           out: packed values, 16-byte aligned, capacity values_cap >= round_up(n,16)
 *   d_values_size      out: number of value bytes (device u64), written by the kernel
 * n == 0 is a no-op that stores 0.  Asynchronous on `stream`.
 */
int flrl_fl_encode(const uint8_t* d_in, size_t n, uint8_t* d_bits, uint8_t* d_values, size_t values_cap,
                   uint64_t* d_values_size, void* d_ws, size_t ws_bytes, flrl_stream_t stream);

/*
 * Replaces the frame-start kernels + decompressCalculateOutput (fl_gpu.cu:687-698, 728-755,
 * 805-808), the device part of gpuDecompress (fl_gpu.cu:537-645).
 *   d_bits[nbits]      nbits must equal ceil(n/128)
 *   d_values[nvalues]  16-byte aligned
 *   d_out[n]           out, 16-byte aligned
 *   d_status           out (device int, may be NULL): 0 ok, else the container was malformed
 *                      (width outside 1..8 or values shorter than the widths imply)
 */
int flrl_fl_decode(const uint8_t* d_bits, size_t nbits, const uint8_t* d_values, size_t nvalues, uint8_t* d_out, size_t n,
                   int* d_status, void* d_ws, size_t ws_bytes, flrl_stream_t stream);

/*
 * flrl_fl_decode with NO allocation anywhere on its path and no workspace argument.  It has no counterpart in the
 * reference; it was added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  The decoder's scratch is a file-scope static
 * __device__ array inside the library, sized for n <= FLRL_STATIC_MAX_BYTES (1 GiB of output per call): 1.0 MiB that
 * every GPU which loads the library holds until the process exits.  Larger n is rejected with FLRL_E_CAPACITY.  Same
 * arguments, result and status as flrl_fl_decode; it calls neither cudaMalloc nor cudaFree nor any allocating routine.  All calls on
 * one GPU share the scratch (what must start at zero is cleared on `stream` by every call): calls that OVERLAP on one
 * GPU race on it — serialise them (one stream), or use flrl_fl_decode with one workspace per call in flight.
 */
#define FLRL_STATIC_MAX_BYTES (1ull << 30)
int flrl_fl_decode_static(const uint8_t* d_bits, size_t nbits, const uint8_t* d_values, size_t nvalues, uint8_t* d_out, size_t n,
                          int* d_status, flrl_stream_t stream);

/* ------------------------------------------------------------------ DEVICE layer: RL ---- */

size_t flrl_rl_workspace_bytes(size_t n);

/*
 * Run-length encode (IMPLEMENTATION-PLAN.md:83-152): a pair starts where a byte differs from
 * its predecessor and every 255 bytes inside a longer run (:125); counts in 1..255.
 *   d_in[n_avail]      the shard is d_in[0..n); bytes [n, n_avail) are look-ahead only, so a
 *                      run that STARTS in the shard gets its full count (n_avail - n <= 255 is
 *                      enough; n_avail == n for a whole stream)
 *   prev_run           number of bytes equal to d_in[0] immediately before the shard (0 for
 *                      a whole stream): position 0 opens a pair only if prev_run % 255 == 0
 *   d_counts,d_values  out, capacity pair_cap each (n is always enough)
 *   d_pairs            out: number of pairs (device u64)
 * A pair belongs to the shard in which it starts, so shard outputs concatenate to the
 * single-stream output byte for byte.
 */
int flrl_rl_encode(const uint8_t* d_in, size_t n, size_t n_avail, uint64_t prev_run, uint8_t* d_counts, uint8_t* d_values,
                   size_t pair_cap, uint64_t* d_pairs, void* d_ws, size_t ws_bytes, flrl_stream_t stream);

size_t flrl_rl_decode_workspace_bytes(size_t pairs);
/*
 * Run-length decode (IMPLEMENTATION-PLAN.md:154-179): out = values[k] repeated counts[k]
 * times; length = sum(counts) (:171).
 *   d_out_size         out: decoded length (device u64)
 *   d_status           out (device int, may be NULL): non-zero if sum(counts) > out_cap
 *                      (nothing is written past out_cap) or a count is 0
 */
int flrl_rl_decode(const uint8_t* d_counts, const uint8_t* d_values, size_t pairs, uint8_t* d_out, size_t out_cap,
                   uint64_t* d_out_size, int* d_status, void* d_ws, size_t ws_bytes, flrl_stream_t stream);

/* -------------------------------------------------------------------------- HOST layer -- */
/*
 * Host buffers in, malloc-owned host buffers out (release with flrl_free, i.e. free(), as
 * the reference's callers do: src/main.cu:126-128).  Inputs are borrowed for the call.
 * Empty input -> all outputs NULL / 0 (fl_gpu.cu:291-294, :539-543).
 * Staging: pinned, double-buffered cudaMemcpyAsync overlapped with the kernels.
 * `ngpus` <= 0 means "all visible devices"; 1 = single GPU.
 */

/* FixedLength::gpuCompress (fl_gpu.cu:289-423); with ngpus > 1 also gpuNCCLCompress
 * (fl_gpu.cu:76-287): 128-aligned contiguous shards (file_io.cu:46-51), one per GPU, NCCL
 * all-gather of the shard value sizes only, peer-copy concatenation. */
int flrl_fl_gpu_compress(const uint8_t* data, size_t size, int ngpus, uint8_t** bits, size_t* bits_size, uint8_t** values,
                         size_t* values_size);
/* FixedLength::gpuDecompress (fl_gpu.cu:537-645). */
int flrl_fl_gpu_decompress(size_t output_size, const uint8_t* bits, size_t bits_size, const uint8_t* values,
                           size_t values_size, int ngpus, uint8_t** out, size_t* out_size);
/* RL counterparts (no reference code; CLI methods `rl` per README.md:25). */
int flrl_rl_gpu_compress(const uint8_t* data, size_t size, int ngpus, uint8_t** counts, uint8_t** values, size_t* pairs);
int flrl_rl_gpu_decompress(const uint8_t* counts, const uint8_t* values, size_t pairs, size_t expected_size, int ngpus,
                           uint8_t** out, size_t* out_size);

/*
 * Caller-provided outputs (extension; no reference counterpart): same behaviour, but results are written into
 * buffers the caller owns.  When those — and the inputs — come from flrl_host_alloc (pinned), every
 * cudaMemcpyAsync goes straight to / from them and the bounce copy is skipped; this is what the CLI's file
 * staging and bench.py's end-to-end leg use.  FLRL_E_CAPACITY if an output does not fit.
 */
int flrl_fl_gpu_compress_into(const uint8_t* data, size_t size, int ngpus, uint8_t* bits, size_t bits_cap, uint8_t* values,
                              size_t values_cap, size_t* bits_size, size_t* values_size);
int flrl_fl_gpu_decompress_into(size_t output_size, const uint8_t* bits, size_t bits_size, const uint8_t* values,
                                size_t values_size, int ngpus, uint8_t* out, size_t out_cap, size_t* out_size);
int flrl_rl_gpu_compress_into(const uint8_t* data, size_t size, int ngpus, uint8_t* counts, uint8_t* values, size_t pair_cap,
                              size_t* pairs);
int flrl_rl_gpu_decompress_into(const uint8_t* counts, const uint8_t* values, size_t pairs, int ngpus, uint8_t* out,
                                size_t out_cap, size_t* out_size);
/* One PART of a longer stream (what a caller that streams a file through bounded buffers needs; the CLI does): encodes
 * data[0 .. size), whose counts may look into the `lookahead` bytes behind it (up to 255 are used), given that
 * `prev_run` bytes equal to data[0] come immediately before data[0].  Part outputs concatenate to the stream's output
 * (the shard semantics of flrl_rl_encode, through the host layer). */
int flrl_rl_gpu_compress_part_into(const uint8_t* data, size_t size, size_t lookahead, uint64_t prev_run, int ngpus, uint8_t* counts,
                                   uint8_t* values, size_t pair_cap, size_t* pairs);

/* ngpus > 1: how shard payloads are concatenated after the NCCL all-gather of their sizes.
 * 0 (default): every GPU copies its payload straight to its offset in the host buffer (8 PCIe links in parallel);
 * 1: every GPU peer-copies (cudaMemcpyPeerAsync over NVLink) into GPU 0's buffer, which is then copied to the
 *    host once — the device-resident merge of gpuNCCLCompress (fl_gpu.cu:221-238, :280). */
void flrl_set_concat_mode(int peer);
/* Frees the per-device caches (buffers, streams, events, pinned staging, the chunk pipeline's too) the host layer keeps
 * between calls. */
void flrl_release(void);
/* Optional: pays the one-time costs of the GPUs a later call with the same `ngpus` will use — contexts, streams, the
 * kernels' modules, and for several GPUs the NCCL communicators (the largest item) — outside the first codec call. */
int flrl_prepare(int ngpus);
/* Chunk size of the host layer's copy / kernel / copy pipeline (default 64 MiB; 0 restores it; rounded to 4 KiB, at
 * least 64 KiB).  A test hook: outputs do not depend on it. */
void flrl_set_pipe_chunk(size_t bytes);

/* CPU methods of the CLI (`fl-cpu`, `rl-cpu`): FixedLength::cpuCompress / cpuDecompress
 * (fl_cpu.cu:9-147) and the RL spec.  These ARE CPU methods by definition; the GPU entry
 * points above never route through them. */
int flrl_fl_cpu_compress(const uint8_t* data, size_t size, uint8_t** bits, size_t* bits_size, uint8_t** values,
                         size_t* values_size);
int flrl_fl_cpu_decompress(size_t output_size, const uint8_t* bits, size_t bits_size, const uint8_t* values,
                           size_t values_size, uint8_t** out, size_t* out_size);
int flrl_rl_cpu_compress(const uint8_t* data, size_t size, uint8_t** counts, uint8_t** values, size_t* pairs);
int flrl_rl_cpu_decompress(const uint8_t* counts, const uint8_t* values, size_t pairs, uint8_t** out, size_t* out_size);

void flrl_free(void* p);
/* Pinned host buffers (cudaHostAlloc): inputs/outputs placed here skip the staging copy. */
void* flrl_host_alloc(size_t bytes);
void flrl_host_free(void* p);

/* Device timing of the last host-layer call on this thread (cudaEvent, ms): kernels only,
 * and whole call including copies.  Mirrors the reference's GpuTimer (src/timers/gpu_timer.cu). */
void flrl_last_timing(float* kernel_ms, float* total_ms);
/* The same per GPU of the last call (g = 0 .. flrl_last_gpu_count() - 1), with the bytes that GPU moved over PCIe
 * (host -> device, device -> host): the reference's per-rank transfer lines (src/timers/cpu_timer_with_transfer.cu:16-40). */
int flrl_last_gpu_count(void);
int flrl_last_timing_gpu(int g, float* kernel_ms, float* total_ms, uint64_t* bytes_in, uint64_t* bytes_out);

#ifdef __cplusplus
}
#endif
#endif /* FLRL_H */
