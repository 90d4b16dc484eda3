// rl_kernels.cuh — launchers of the run-length kernels (rl_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace flrl {

size_t rl_encode_tiles(size_t n);
size_t rl_encode_workspace_bytes(size_t n);
size_t rl_decode_tiles(size_t pairs);
size_t rl_decode_workspace_bytes(size_t pairs);

// d_in, d_counts, d_values 16-byte aligned.  See include/flrl.h for the shard semantics.
cudaError_t rl_encode_launch(const uint8_t* d_in, size_t n, size_t n_avail, uint64_t prev_run, uint8_t* d_counts,
                             uint8_t* d_values, size_t pair_cap, uint64_t* d_pairs, void* d_ws, cudaStream_t stream);
cudaError_t rl_decode_launch(const uint8_t* d_counts, const uint8_t* d_values, size_t pairs, uint8_t* d_out, size_t out_cap,
                             uint64_t* d_out_size, void* d_ws, cudaStream_t stream);
const int* rl_decode_status(const void* d_ws, size_t pairs);
// *d_sum = sum of the n bytes at d_p (16-byte aligned)
cudaError_t sum_bytes_launch(const uint8_t* d_p, size_t n, uint64_t* d_sum, cudaStream_t stream);

// d_sums[i] = sum of d_p[i*seg .. min((i+1)*seg, n)) for i < nseg = ceil(n / seg); d_sums[nseg] = 1 if a byte is 0, | 2 if a
// byte is > 8 (FL widths must be 1..8, RL counts >= 1).  d_p 16-byte aligned, seg a multiple of 16; d_sums holds nseg + 1 words.
cudaError_t segment_sums_launch(const uint8_t* d_p, size_t n, size_t seg, uint64_t* d_sums, cudaStream_t stream);

}  // namespace flrl
