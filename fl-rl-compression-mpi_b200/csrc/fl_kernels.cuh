// fl_kernels.cuh — launchers of the fixed-length kernels (fl_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace flrl {

size_t fl_tiles(size_t n);
size_t fl_workspace_bytes(size_t n);

// d_in, d_values 16-byte aligned; d_values capacity >= round_up(n, 16); d_bits >= ceil(n/128).
cudaError_t fl_encode_launch(const uint8_t* d_in, size_t n, uint8_t* d_bits, uint8_t* d_values, uint64_t* d_values_size,
                             void* d_ws, cudaStream_t stream);
// nframes == ceil(n/128).  The int at fl_decode_status(d_ws, n) is 0 on success after the stream drains.
cudaError_t fl_decode_launch(const uint8_t* d_bits, size_t nframes, const uint8_t* d_values, size_t nvalues, uint8_t* d_out,
                             size_t n, void* d_ws, cudaStream_t stream);
const int* fl_decode_status(const void* d_ws, size_t n);

}  // namespace flrl
