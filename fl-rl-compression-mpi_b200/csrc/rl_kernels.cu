// rl_kernels.cu — run-length codec for sm_100a.
//
// The reference documents RL (IMPLEMENTATION-PLAN.md:81-179: mask -> scan -> start indices ->
// split of runs > 255 -> emit; decode = prescan + binary search) but ships no code for it.
// Both directions are ONE pass here:
//
//   encode: TMA bulk load of an 8 KiB tile (+16 B left halo, +256 B look-ahead) -> run-start
//           bit mask per 32-byte lane segment (SIMD-in-word compares) -> block max-scan for the
//           start S of the run that enters each segment (+ look-back #1 across tiles) -> the
//           single possible 255-split position of each segment, (i - S) % 255 == 0 -> block
//           sum-scan of pair counts (+ look-back #2 for the tile's first pair slot) -> compaction
//           of (count, value) bytes through shared memory -> 16-byte aligned coalesced stores.
//           Replaces the plan's iterative "rescan while any run > 255" loop (:125-147).
//   decode: TMA bulk load of a 4096-pair tile -> block scan of the counts (+ look-back for the
//           tile's output offset) -> every thread owns 16-byte aligned output chunks, binary-
//           searches the scanned offsets in shared memory for the pair covering the chunk's
//           first byte (:164-179) and builds the 16 bytes in registers -> 128-bit stores.
//
// A pair belongs to the tile (and the shard) in which it STARTS; its count looks ahead up to
// 255 bytes, so tiles and shards concatenate to the single-stream output byte for byte.
#include "rl_kernels.cuh"

#include <algorithm>

#include "flrl_bits.cuh"
#include "lookback.cuh"

namespace flrl {

// high bit of each byte of the result is set iff that byte of x is non-zero
__device__ __forceinline__ uint32_t nz_bytes(uint32_t x)
{
    return (((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x) & 0x80808080u;
}

static __device__ __forceinline__ uint32_t byte_sum4(uint32_t w)
{
    const uint32_t h = (w & 0x00ff00ffu) + ((w >> 8) & 0x00ff00ffu);
    return (h & 0xffffu) + (h >> 16);
}

// ---- sum of a byte array (shard offsets of the decoders: sum of widths / sum of counts) ------
__global__ void __launch_bounds__(256) sum_bytes_kernel(const uint8_t* __restrict__ p, size_t n, unsigned long long* __restrict__ out)
{
    unsigned long long acc = 0;
    const size_t n16 = n / 16;
    const uint4* q = reinterpret_cast<const uint4*>(p);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 v = q[i];
        acc += byte_sum4(v.x) + byte_sum4(v.y) + byte_sum4(v.z) + byte_sum4(v.w);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 15)) acc += p[n16 * 16 + threadIdx.x];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if ((threadIdx.x & 31) == 0 && acc) atomicAdd(out, acc);
}

cudaError_t sum_bytes_launch(const uint8_t* d_p, size_t n, uint64_t* d_sum, cudaStream_t stream)
{
    cudaError_t e = cudaMemsetAsync(d_sum, 0, sizeof(uint64_t), stream);
    if (e != cudaSuccess || n == 0) return e;
    const unsigned blocks = (unsigned)std::min<size_t>((n / 16 + 255) / 256 + 1, 148 * 8);
    sum_bytes_kernel<<<blocks, 256, 0, stream>>>(d_p, n, reinterpret_cast<unsigned long long*>(d_sum));
    return cudaGetLastError();
}

}  // namespace flrl
