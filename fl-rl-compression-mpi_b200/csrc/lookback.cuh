// lookback.cuh — single-pass decoupled look-back over per-tile aggregates (Merrill & Garland).
//
// One 64-bit descriptor per tile: bits 63:62 = state (0 invalid, 1 aggregate, 2 inclusive
// prefix), bits 61:0 = value.  State and value share one word, so a relaxed gpu-scope
// store/load pair is enough — no fences.  The descriptor array must be zeroed before launch.
// Tiles are taken in blockIdx order (or from an in-order counter), so every predecessor of
// a running tile has already started: the wait below cannot deadlock.
//
// Replaces thrust::exclusive_scan over uint64 frame starts
// (/root/reference/src/fl/fl_gpu.cu:805-808 and the init kernel :687-698): no 8-byte-per-
// frame array is ever materialised.  Users: fl_decode.cu (one look-back per block of the offsets
// pre-pass) and rl_encode.cu (start of the run that enters a tile).
#pragma once
#include "flrl_ptx.cuh"

namespace flrl {

static constexpr uint64_t LB_AGG = 1ull << 62;
static constexpr uint64_t LB_INC = 2ull << 62;
static constexpr uint64_t LB_MASK = (1ull << 62) - 1;

__device__ __forceinline__ uint64_t warp_sum_u64(uint64_t v)
{
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

}  // namespace flrl
