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
// frame array is ever materialised.
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

// Called by one full warp.  Returns the sum of the aggregates of tiles [0, tile).
__device__ __forceinline__ uint64_t lookback_exclusive_sum(const uint64_t* desc, long long tile, int lane)
{
    uint64_t excl = 0;
    long long j = tile - 1 - lane;
    while (true) {
        uint64_t v = LB_INC;  // tiles before 0: inclusive prefix 0
        if (j >= 0) v = ld_relaxed_u64(desc + j);
        while (__any_sync(0xffffffffu, (v >> 62) == 0)) {
            if ((v >> 62) == 0) v = ld_relaxed_u64(desc + j);
        }
        const unsigned inc = __ballot_sync(0xffffffffu, (v >> 62) == 2);
        uint64_t val = v & LB_MASK;
        if (inc) {
            const int first = __ffs(inc) - 1;  // nearest predecessor holding an inclusive prefix
            if (lane > first) val = 0;
            excl += warp_sum_u64(val);
            return excl;
        }
        excl += warp_sum_u64(val);
        j -= 32;
    }
}

}  // namespace flrl
