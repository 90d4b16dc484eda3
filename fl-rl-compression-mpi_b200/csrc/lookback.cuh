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
//
// Why it is wide: every tile between the newest inclusive prefix and a tile that starts its
// walk is itself still walking, so the steady-state tile rate is bounded by
// (tiles inspected per round) / (time per round).  A 32-wide, one-row-at-a-time walk
// measured 6.5 us per tile on B200 (1 GiB FL encode, ~1000 tiles resident) and capped the
// whole kernel.  Here a round inspects 32*LB_W predecessors: LB_W coalesced 256-byte loads
// are in flight together (one L2 round trip) and the whole round is evaluated with 2*LB_W
// votes, one REDUX and one 64-bit shuffle.  Only lanes that saw an invalid descriptor
// re-poll, so the spin does not flood L2.
// Aggregates must be < 2^24 (the round's partial sum is kept in 32 bits).
static constexpr int LB_W = 8;

__device__ __forceinline__ uint64_t lookback_exclusive_sum(const uint64_t* desc, long long tile, int lane, unsigned sleep_ns = 0)
{
    uint64_t excl = 0;
    long long j0 = tile - 1 - lane;  // row k covers j0 - 32k; nearest-first order is (k, lane)
    while (true) {
        uint64_t v[LB_W];
#pragma unroll
        for (int k = 0; k < LB_W; ++k) v[k] = (j0 - 32 * k >= 0) ? ld_relaxed_u64(desc + (j0 - 32 * k)) : LB_INC;
        int kstar;
        unsigned incm;
        while (true) {
            kstar = LB_W;
            incm = 0;
            bool bad = false;
#pragma unroll
            for (int k = 0; k < LB_W; ++k) {
                const unsigned i_k = __ballot_sync(0xffffffffu, (v[k] >> 62) == 2);
                const unsigned n_k = __ballot_sync(0xffffffffu, (v[k] >> 62) == 0);
                if (kstar == LB_W) {
                    if (i_k) {
                        kstar = k;
                        incm = i_k;
                        bad = bad || (n_k & ((2u << (__ffs(i_k) - 1)) - 1u)) != 0;
                    } else {
                        bad = bad || n_k != 0;
                    }
                }
            }
            if (!bad) break;
            if (sleep_ns) __nanosleep(sleep_ns);  // back off: hundreds of warps poll the same few lines
#pragma unroll
            for (int k = 0; k < LB_W; ++k)
                if (k <= kstar && (v[k] >> 62) == 0) v[k] = ld_relaxed_u64(desc + (j0 - 32 * k));
        }
        uint32_t small = 0;
#pragma unroll
        for (int k = 0; k < LB_W; ++k)
            if (k < kstar) small += (uint32_t)v[k];  // aggregates: state bits are above bit 31
        if (kstar < LB_W) {
            const int first = __ffs(incm) - 1;
            uint64_t vs = 0;
#pragma unroll
            for (int k = 0; k < LB_W; ++k)
                if (k == kstar) vs = v[k] & LB_MASK;
            if (lane < first) small += (uint32_t)vs;
            const uint64_t big = __shfl_sync(0xffffffffu, vs, first);
            return excl + big + __reduce_add_sync(0xffffffffu, small);
        }
        excl += __reduce_add_sync(0xffffffffu, small);
        j0 -= 32 * LB_W;
    }
}

}  // namespace flrl
