// rl_scan.cuh — the grid-wide prefix engine of the run-length kernels (encoder: pairs per tile -> first pair slot;
// decoder: bytes per tile -> first output byte).
#pragma once
#include "flrl_ptx.cuh"

namespace flrl {

// The scanner (one warp of the grid's first CTA): exclusive prefix sum of the tiles' posted totals, in tile order.
// agg[t] = 0x80000000 | total of tile t (0 = not yet posted); inc[t] = 1 << 63 | sum of the totals of tiles 0 .. t-1.
// A lane owns 8 consecutive tiles (two 16-byte loads, a serial prefix, eight 8-byte results written as 16-byte
// stores), the warp 256; the next 256 counts are already in flight while the current ones are
// summed, and only the loads that still held an empty entry are repeated.  It must run several times faster
// than the tiles arrive: whenever it has waited for a straggler, everything behind that tile waits with it.
struct ScanRegs {
    uint4 a, b;
};
__device__ __forceinline__ uint4 ld_relaxed_u32x4(const uint32_t* p)
{
    uint4 v;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64x2(uint64_t* p, uint64_t x, uint64_t y)
{
    asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(x), "l"(y) : "memory");
}
__device__ __forceinline__ bool all_posted(const uint4& v) { return ((v.x & v.y & v.z & v.w) >> 31) != 0; }
__device__ __forceinline__ void scan_pair_counts(const uint32_t* __restrict__ agg, uint64_t* __restrict__ inc, long long tiles, int lane)
{
    // the arrays are padded to a multiple of 256 entries, so whole chunks can be read and written; entries at or past
    // `tiles` are never posted and count as 0
    uint64_t base = 0;
    const long long chunks = (tiles + 255) / 256;
    ScanRegs cur, nxt;
    cur.a = ld_relaxed_u32x4(agg + 8 * lane);
    cur.b = ld_relaxed_u32x4(agg + 8 * lane + 4);
    for (long long c = 0; c < chunks; ++c) {
        const long long t0 = c * 256 + 8 * lane;
        if (c + 1 < chunks) {
            nxt.a = ld_relaxed_u32x4(agg + t0 + 256);
            nxt.b = ld_relaxed_u32x4(agg + t0 + 260);
        }
        // An answer must never wait for a LATER tile (the CTA that holds that tile may itself be waiting for this
        // answer): `front` = entries of the chunk already answered; every round answers all entries up to the first
        // one that is still empty.  Normally the whole chunk is there at the first look (one round, vector stores).
        uint32_t front = 0;
        while (true) {
            uint32_t v[8] = {cur.a.x, cur.a.y, cur.a.z, cur.a.w, cur.b.x, cur.b.y, cur.b.z, cur.b.w};
            uint32_t lv = 0;  // leading entries of this lane that are posted (entries past the last tile count as posted)
            bool open = true;
            uint32_t ex[8], run = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const bool real = t0 + j < tiles;
                open = open && (!real || (v[j] >> 31));
                lv += open ? 1u : 0u;
                ex[j] = run;
                run += (open && real) ? (v[j] & 0x7fffffffu) : 0u;
            }
            const unsigned fullm = __ballot_sync(0xffffffffu, lv == 8);
            const uint32_t r = fullm == 0xffffffffu ? 32u : (uint32_t)__ffs(~fullm) - 1u;  // leading lanes that are complete
            const uint32_t lv_r = __shfl_sync(0xffffffffu, lv, r & 31u);
            const uint32_t new_front = r == 32u ? 256u : 8u * r + lv_r;
            // exclusive prefix over lanes 0 .. r (lane r with its leading posted entries only)
            uint32_t s = (uint32_t)lane <= r ? run : 0u;
            const uint32_t mine = s;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t o = __shfl_up_sync(0xffffffffu, s, d);
                if (lane >= d) s += o;
            }
            const uint64_t lb = base + (s - mine);
            if (front == 0 && new_front == 256u) {
#pragma unroll
                for (int j = 0; j < 8; j += 2)
                    st_relaxed_u64x2(inc + t0 + j, (1ull << 63) | (lb + ex[j]), (1ull << 63) | (lb + ex[j + 1]));
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const uint32_t pos = 8u * lane + j;
                    if (pos >= front && pos < new_front) st_relaxed_u64(inc + t0 + j, (1ull << 63) | (lb + ex[j]));
                }
            }
            front = new_front;
            if (front == 256u) {
                base += __shfl_sync(0xffffffffu, s, 31);
                break;
            }
            if (lv < 8) {  // poll the groups that still hold an empty entry
                if (lv < 4) cur.a = ld_relaxed_u32x4(agg + t0);
                cur.b = ld_relaxed_u32x4(agg + t0 + 4);
            }
        }
        cur = nxt;
    }
}

}  // namespace flrl
