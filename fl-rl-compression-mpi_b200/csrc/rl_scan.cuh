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
    // the arrays are padded to a multiple of 256 entries (rl_encode_tiles), so whole chunks can be read and written;
    // entries at or past `tiles` are never posted and count as 0
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
        const bool real_a = t0 + 4 <= tiles, real_b = t0 + 8 <= tiles;  // groups that hold only real tiles
        while (true) {
            // a group that reaches past the last tile is checked entry by entry
            bool ok = true;
            if (real_a) ok = all_posted(cur.a);
            else {
                if (t0 + 0 < tiles) ok = ok && (cur.a.x >> 31);
                if (t0 + 1 < tiles) ok = ok && (cur.a.y >> 31);
                if (t0 + 2 < tiles) ok = ok && (cur.a.z >> 31);
            }
            bool okb = true;
            if (real_b) okb = all_posted(cur.b);
            else {
                if (t0 + 4 < tiles) okb = okb && (cur.b.x >> 31);
                if (t0 + 5 < tiles) okb = okb && (cur.b.y >> 31);
                if (t0 + 6 < tiles) okb = okb && (cur.b.z >> 31);
            }
            if (!__any_sync(0xffffffffu, !(ok && okb))) break;
            if (!ok) cur.a = ld_relaxed_u32x4(agg + t0);
            if (!okb) cur.b = ld_relaxed_u32x4(agg + t0 + 4);
        }
        uint32_t v[8] = {cur.a.x, cur.a.y, cur.a.z, cur.a.w, cur.b.x, cur.b.y, cur.b.z, cur.b.w};
        uint32_t ex[8], run = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            ex[j] = run;
            run += (t0 + j < tiles) ? (v[j] & 0x7fffffffu) : 0u;
        }
        uint32_t s = run;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, s, d);
            if (lane >= d) s += o;
        }
        const uint64_t lb = base + (s - run);
#pragma unroll
        for (int j = 0; j < 8; j += 2)
            st_relaxed_u64x2(inc + t0 + j, (1ull << 63) | (lb + ex[j]), (1ull << 63) | (lb + ex[j + 1]));
        base += __shfl_sync(0xffffffffu, s, 31);
        cur = nxt;
    }
}

}  // namespace flrl
