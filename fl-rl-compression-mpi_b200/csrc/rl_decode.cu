// rl_decode.cu — run-length decoder for sm_100a: one launch, counts and values read once, output written once.
//
// The reference documents RL decode as "prescan of the counts + binary search" (IMPLEMENTATION-PLAN.md:154-179)
// and ships no code.  Structure here:
//
//   * persistent CTAs (grid = what fits on the device) take tiles of 4096 pairs (fewer when that would leave a CTA with too few tiles) from an atomic counter;
//   * a service warp per CTA runs AHEAD of the compute warps through a ring of tile slots: tile id, TMA bulk
//     loads of the tile's counts and values, sum of the counts (= the tile's output bytes; a count of 0 is
//     malformed and flagged) posted to the grid's scanner warp (rl_scan.cuh), which answers with the tile's first
//     output byte.  By the time the compute warps reach the tile, that answer is normally there: no tile
//     waits for a look-back;
//   * 8 compute warps: block scan of the counts -> exclusive starts in shared memory (":164-171") + a table
//     "pair that covers output byte b << sh" -> every thread owns 64-byte units aligned in the OUTPUT
//     address space, finds the pair covering the unit's first byte by a binary search (":172-179") that the table narrows to the
//     pairs of one block (0-1 steps instead of 12), walks the pairs of the unit and XORs value deltas
//     into 64-bit words -> 256-bit stores (STG.256: whole 32-byte sectors);
//   * tiles of SHORT runs (mean count < 12): that walk diverges — every 16-byte piece costs the maximum number of pairs any
//     lane has in its piece — so they are expanded from a bit mask of the output positions at which a pair starts (built
//     with shared-memory atomicOr where offs[] would be): per 4 output bytes one 16-entry table look-up (PRMT selector +
//     number of starts behind the 4 flags), one unaligned 4-byte read of the values and two PRMTs, no divergence
//     (random bytes: 6.6 -> 2.3 ms per GiB; mean 4: 0.98 -> 0.91 ms).  tools/model_rld.py is the same arithmetic on the CPU.
#include <stdlib.h>

#include <algorithm>

#include "flrl_bits.cuh"
#include "rl_kernels.cuh"
#include "rl_scan.cuh"

namespace flrl {

#ifndef RLD_SLOTS_N
#define RLD_SLOTS_N 3
#endif
#ifndef RLD_CTAS
#define RLD_CTAS 4
#endif
#ifndef RLD_SHORT_MEAN
#define RLD_SHORT_MEAN 12  // tiles whose mean count is below this take the start-bit expand (0: never)
#endif
static constexpr int RLD_WARPS = 8;                        // compute warps
static constexpr int RLD_THREADS = RLD_WARPS * 32;
static constexpr int RLD_ALL = RLD_THREADS + 32;            // + service warp
static constexpr int RLD_PER = 16;                          // pairs per thread
static constexpr int RLD_PAIRS = RLD_THREADS * RLD_PER;     // 4096 pairs per tile (<= 1,044,480 output bytes)
static constexpr int RLD_SLOTS = RLD_SLOTS_N;               // tiles the service warp may be ahead
static constexpr int RLD_CTAS_PER_SM = RLD_CTAS;
static constexpr uint32_t BAR_D = 1;                        // named barrier of the compute warps

// high bit of each byte of the result is set iff that byte of x is non-zero
__device__ __forceinline__ uint32_t nz_bytes_d(uint32_t x) { return (((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x) & 0x80808080u; }
__device__ __forceinline__ uint32_t byte_sum4_d(uint32_t w)
{
    const uint32_t h = (w & 0x00ff00ffu) + ((w >> 8) & 0x00ff00ffu);
    return (h & 0xffffu) + (h >> 16);
}

struct RldSlotMeta {
    long long tile;     // -1: no more tiles
    uint32_t np;        // pairs in the tile
    uint32_t total;     // output bytes of the tile
};

__global__ void __launch_bounds__(RLD_ALL, RLD_CTAS_PER_SM)
rl_decode_kernel(const uint8_t* __restrict__ counts, const uint8_t* __restrict__ values, size_t pairs, uint8_t* __restrict__ out,
                 size_t out_cap, uint64_t* __restrict__ d_out_size, uint64_t* __restrict__ inc, uint32_t* __restrict__ agg,
                 uint32_t* __restrict__ next_tile, int* __restrict__ err, long long tiles, uint32_t tile_pairs)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* const slots = smem;                                                      // RLD_SLOTS x [counts 4096][values 4096]
    uint32_t* const offs = reinterpret_cast<uint32_t*>(smem + RLD_SLOTS * 2 * RLD_PAIRS);  // exclusive starts; offs[np] = total
    uint16_t* const tbl = reinterpret_cast<uint16_t*>(offs + RLD_PAIRS + 4);          // pair covering output byte 256 b
    __shared__ __align__(8) uint64_t loaded[RLD_SLOTS], full[RLD_SLOTS], slot_free[RLD_SLOTS];
    __shared__ RldSlotMeta meta[RLD_SLOTS];
    __shared__ uint32_t wsum[RLD_WARPS];
    __shared__ uint64_t obase_s[2];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // the first CTA that RUNS is the grid's scanner (role counter next to the tile counter): nothing depends on which
    // CTA the hardware places first, nor on the whole grid being resident at once
    __shared__ uint32_t scanner_role;
    if (tid == 0) scanner_role = atomicAdd(next_tile + 2, 1u);
    __syncthreads();
    if (scanner_role == 0) {
        if (warp == 0) scan_pair_counts(agg, inc, tiles, lane);
        return;
    }
    if (tid == 0) {
        for (int s = 0; s < RLD_SLOTS; ++s) {
            mbar_init(&loaded[s], 1);
            mbar_init(&full[s], 1);
            mbar_init(&slot_free[s], RLD_WARPS);
        }
        fence_mbar_init();
    }
    // short-run expand: for the four start flags (b1 b2 b3 b4) behind a byte, the PRMT selector that spreads up to four
    // consecutive values over the next 4 output bytes (byte i takes value b1 + .. + bi) | number of flags << 16
    __shared__ uint32_t ent[16];
    if (tid < 16) {
        const uint32_t b1 = tid & 1u, b2 = (tid >> 1) & 1u, b3 = (tid >> 2) & 1u;
        ent[tid] = (b1 << 4) | ((b1 + b2) << 8) | ((b1 + b2 + b3) << 12) | ((uint32_t)__popc(tid) << 16);
    }
    __syncthreads();

    if (warp == RLD_WARPS) {
        // ================================ service warp ===========================================
        for (uint32_t q = 0;; ++q) {
            const int s = (int)(q % RLD_SLOTS);
            const uint32_t use = q / RLD_SLOTS;
            if (use) mbar_wait(&slot_free[s], (use - 1) & 1u);
            long long tile = 0;
            if (lane == 0) tile = atomicAdd(next_tile, 1u);
            tile = __shfl_sync(0xffffffffu, tile, 0);
            if (tile >= tiles) {
                if (lane == 0) {
                    meta[s].tile = -1;
                    mbar_arrive(&full[s]);
                }
                break;
            }
            const size_t p0 = (size_t)tile * tile_pairs;  // tile_pairs <= RLD_PAIRS, a multiple of 16
            const uint32_t np = (uint32_t)(pairs - p0 < (size_t)tile_pairs ? pairs - p0 : tile_pairs);
            const uint32_t np16 = np & ~15u;
            uint8_t* cnt = slots + s * 2 * RLD_PAIRS;
            uint8_t* val = cnt + RLD_PAIRS;
            if (lane < 16 && np16 + lane < np) {  // the ragged end (< 16 pairs): plain loads
                cnt[np16 + lane] = counts[p0 + np16 + lane];
                val[np16 + lane] = values[p0 + np16 + lane];
            }
            __syncwarp();
            if (lane == 0) {
                if (np16) {
                    mbar_expect_tx(&loaded[s], 2 * np16);
                    tma_load_1d(cnt, counts + p0, np16, &loaded[s]);
                    tma_load_1d(val, values + p0, np16, &loaded[s]);
                } else {
                    mbar_arrive(&loaded[s]);
                }
            }
            mbar_wait(&loaded[s], use & 1u);
            // the tile's output bytes: every lane sums 128 counts
            uint32_t sum = 0;
            bool zero = false;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int at = (i * 32 + lane) * 16;  // 16 counts per load
                uint4 v = *reinterpret_cast<const uint4*>(cnt + at);
                uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int valid = (int)np - (at + 4 * k);
                    const uint32_t keep = valid >= 4 ? ~0u : (valid <= 0 ? 0u : ((1u << (8 * valid)) - 1u));
                    w[k] &= keep;
                    zero = zero || ((~nz_bytes_d(w[k]) & 0x80808080u & keep) != 0);  // a count of 0 is malformed
                    sum += byte_sum4_d(w[k]);
                }
            }
            if (zero) atomicExch(err, 2);
            sum = __reduce_add_sync(0xffffffffu, sum);
            if (lane == 0) {
                st_relaxed_u32(agg + tile, 0x80000000u | sum);  // to the scanner
                meta[s].tile = tile;
                meta[s].np = np;
                meta[s].total = sum;
                mbar_arrive(&full[s]);
            }
        }
        return;
    }

    // ================================= compute warps =======================================
    long long pref_tile = -1;  // thread 0: tile whose scanner answer was asked for ahead of time, and the answer (or 0)
    uint64_t g_pref = 0;
    for (uint32_t k = 0;; ++k) {
        const int s = (int)(k % RLD_SLOTS);
        mbar_wait(&full[s], (k / RLD_SLOTS) & 1u);
        const long long tile = *reinterpret_cast<const volatile long long*>(&meta[s].tile);
        if (tile < 0) break;
        const uint32_t np = meta[s].np, total = meta[s].total;
        const uint8_t* cnt = slots + s * 2 * RLD_PAIRS;
        const uint8_t* val = cnt + RLD_PAIRS;
        // block table: tbl[b] = pair that covers output byte b << sh of the tile; the block size is the smallest power of
        // two >= 16 for which the table has at most 4096 entries, so a block holds about one pair on long runs
        // and is a single 16-byte chunk on short ones (no search at all)
        const uint32_t sh = total <= 16u * RLD_PAIRS ? 4u : 32u - (uint32_t)__clz((total - 1u) >> 12);

        // ---- scan of the counts: 16 per thread
        uint32_t c[4];
        {
            const uint4 q = *reinterpret_cast<const uint4*>(cnt + tid * RLD_PER);
            c[0] = q.x; c[1] = q.y; c[2] = q.z; c[3] = q.w;
        }
        const int valid = (int)np - tid * RLD_PER;  // pairs of this thread that exist
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int v = valid - 4 * j;
            c[j] &= v >= 4 ? ~0u : (v <= 0 ? 0u : ((1u << (8 * v)) - 1u));
        }
        const uint32_t mine = byte_sum4_d(c[0]) + byte_sum4_d(c[1]) + byte_sum4_d(c[2]) + byte_sum4_d(c[3]);
        uint32_t isum = mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, isum, d);
            if (lane >= d) isum += o;
        }
        if (lane == 31) wsum[warp] = isum;
        if (tid == 0) {
            // the tile's first output byte: asked for by the service warp a few tiles ago; normally already fetched while
            // the previous tile was expanded (pref_*), so that its L2 round trip is not on every warp's path to barrier A
            uint64_t g = 1ull << 63;
            if (tile > 0) {
                g = tile == pref_tile ? g_pref : 0ull;
                while ((g >> 63) == 0) {
                    g = ld_relaxed_u64(inc + tile);
                    if ((g >> 63) == 0) __nanosleep(40);
                }
            }
            g &= ~(1ull << 63);
            obase_s[k & 1u] = g;
            if (tile == tiles - 1) *d_out_size = g + total;
            if (g + total > out_cap) atomicExch(err, 1);
        }
        bar_sync(BAR_D, RLD_THREADS);  // A: wsum, obase; everybody is done with the previous tile's offs / tbl
        // Short runs (mean count < RLD_SHORT_MEAN; CTA-uniform): the walk over pairs below diverges (every 16-byte piece costs
        // the MAXIMUM number of pairs any lane has in its piece), so such tiles are expanded from a bit mask of the output
        // positions at which a pair starts (it takes the place of offs[]): constant work per output byte, no divergence.
        const bool short_runs = RLD_SHORT_MEAN && total < (uint32_t)RLD_SHORT_MEAN * np;  // => total < 16 * RLD_PAIRS, sh == 4
        uint32_t* const bm = offs;  // bit x: a pair starts at output byte x of the tile (x >= 1); zero up to total + 64 and beyond
        if (short_runs) {
            const uint32_t nw = ((total + 64u) >> 5) + 3u;
            for (uint32_t i = tid; i < nw; i += RLD_THREADS) bm[i] = 0u;
            bar_sync(BAR_D, RLD_THREADS);
        }
        uint32_t run = isum - mine;
#pragma unroll
        for (int j = 0; j < RLD_WARPS; ++j)
            if (j < warp) run += wsum[j];
        {
            uint32_t o[16];
            uint32_t bw = run >> 5, bits = 0;  // short runs: start bits of the word being filled
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                o[j] = run;
                const uint32_t cj = (c[j >> 2] >> (8 * (j & 3))) & 0xffu;
                for (uint32_t b = (run + (1u << sh) - 1u) >> sh; (b << sh) < run + cj; ++b) tbl[b] = (uint16_t)(tid * RLD_PER + j);
                if (short_runs && cj && run) {  // position 0 is not marked: the pair covering byte 0 is pair 0
                    if ((run >> 5) != bw) {
                        if (bits) atomicOr(&bm[bw], bits);
                        bw = run >> 5;
                        bits = 0;
                    }
                    bits |= 1u << (run & 31u);
                }
                run += cj;
            }
            if (short_runs) {
                if (bits) atomicOr(&bm[bw], bits);
            } else {
                uint4* dst = reinterpret_cast<uint4*>(offs + tid * RLD_PER);
                dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
                dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
                dst[2] = make_uint4(o[8], o[9], o[10], o[11]);
                dst[3] = make_uint4(o[12], o[13], o[14], o[15]);
                if (tid == RLD_THREADS - 1) offs[RLD_PAIRS] = run;
            }
        }
        bar_sync(BAR_D, RLD_THREADS);  // B: offs, tbl
        if (tid == 0) {  // the next tile's first output byte, if the service warp has that tile ready and the scanner has answered
            const int s2 = (int)((k + 1u) % RLD_SLOTS);
            pref_tile = -1;
            if (mbar_try_wait(&full[s2], ((k + 1u) / RLD_SLOTS) & 1u)) {
                const long long t2 = *reinterpret_cast<const volatile long long*>(&meta[s2].tile);
                if (t2 > 0) {
                    pref_tile = t2;
                    g_pref = ld_relaxed_u64(inc + t2);
                }
            }
        }

        // ---- expand: every thread owns 64-byte units (two 32-byte stores, 32-byte aligned in the OUTPUT address space):
        // one search per unit, then a walk over the pairs of its four 16-byte pieces
        const uint64_t O = obase_s[k & 1u];
        const uint64_t lim = out_cap > O ? (out_cap - O < total ? out_cap - O : total) : 0;  // bytes of this tile that fit
        const uint32_t tot = (uint32_t)lim;
        const uint32_t mis = (uint32_t)((reinterpret_cast<uintptr_t>(out) + O) & 31u);
        const uint32_t nunits = tot ? (mis + tot + 63u) >> 6 : 0u;
        if (short_runs) {
            // from the pair that covers the unit's first byte on, every 4 output
            // bytes cost one table entry (selector + number of starts), one unaligned 4-byte read of the values and two PRMTs
            // (tools/model_rld.py is the same arithmetic on the CPU)
            for (uint32_t q = tid; q < nunits; q += RLD_THREADS) {
                const int p = (int)(64u * q) - (int)mis;  // tile-relative output position of the unit's first byte (< 0: first unit only)
                // pair covering s0 = tbl[s0 >> 4] + start bits in ((s0 & ~15) + 1 .. s0].  Measured slower: a binary search over the
                // threads' first offsets instead of tbl[] (eight dependent shared-memory reads per unit: 0.99 ms per GiB on mean 4
                // against 0.91), and a table per 64-byte unit instead of per 16 bytes (1.01 ms).
                const uint32_t s0 = p < 0 ? 0u : (uint32_t)p;
                const uint32_t blk = s0 >> 4, B = blk << 4;
                const uint32_t wB = (B & 16u) ? __funnelshift_r(bm[B >> 5], bm[(B >> 5) + 1u], 16) : bm[B >> 5];
                uint32_t cur = (uint32_t)tbl[blk] + (uint32_t)__popc(wB & ((2u << (s0 - B)) - 2u));
                uint32_t lo, hi;  // start bits of positions p + 1 .. p + 64
                if (p >= 0) {
                    const uint32_t w0 = (uint32_t)(p + 1) >> 5, sft = (uint32_t)(p + 1) & 31u;
                    const uint32_t a = bm[w0], b = bm[w0 + 1u], c2 = bm[w0 + 2u];
                    lo = __funnelshift_r(a, b, sft);
                    hi = __funnelshift_r(b, c2, sft);
                } else {
                    const uint32_t kk = (uint32_t)(-(p + 1));  // 0 .. 30: positions p + 1 .. 0 hold no start
                    const uint32_t a = bm[0], b = bm[1];
                    lo = a << kk;
                    hi = __funnelshift_l(a, b, kk);
                }
                uint32_t w[16];
#pragma unroll
                for (int wq = 0; wq < 16; ++wq) {
                    const uint32_t e = ent[((wq < 8 ? lo : hi) >> (4 * (wq & 7))) & 15u];
                    const uint32_t* vp = reinterpret_cast<const uint32_t*>(val + (cur & ~3u));
                    const uint32_t v4 = __byte_perm(vp[0], vp[1], 0x3210u + 0x1111u * (cur & 3u));  // values cur .. cur + 3
                    w[wq] = __byte_perm(v4, 0u, e & 0xffffu);
                    cur += e >> 16;
                }
                uint8_t* dst = out + O + p;
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    const int ps = p + 32 * half;
                    if (ps >= 0 && (uint32_t)ps + 32u <= tot) {
                        asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst + 32 * half), "r"(w[8 * half]),
                                     "r"(w[8 * half + 1]), "r"(w[8 * half + 2]), "r"(w[8 * half + 3]), "r"(w[8 * half + 4]),
                                     "r"(w[8 * half + 5]), "r"(w[8 * half + 6]), "r"(w[8 * half + 7])
                                     : "memory");
                    } else {
#pragma unroll
                        for (int bi = 0; bi < 32; ++bi) {
                            const int pos = ps + bi;
                            if (pos >= 0 && (uint32_t)pos < tot) dst[32 * half + bi] = (uint8_t)(w[8 * half + (bi >> 2)] >> (8 * (bi & 3)));
                        }
                    }
                }
            }
        } else
        for (uint32_t q = tid; q < nunits; q += RLD_THREADS) {
            const int p = (int)(64u * q) - (int)mis;  // tile-relative output position of the unit's first byte
            const uint32_t lo = p < 0 ? 0u : (uint32_t)p;
            // j = last pair with offs[j] <= lo  (IMPLEMENTATION-PLAN.md:164-179, binary search)
            uint32_t a = 0, b = np;  // invariant: offs[a] <= lo, offs[b] > lo (offs[np] = total > lo)
            {
                const uint32_t blk = lo >> sh;
                a = tbl[blk];
                if (((blk + 1u) << sh) < total) b = (uint32_t)tbl[blk + 1u] + 1u;
            }
            while (b - a > 1) {
                const uint32_t mid = (a + b) >> 1;
                if (offs[mid] <= lo) a = mid; else b = mid;
            }
            uint32_t j = a;
            uint32_t pv = val[j];
            uint32_t end = offs[j + 1];
            uint8_t* dst = out + O + p;
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                uint64_t w[4];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int ph = p + 32 * half + 16 * h;  // first byte of the 16-byte piece
                    const uint32_t hi = ph + 16 <= 0 ? 0u : ((uint32_t)(ph + 16) < tot ? (uint32_t)(ph + 16) : tot);
                    const uint32_t v4 = pv * 0x01010101u;
                    uint64_t x_lo = ((uint64_t)v4 << 32) | v4, x_hi = x_lo;
                    while (end < hi) {
                        ++j;
                        const uint32_t nv = val[j];
                        const uint32_t d4 = (nv ^ pv) * 0x01010101u;
                        const uint64_t d8 = ((uint64_t)d4 << 32) | d4;
                        const uint32_t at = end - (uint32_t)ph;  // 0..15: first byte of the piece owned by pair j
                        x_lo ^= shl64c(d8, 8 * at);
                        x_hi ^= at < 8 ? d8 : shl64c(d8, 8 * at - 64);
                        pv = nv;
                        end = offs[j + 1];
                    }
                    w[2 * h] = x_lo;
                    w[2 * h + 1] = x_hi;
                }
                const int ps = p + 32 * half;
                if (ps >= 0 && (uint32_t)ps + 32u <= tot) {
                    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(dst + 32 * half), "r"((uint32_t)w[0]),
                                 "r"((uint32_t)(w[0] >> 32)), "r"((uint32_t)w[1]), "r"((uint32_t)(w[1] >> 32)), "r"((uint32_t)w[2]),
                                 "r"((uint32_t)(w[2] >> 32)), "r"((uint32_t)w[3]), "r"((uint32_t)(w[3] >> 32))
                                 : "memory");
                } else {
                    for (int bi = 0; bi < 32; ++bi) {
                        const int pos = ps + bi;
                        if (pos >= 0 && (uint32_t)pos < tot) dst[32 * half + bi] = (uint8_t)(w[bi >> 3] >> (8 * (bi & 7)));
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&slot_free[s]);  // this warp is done with the slot's counts / values
    }
}

// =============================================================================================
//                                        launchers
// =============================================================================================
// A tile is 4096 pairs unless that leaves the persistent CTAs with fewer than ~6 tiles each (long runs: a tile of 4096
// pairs is up to 1 MiB of output, and 1024 such tiles over 591 CTAs end 14 % late): then 2048 or 1024.
static constexpr int RLD_MIN_PAIRS = 1024;
static uint32_t rl_decode_tile_pairs(size_t pairs, size_t workers)
{
    uint32_t tp = RLD_PAIRS;
    while (tp > RLD_MIN_PAIRS && pairs / tp < 6 * workers) tp /= 2;
    return tp;
}
// sized for the smallest tile, padded to whole scanner chunks
size_t rl_decode_tiles(size_t pairs) { return ((pairs + RLD_MIN_PAIRS - 1) / RLD_MIN_PAIRS + 255) / 256 * 256; }
// workspace: inc[tiles] u64 | agg[tiles] u32 | tile counter u32 | status int
size_t rl_decode_workspace_bytes(size_t pairs) { return rl_decode_tiles(pairs) * 12 + 16; }
const int* rl_decode_status(const void* d_ws, size_t pairs)
{
    return reinterpret_cast<const int*>(reinterpret_cast<const uint8_t*>(d_ws) + rl_decode_tiles(pairs) * 12 + 4);
}

cudaError_t rl_decode_launch(const uint8_t* d_counts, const uint8_t* d_values, size_t pairs, uint8_t* d_out, size_t out_cap,
                             uint64_t* d_out_size, void* d_ws, cudaStream_t stream)
{
    const size_t tmax = rl_decode_tiles(pairs);
    uint64_t* inc = reinterpret_cast<uint64_t*>(d_ws);
    uint32_t* agg = reinterpret_cast<uint32_t*>(inc + tmax);
    uint32_t* next_tile = agg + tmax;
    int* err = reinterpret_cast<int*>(next_tile + 1);
    cudaError_t e = cudaMemsetAsync(d_ws, 0, rl_decode_workspace_bytes(pairs), stream);
    if (e != cudaSuccess) return e;
    constexpr int smem = RLD_SLOTS * 2 * RLD_PAIRS + (RLD_PAIRS + 4) * 4 + (RLD_PAIRS + 8) * 2;
    // tiles are taken from a counter by CTAs that run, and the first CTA that runs is the scanner: no co-residency is required
    int dev = 0;
    e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    static int resident[64];
    int cap = dev < 64 ? resident[dev] : 0;
    if (!cap) {
        e = cudaFuncSetAttribute((const void*)rl_decode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        int sms = 0, per_sm = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, (const void*)rl_decode_kernel, RLD_ALL, smem);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) return cudaErrorLaunchOutOfResources;
        cap = sms * std::max(1, std::min(per_sm, RLD_CTAS_PER_SM));
        if (cap < 2) return cudaErrorLaunchOutOfResources;
        if (dev < 64) resident[dev] = cap;
    }
    const uint32_t tp = rl_decode_tile_pairs(pairs, (size_t)cap - 1);
    const size_t tiles = (pairs + tp - 1) / tp;
    const unsigned grid = (unsigned)std::min<size_t>(tiles, (size_t)cap - 1) + 1;  // the scanner CTA + the workers
    rl_decode_kernel<<<grid, RLD_ALL, smem, stream>>>(d_counts, d_values, pairs, d_out, out_cap, d_out_size, inc, agg, next_tile, err,
                                                      (long long)tiles, tp);
    return cudaGetLastError();
}

}  // namespace flrl
