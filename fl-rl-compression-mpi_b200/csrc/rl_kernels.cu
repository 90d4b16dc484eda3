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

static constexpr int RL_THREADS = 256;
static constexpr int RL_SEG = 32;                       // input bytes per thread
static constexpr int RL_TILE = RL_THREADS * RL_SEG;     // 8 KiB
static constexpr int RL_HALO = 256;                     // look-ahead: a count never exceeds 255
static constexpr int RL_INBUF = 16 + RL_TILE + RL_HALO + 16;
static constexpr int RL_STAGE = RL_TILE + 32;           // compacted counts / values (+ over-read slack)
static constexpr uint64_t RL_BIAS = 255;                // run starts are stored as S + 255 >= 0

// ---- run-start flags of 8 consecutive words -------------------------------------------------
// bit (4k + j) of the result: byte j of word k differs from the byte before it.
// high bit of each byte of t is set iff that byte of x is non-zero; the multiply gathers the
// four (eight, for a word pair) byte flags into adjacent bits (no two partial products share
// a bit, so there are no carries).
__device__ __forceinline__ uint32_t nz_bytes(uint32_t x)
{
    return (((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x) & 0x80808080u;
}
__device__ __forceinline__ uint32_t start_mask32(const uint32_t (&w)[8], uint32_t prev_word)
{
    uint32_t m = 0, prev = prev_word;
#pragma unroll
    for (int k = 0; k < 8; k += 2) {
        const uint32_t a = nz_bytes(w[k] ^ __byte_perm(prev, w[k], 0x6543));
        const uint32_t b = nz_bytes(w[k + 1] ^ __byte_perm(w[k], w[k + 1], 0x6543));
        const uint32_t g = (a >> 7) | (b >> 3);  // bit 0 / bit 4 of every byte
        m |= ((g * 0x00204081u) >> 21 & 0xffu) << (4 * k);
        prev = w[k + 1];
    }
    return m;
}

// look-back #1: start (biased) of the run that is open at the end of tile `tile - 1`.
// d1 states: 0 not yet known, 1 "no run starts in this tile", 2 final value.
__device__ __forceinline__ uint64_t lookback_run_start(const uint64_t* d1, long long tile, int lane)
{
    long long j = tile - 1 - lane;
    while (true) {
        const uint64_t v = (j >= 0) ? ld_relaxed_u64(d1 + j) : LB_INC;
        const unsigned fin = __ballot_sync(0xffffffffu, (v >> 62) == 2);
        const unsigned inv = __ballot_sync(0xffffffffu, (v >> 62) == 0);
        if (fin) {
            const int f = __ffs(fin) - 1;
            if (inv & ((1u << f) - 1u)) continue;  // a nearer tile has not published yet
            return __shfl_sync(0xffffffffu, v & LB_MASK, f);
        }
        if (inv) continue;
        j -= 32;
    }
}

// 16 staged bytes starting at byte offset u (any alignment) of a 4-byte aligned shared buffer
__device__ __forceinline__ uint4 lds_unaligned16(const uint8_t* buf, uint32_t u)
{
    const uint32_t* w = reinterpret_cast<const uint32_t*>(buf) + (u >> 2);
    const uint32_t sel = 0x3210u + 0x1111u * (u & 3u);
    const uint32_t a = w[0], b = w[1], c = w[2], d = w[3], e = w[4];
    return make_uint4(__byte_perm(a, b, sel), __byte_perm(b, c, sel), __byte_perm(c, d, sel), __byte_perm(d, e, sel));
}

// cnt staged bytes -> dst[g0 .. g0 + cnt), coalesced 16-byte stores on 16-byte aligned global
// addresses, byte stores on the two ragged ends; nothing at or past index `cap` is written.
__device__ __forceinline__ void store_compacted(uint8_t* __restrict__ dst, uint64_t g0, uint32_t cnt, uint64_t cap,
                                                const uint8_t* stage, int tid)
{
    const uint32_t a = (uint32_t)((reinterpret_cast<uintptr_t>(dst) + g0) & 15u);
    const uint32_t nchunks = (a + cnt + 15u) >> 4;
    for (uint32_t q = tid; q < nchunks; q += RL_THREADS) {
        const int u = (int)(16u * q) - (int)a;  // staging index of the chunk's first byte
        if (u >= 0 && (uint32_t)u + 16u <= cnt && g0 + u + 16u <= cap) {
            *reinterpret_cast<uint4*>(dst + g0 + u) = lds_unaligned16(stage, (uint32_t)u);
        } else {
            const int lo = u < 0 ? 0 : u;
            const int hi = (uint32_t)u + 16u < cnt ? u + 16 : (int)cnt;
            for (int k = lo; k < hi; ++k)
                if (g0 + k < cap) dst[g0 + k] = stage[k];
        }
    }
}

// =============================================================================================
//                                         ENCODE
// =============================================================================================
struct RlEncShared {
    int wmax[8];        // per-warp last run start (tile relative), -1 none
    uint32_t wsum[8];   // per-warp pair count
    int wmin[8];        // per-warp first boundary position (tile relative)
    int sentinel;       // first run start in the look-ahead halo (tile relative), or RL_TILE + RL_HALO
    uint32_t d0;        // (tile base - start of the run entering the tile) % 255
    uint64_t gbase;     // first pair slot of the tile
};

__global__ void __launch_bounds__(RL_THREADS, 4)
rl_encode_kernel(const uint8_t* __restrict__ in, size_t n, size_t n_avail, uint32_t prev_phase, int has_prev,
                 uint8_t* __restrict__ counts, uint8_t* __restrict__ values, size_t pair_cap, uint64_t* __restrict__ d_pairs,
                 uint64_t* __restrict__ d1, uint64_t* __restrict__ d2)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* inbuf = smem;                       // [16 left halo][tile][halo][pad]; tile byte i at inbuf[16 + i]
    uint8_t* cnt8 = smem + RL_INBUF;             // compacted counts
    uint8_t* val8 = cnt8 + RL_STAGE;             // compacted values
    __shared__ __align__(8) uint64_t bar;
    __shared__ RlEncShared sh;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long tile = blockIdx.x;
    const size_t tbase = (size_t)tile * RL_TILE;
    const bool last_tile = tile == (long long)gridDim.x - 1;

    // ---- stage the tile: one bulk copy for the 16-byte multiple, scalar loads for the ragged end
    const size_t lo = tile > 0 ? tbase - 16 : tbase;
    const size_t want = tbase + RL_TILE + RL_HALO < n_avail ? tbase + RL_TILE + RL_HALO : n_avail;
    const size_t hi16 = want & ~(size_t)15;
    if (tid == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
        if (hi16 > lo) {
            mbar_expect_tx(&bar, (uint32_t)(hi16 - lo));
            tma_load_1d(inbuf + (tile > 0 ? 0 : 16), in + lo,
                        (uint32_t)(hi16 - lo), &bar);
        }
    }
    if (tid >= 32 && tid < 48) {
        const size_t g = (hi16 > lo ? hi16 : lo) + (tid - 32);
        if (g < want) inbuf[16 + (g - tbase)] = in[g];
    }
    if (tile == 0 && tid == 48) inbuf[15] = has_prev ? in[0] : (uint8_t)~in[0];
    __syncthreads();
    if (hi16 > lo) mbar_wait(&bar, 0);

    // ---- run-start mask of this thread's 32 bytes
    const int off = tid * RL_SEG;
    uint32_t w[8];
    {
        const uint4 a = *reinterpret_cast<const uint4*>(inbuf + 16 + off);
        const uint4 b = *reinterpret_cast<const uint4*>(inbuf + 32 + off);
        w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    }
    uint32_t m = start_mask32(w, *reinterpret_cast<const uint32_t*>(inbuf + 12 + off));
    const size_t gpos = tbase + off;
    // positions at or past the end of the available data close the last run; positions >= n are not emitted
    const uint32_t k_av = n_avail > gpos ? (n_avail - gpos < 32 ? (uint32_t)(n_avail - gpos) : 32u) : 0u;
    const uint32_t k_em = n > gpos ? (n - gpos < 32 ? (uint32_t)(n - gpos) : 32u) : 0u;
    if (k_av < 32) m |= ~0u << k_av;
    const uint32_t emit = k_em < 32 ? ((1u << k_em) - 1u) : ~0u;

    // warp 7 also finds the first run start inside the look-ahead halo
    if (warp == 7) {
        int first = RL_HALO;
        if (lane < RL_HALO / RL_SEG) {
            const int hoff = RL_TILE + lane * RL_SEG;
            uint32_t hw[8];
            const uint4 a = *reinterpret_cast<const uint4*>(inbuf + 16 + hoff);
            const uint4 b = *reinterpret_cast<const uint4*>(inbuf + 32 + hoff);
            hw[0] = a.x; hw[1] = a.y; hw[2] = a.z; hw[3] = a.w; hw[4] = b.x; hw[5] = b.y; hw[6] = b.z; hw[7] = b.w;
            uint32_t hm = start_mask32(hw, *reinterpret_cast<const uint32_t*>(inbuf + 12 + hoff));
            const size_t hg = tbase + hoff;
            const uint32_t hk = n_avail > hg ? (n_avail - hg < 32 ? (uint32_t)(n_avail - hg) : 32u) : 0u;
            if (hk < 32) hm |= ~0u << hk;
            if (hm) first = lane * RL_SEG + (__ffs(hm) - 1);
        }
        first = __reduce_min_sync(0xffffffffu, first);
        if (lane == 0) sh.sentinel = RL_TILE + first;
    }

    // ---- S: last run start before this segment (block-wide exclusive max-scan)
    int mine = m ? off + 31 - __clz(m) : -1;
    int inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc = max(inc, o);
    }
    int s_rel = __shfl_up_sync(0xffffffffu, inc, 1);
    if (lane == 0) s_rel = -1;
    if (lane == 31) sh.wmax[warp] = inc;
    __syncthreads();
    {
        int pm = -1;
#pragma unroll
        for (int k = 0; k < 7; ++k)
            if (k < warp) pm = max(pm, sh.wmax[k]);
        s_rel = max(s_rel, pm);
    }
    if (warp == 0) {
        // look-back #1: only tiles whose first byte continues an earlier run need it
        int tl = -1;
#pragma unroll
        for (int k = 0; k < 8; ++k) tl = max(tl, sh.wmax[k]);
        const bool need = (__shfl_sync(0xffffffffu, m, 0) & 1u) == 0;  // thread 0's first byte is not a run start
        if (tl >= 0 && lane == 0) st_relaxed_u64(d1 + tile, LB_INC | (uint64_t)(tbase + tl + RL_BIAS));
        if (need) {
            uint64_t sb;  // biased start of the entering run
            if (tile == 0) {
                sb = RL_BIAS - prev_phase;
            } else {
                if (tl < 0 && lane == 0) st_relaxed_u64(d1 + tile, LB_AGG);
                sb = lookback_run_start(d1, tile, lane);
            }
            if (lane == 0) {
                if (tl < 0) st_relaxed_u64(d1 + tile, LB_INC | sb);
                sh.d0 = (uint32_t)((tbase + RL_BIAS - sb) % 255u);
            }
        }
    }
    __syncthreads();

    // ---- the one possible 255-split inside the leading part of the segment
    const uint32_t lead = m ? (uint32_t)(__ffs(m) - 1) : 32u;  // bytes that continue the entering run
    uint32_t bm = m;                                           // pair boundaries: run starts + the split
    if (lead) {
        const uint32_t phase = s_rel >= 0 ? (uint32_t)(off - s_rel) % 255u : (sh.d0 + (uint32_t)off) % 255u;
        const uint32_t j = phase ? 255u - phase : 0u;
        if (j < lead) bm |= 1u << j;
    }
    const uint32_t e = bm & emit;
    const uint32_t c = __popc(e);

    // ---- pair slots (exclusive sum-scan) and the next boundary after this segment (suffix min)
    uint32_t isum = c;
    int nxt = bm ? off + (__ffs(bm) - 1) : 0x7fffffff;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, isum, d);
        const int q = __shfl_down_sync(0xffffffffu, nxt, d);
        if (lane >= d) isum += o;
        if (lane + d < 32) nxt = min(nxt, q);
    }
    if (lane == 31) sh.wsum[warp] = isum;
    if (lane == 0) sh.wmin[warp] = nxt;
    int nxt_after = __shfl_down_sync(0xffffffffu, nxt, 1);
    if (lane == 31) nxt_after = 0x7fffffff;
    __syncthreads();
    uint32_t slot = isum - c, total = 0;
    {
        nxt_after = min(nxt_after, sh.sentinel);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const uint32_t s = sh.wsum[k];
            if (k < warp) slot += s;
            if (k > warp) nxt_after = min(nxt_after, sh.wmin[k]);
            total += s;
        }
    }
    if (warp == 0) {
        // look-back #2: first pair slot of the tile
        uint64_t g = 0;
        if (tile > 0) {
            if (lane == 0) st_relaxed_u64(d2 + tile, LB_AGG | total);
            g = lookback_exclusive_sum(d2, tile, lane);
        }
        if (lane == 0) {
            st_relaxed_u64(d2 + tile, LB_INC | (g + total));
            sh.gbase = g;
            if (last_tile) *d_pairs = g + total;
        }
    }

    // ---- compaction: (count, value) of every pair that starts in this segment
    {
        uint32_t bits = e, k = slot;
        int prevb = -1;
        while (bits) {
            const int b = __ffs(bits) - 1;
            bits &= bits - 1;
            val8[k] = inbuf[16 + off + b];
            if (prevb >= 0) cnt8[k - 1] = (uint8_t)(b - prevb);
            prevb = b;
            ++k;
        }
        if (prevb >= 0) {
            // the segment's last pair ends at the next boundary (this segment's non-emitted ones included)
            const uint32_t later = bm & ~emit & ~((2u << prevb) - 1u);
            const int nb = later ? off + (__ffs(later) - 1) : nxt_after;
            const int len = nb - (off + prevb);
            cnt8[k - 1] = (uint8_t)(len < 255 ? len : 255);
        }
    }
    __syncthreads();
    const uint64_t g0 = sh.gbase;
    store_compacted(counts, g0, total, pair_cap, cnt8, tid);
    store_compacted(values, g0, total, pair_cap, val8, tid);
}

// =============================================================================================
//                                         DECODE
// =============================================================================================
static constexpr int RLD_THREADS = 256;
static constexpr int RLD_PER = 16;                          // pairs per thread
static constexpr int RLD_PAIRS = RLD_THREADS * RLD_PER;     // 4096 pairs per tile (<= 1,044,480 output bytes)

__device__ __forceinline__ uint32_t byte_sum4(uint32_t w)
{
    const uint32_t h = (w & 0x00ff00ffu) + ((w >> 8) & 0x00ff00ffu);
    return (h & 0xffffu) + (h >> 16);
}

__global__ void __launch_bounds__(RLD_THREADS, 4)
rl_decode_kernel(const uint8_t* __restrict__ counts, const uint8_t* __restrict__ values, size_t pairs, uint8_t* __restrict__ out,
                 size_t out_cap, uint64_t* __restrict__ d_out_size, uint64_t* __restrict__ desc, int* __restrict__ err)
{
    __shared__ __align__(128) uint8_t cnt[RLD_PAIRS];
    __shared__ __align__(128) uint8_t val[RLD_PAIRS];
    __shared__ __align__(16) uint32_t offs[RLD_PAIRS + 4];  // exclusive starts inside the tile; offs[np] = total
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t wsum[8];
    __shared__ uint64_t obase;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long tile = blockIdx.x;
    const size_t p0 = (size_t)tile * RLD_PAIRS;
    const uint32_t np = (uint32_t)(pairs - p0 < (size_t)RLD_PAIRS ? pairs - p0 : RLD_PAIRS);
    const uint32_t np16 = np & ~15u;
    const bool last_tile = tile == (long long)gridDim.x - 1;

    if (tid == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
        if (np16) {
            mbar_expect_tx(&bar, 2 * np16);
            tma_load_1d(cnt, counts + p0, np16, &bar);
            tma_load_1d(val, values + p0, np16, &bar);
        }
    }
    if (tid >= 32 && tid < 48 && np16 + (tid - 32) < np) {
        cnt[np16 + tid - 32] = counts[p0 + np16 + tid - 32];
        val[np16 + tid - 32] = values[p0 + np16 + tid - 32];
    }
    __syncthreads();
    if (np16) mbar_wait(&bar, 0);

    // ---- scan of the counts: 16 per thread
    uint32_t c[4];
    {
        const uint4 q = *reinterpret_cast<const uint4*>(cnt + tid * RLD_PER);
        c[0] = q.x; c[1] = q.y; c[2] = q.z; c[3] = q.w;
    }
    const int valid = (int)np - tid * RLD_PER;  // pairs of this thread that exist
    bool zero = false;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int v = valid - 4 * k;
        const uint32_t keep = v >= 4 ? ~0u : (v <= 0 ? 0u : ((1u << (8 * v)) - 1u));
        c[k] &= keep;
        zero = zero || ((~nz_bytes(c[k]) & 0x80808080u & keep) != 0);  // a count of 0 is malformed
    }
    if (zero) atomicExch(err, 2);
    const uint32_t mine = byte_sum4(c[0]) + byte_sum4(c[1]) + byte_sum4(c[2]) + byte_sum4(c[3]);
    uint32_t inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    uint32_t run = inc - mine, total = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const uint32_t s = wsum[k];
        if (k < warp) run += s;
        total += s;
    }
    {
        uint32_t o[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            o[k] = run;
            run += (c[k >> 2] >> (8 * (k & 3))) & 0xffu;
        }
        uint4* dst = reinterpret_cast<uint4*>(offs + tid * RLD_PER);
        dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
        dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
        dst[2] = make_uint4(o[8], o[9], o[10], o[11]);
        dst[3] = make_uint4(o[12], o[13], o[14], o[15]);
        if (tid == RLD_THREADS - 1) offs[RLD_PAIRS] = run;
    }
    if (warp == 0) {
        uint64_t g = 0;
        if (tile > 0) {
            if (lane == 0) st_relaxed_u64(desc + tile, LB_AGG | total);
            g = lookback_exclusive_sum(desc, tile, lane);
        }
        if (lane == 0) {
            st_relaxed_u64(desc + tile, LB_INC | (g + total));
            obase = g;
            if (last_tile) *d_out_size = g + total;
            if (g + total > out_cap) atomicExch(err, 1);
        }
    }
    __syncthreads();

    // ---- expand: 16-byte chunks aligned in the OUTPUT address space
    const uint64_t O = obase;
    const uint64_t lim = out_cap > O ? (out_cap - O < total ? out_cap - O : total) : 0;  // bytes of this tile that fit
    if (lim == 0) return;
    const uint32_t tot = (uint32_t)lim;
    const uint32_t mis = (uint32_t)((reinterpret_cast<uintptr_t>(out) + O) & 15u);
    const uint32_t nchunks = (mis + tot + 15u) >> 4;
    uint32_t jlo = 0;
    for (uint32_t q = tid; q < nchunks; q += RLD_THREADS) {
        const int p = (int)(16u * q) - (int)mis;  // tile-relative output position of the chunk's first byte
        const uint32_t lo = p < 0 ? 0u : (uint32_t)p;
        const uint32_t hi = (uint32_t)(p + 16) < tot ? (uint32_t)(p + 16) : tot;
        // j = last pair with offs[j] <= lo  (IMPLEMENTATION-PLAN.md:164-179, binary search)
        uint32_t a = jlo, b = np;  // invariant: offs[a] <= lo, offs[b] > lo (offs[np] = total > lo)
        while (b - a > 1) {
            const uint32_t mid = (a + b) >> 1;
            if (offs[mid] <= lo) a = mid; else b = mid;
        }
        uint32_t j = a;
        jlo = j;
        uint32_t pv = val[j];
        uint32_t end = offs[j + 1];
        uint64_t x_lo, x_hi;
        {
            const uint32_t v4 = pv * 0x01010101u;
            x_lo = ((uint64_t)v4 << 32) | v4;
            x_hi = x_lo;
        }
        while (end < hi) {
            ++j;
            const uint32_t nv = val[j];
            const uint32_t d4 = (nv ^ pv) * 0x01010101u;
            const uint64_t d8 = ((uint64_t)d4 << 32) | d4;
            const uint32_t at = end - (uint32_t)p;  // 1..15: first byte of the chunk owned by pair j
            x_lo ^= shl64c(d8, 8 * at);
            x_hi ^= at < 8 ? d8 : shl64c(d8, 8 * at - 64);
            pv = nv;
            end = offs[j + 1];
        }
        uint8_t* dst = out + O + p;
        if (p >= 0 && (uint32_t)p + 16u <= tot) {
            *reinterpret_cast<uint4*>(dst) = make_uint4((uint32_t)x_lo, (uint32_t)(x_lo >> 32), (uint32_t)x_hi, (uint32_t)(x_hi >> 32));
        } else {
            for (uint32_t k = lo; k < hi; ++k) {
                const uint32_t bi = k - (uint32_t)p;
                dst[bi] = (uint8_t)((bi < 8 ? x_lo >> (8 * bi) : x_hi >> (8 * bi - 64)) & 0xff);
            }
        }
    }
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

// =============================================================================================
//                                        launchers
// =============================================================================================
size_t rl_encode_tiles(size_t n) { return (n + RL_TILE - 1) / RL_TILE; }
size_t rl_encode_workspace_bytes(size_t n) { return (2 * rl_encode_tiles(n) + 2) * sizeof(uint64_t); }
size_t rl_decode_tiles(size_t pairs) { return (pairs + RLD_PAIRS - 1) / RLD_PAIRS; }
size_t rl_decode_workspace_bytes(size_t pairs) { return (rl_decode_tiles(pairs) + 2) * sizeof(uint64_t); }
const int* rl_decode_status(const void* d_ws, size_t pairs)
{
    return reinterpret_cast<const int*>(reinterpret_cast<const uint64_t*>(d_ws) + rl_decode_tiles(pairs));
}

cudaError_t rl_encode_launch(const uint8_t* d_in, size_t n, size_t n_avail, uint64_t prev_run, uint8_t* d_counts,
                             uint8_t* d_values, size_t pair_cap, uint64_t* d_pairs, void* d_ws, cudaStream_t stream)
{
    const size_t tiles = rl_encode_tiles(n);
    uint64_t* d1 = reinterpret_cast<uint64_t*>(d_ws);
    uint64_t* d2 = d1 + tiles;
    cudaError_t e = cudaMemsetAsync(d1, 0, 2 * tiles * sizeof(uint64_t), stream);
    if (e != cudaSuccess) return e;
    constexpr int smem = RL_INBUF + 2 * RL_STAGE;
    rl_encode_kernel<<<(unsigned)tiles, RL_THREADS, smem, stream>>>(d_in, n, n_avail, (uint32_t)(prev_run % 255u), prev_run != 0,
                                                                   d_counts, d_values, pair_cap, d_pairs, d1, d2);
    return cudaGetLastError();
}

cudaError_t rl_decode_launch(const uint8_t* d_counts, const uint8_t* d_values, size_t pairs, uint8_t* d_out, size_t out_cap,
                             uint64_t* d_out_size, void* d_ws, cudaStream_t stream)
{
    const size_t tiles = rl_decode_tiles(pairs);
    uint64_t* desc = reinterpret_cast<uint64_t*>(d_ws);
    int* err = reinterpret_cast<int*>(desc + tiles);
    cudaError_t e = cudaMemsetAsync(desc, 0, (tiles + 1) * sizeof(uint64_t), stream);
    if (e != cudaSuccess) return e;
    rl_decode_kernel<<<(unsigned)tiles, RLD_THREADS, 0, stream>>>(d_counts, d_values, pairs, d_out, out_cap, d_out_size, desc, err);
    return cudaGetLastError();
}

}  // namespace flrl
