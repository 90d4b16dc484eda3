// rl_encode.cu — run-length encoder for sm_100a: one launch, input read once.
//
// The reference documents RL (IMPLEMENTATION-PLAN.md:83-152: mask -> scan -> start indices -> split of
// runs > 255 -> emit) but ships no code.  Structure here:
//
//   * one CTA per 32 KiB tile (4 sub-tiles of 8 KiB, each its own TMA bulk load + mbarrier, so the first
//     sub-tile is processed while the others are still in flight); 4 CTAs resident per SM;
//   * 8 compute warps, thread = one 32-byte segment per sub-tile: run-start bit mask (XOR with the
//     byte-shifted word, SWAR non-zero-byte test, IDP.4A gathers the byte flags into mask bits) -> nearest
//     earlier run start S by ballot (no shuffle scan) -> the one possible 255-split of the segment,
//     (i - S) % 255 == 0 (the plan's ":125" rule, replacing its iterative rescan loop) -> warp sum-scan of
//     pair counts -> compaction of values and LAGGED counts (a pair's count is written when the NEXT
//     boundary is found: no look-ahead halo, no suffix scan) into shared staging -> 16-byte stores aligned
//     in the GLOBAL address space;
//   * a 9th warp runs both decoupled look-backs off the compute warps' critical path: #1 = start of the
//     run that enters the tile (gives the splits that fall before the tile's first run start, and the count
//     of the PREVIOUS tile's last pair, which this tile writes), #2 = the tile's first pair slot;
//   * two CTA barriers per tile; the compute warps never wait for another CTA before the store phase.
//
// A pair belongs to the tile (and the shard) in which it STARTS; its count looks ahead up to 255 bytes, so
// tiles and shards concatenate to the single-stream output byte for byte.
#include <stdlib.h>

#include <algorithm>

#include "flrl_bits.cuh"
#include "lookback.cuh"
#include "rl_kernels.cuh"

namespace flrl {

static constexpr int RL_THREADS = 256;                  // compute threads
static constexpr int RLE_THREADS = RL_THREADS + 32;     // + look-back warp
static constexpr int RL_SEG = 32;                       // input bytes per thread per sub-tile
static constexpr int RL_SUB = RL_THREADS * RL_SEG;      // 8 KiB
static constexpr int RL_NSUB = 4;
static constexpr int RL_TILE = RL_SUB * RL_NSUB;        // 32 KiB
static constexpr int RL_NCHUNK = RL_NSUB * 8;           // warp chunks (1 KiB) per tile = 32
static constexpr int RL_INBUF = 16 + RL_TILE + 16;      // [16 left halo][tile][pad]
static constexpr int RL_CAP = 11264;                    // staged pairs per round
static constexpr int RL_STAGE = RL_CAP + 32;            // + over-read slack of the unaligned 16-byte reads
static constexpr int RLE_CTAS_PER_SM = 4;
static constexpr uint64_t RL_BIAS = 255;                // run starts are stored as S + 255 >= 0
static constexpr uint32_t BAR_ALL = 1;                  // compute warps + look-back warp
static constexpr uint32_t BAR_COMPUTE = 2;              // compute warps only

// ---- run-start flags of 32 consecutive bytes -------------------------------------------------
// bit (4k + j) of the result: byte j of word k differs from the byte before it.
// nz_flags: 0x80 in every non-zero byte.  IDP.4A with weights 1,2,4,8 / 16,32,64,128 turns the eight byte
// flags of a word pair into 128 * (8-bit mask); four pairs are merged with one shift and three IMADs.
__device__ __forceinline__ uint32_t nz_flags(uint32_t x)
{
    return (((x | 0x80808080u) - 0x01010101u) | x) & 0x80808080u;
}
__device__ __forceinline__ uint32_t start_mask32(const uint32_t (&w)[8], uint32_t prev_word)
{
    uint32_t v[4], prev = prev_word;
#pragma unroll
    for (int k = 0; k < 8; k += 2) {
        const uint32_t a = nz_flags(w[k] ^ __byte_perm(prev, w[k], 0x6543));
        const uint32_t b = nz_flags(w[k + 1] ^ __byte_perm(w[k], w[k + 1], 0x6543));
        v[k >> 1] = __dp4a(a, 0x08040201u, __dp4a(b, 0x80402010u, 0u));
        prev = w[k + 1];
    }
    return (v[0] >> 7) + v[1] * 2u + v[2] * 512u + v[3] * 131072u;
}
__device__ __forceinline__ uint32_t segment_mask(const uint8_t* inbuf, int off)
{
    const uint4 a = *reinterpret_cast<const uint4*>(inbuf + 16 + off);
    const uint4 b = *reinterpret_cast<const uint4*>(inbuf + 32 + off);
    const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    return start_mask32(w, *reinterpret_cast<const uint32_t*>(inbuf + 12 + off));
}

// look-back #1: biased start of the last run that begins before tile `tile`.
// d1 states: 0 not yet known, 1 "no run starts in this tile", 2 final value.
__device__ __forceinline__ uint64_t lookback_run_start(const uint64_t* d1, long long tile, int lane)
{
    long long j = tile - 1 - lane;
    uint64_t v = (j >= 0) ? ld_relaxed_u64(d1 + j) : LB_INC;
    while (true) {
        const unsigned fin = __ballot_sync(0xffffffffu, (v >> 62) == 2);
        const unsigned inv = __ballot_sync(0xffffffffu, (v >> 62) == 0);
        const unsigned before = fin ? ((1u << (__ffs(fin) - 1)) - 1u) : 0xffffffffu;  // lanes nearer than the first final one
        if (inv & before) {
            if ((v >> 62) == 0) v = ld_relaxed_u64(d1 + j);  // only the lanes that saw nothing re-poll
            continue;
        }
        if (fin) return __shfl_sync(0xffffffffu, v & LB_MASK, __ffs(fin) - 1);
        j -= 32;
        v = (j >= 0) ? ld_relaxed_u64(d1 + j) : LB_INC;
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

template <typename T>
__device__ __forceinline__ T dmin(T a, T b) { return a < b ? a : b; }
template <typename T>
__device__ __forceinline__ T dmax(T a, T b) { return a > b ? a : b; }

struct RlEncShared {
    int chunk_first[RL_NCHUNK];   // first run start in each 1 KiB warp chunk (tile relative), RL_TILE if none
    int chunk_last[RL_NCHUNK];    // last run start in each warp chunk (tile relative), -1 if none
    uint32_t chunk_cnt[RL_NCHUNK];// boundaries in each warp chunk: low 16 bits all, high 16 bits emitted (position < n)
    // look-back warp -> compute warps
    uint64_t g0;                  // first pair slot of the tile
    uint32_t nl;                  // emitted pairs that start at 255-splits before the tile's first run start
    uint32_t lead_first;          // tile-relative position of the first such split
    uint32_t first_start;         // tile-relative position of the tile's first run start (RL_TILE if none)
    uint32_t nl_all;              // the same splits, emitted or not
};

__global__ void __launch_bounds__(RLE_THREADS, RLE_CTAS_PER_SM)
rl_encode_kernel(const uint8_t* __restrict__ in, size_t n, size_t n_avail, size_t n_scan, uint32_t prev_phase, int has_prev,
                 uint8_t* __restrict__ counts, uint8_t* __restrict__ values, size_t pair_cap, uint64_t* __restrict__ d_pairs,
                 uint64_t* __restrict__ d1, uint64_t* __restrict__ d2, long long tiles)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* const inbuf = smem;                       // [16 left halo][tile][pad]
    uint8_t* const cnt8 = smem + RL_INBUF;             // staged counts
    uint8_t* const val8 = cnt8 + RL_STAGE;             // staged values
    __shared__ __align__(8) uint64_t full[RL_NSUB], base_ready;
    __shared__ RlEncShared sh;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long tile = blockIdx.x;
    const size_t tbase = (size_t)tile * RL_TILE;
    const bool last_tile = tile == tiles - 1;
    // tile-relative limits: positions >= rel_emit are not emitted; position rel_scan is a forced boundary (the end
    // of the data, or far enough into the look-ahead that every emitted pair has its count), nothing after it counts
    const uint32_t rel_emit = n > tbase ? (uint32_t)dmin((size_t)(n - tbase), (size_t)RL_TILE) : 0u;
    const uint32_t rel_scan = (uint32_t)dmin((size_t)(n_scan - tbase), (size_t)RL_TILE);  // == RL_TILE: not in this tile
    const bool ragged = rel_emit < (uint32_t)RL_TILE || rel_scan < (uint32_t)RL_TILE;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < RL_NSUB; ++s) mbar_init(&full[s], 1);
        mbar_init(&base_ready, 1);
        fence_mbar_init();
    }
    __syncthreads();

    if (warp == RL_THREADS / 32) {
        // ============================ loads, then the two look-backs ==========================
        {
            const size_t want = dmin((size_t)tbase + RL_TILE, n_avail);  // bytes of the tile that exist
            const size_t hi16 = want & ~(size_t)15;
            const size_t lo0 = tile > 0 ? tbase - 16 : tbase;
            // the ragged end (< 16 bytes) and the byte "before" the stream: plain loads
            if (lane < 15) {
                const size_t g = dmax(hi16, lo0) + lane;
                if (g < want) inbuf[16 + (g - tbase)] = in[g];
            }
            if (tile == 0 && lane == 15) inbuf[15] = has_prev ? in[0] : (uint8_t)~in[0];
            __syncwarp();
            if (lane < RL_NSUB) {
                const size_t lo = lane == 0 ? lo0 : tbase + (size_t)lane * RL_SUB;
                const size_t hi = dmin((size_t)tbase + (size_t)(lane + 1) * RL_SUB, hi16);
                if (hi > lo) {
                    mbar_expect_tx(&full[lane], (uint32_t)(hi - lo));
                    tma_load_1d(inbuf + 16 + ((long long)lo - (long long)tbase), in + lo, (uint32_t)(hi - lo), &full[lane]);
                } else {
                    mbar_arrive(&full[lane]);
                }
            }
        }
        bar_sync(BAR_ALL, RLE_THREADS);  // #1: chunk_first / chunk_last are posted
        const int cf = sh.chunk_first[lane], cl = sh.chunk_last[lane];
        const int F_t = __reduce_min_sync(0xffffffffu, cf);   // first run start of the tile, RL_TILE if none
        const int L_t = __reduce_max_sync(0xffffffffu, cl);   // last run start, -1 if none
        uint64_t sbB = prev_phase ? RL_BIAS - prev_phase : 0;   // biased start of the run that enters the tile (< RL_BIAS)
        if (tile > 0) {
            if (lane == 0) st_relaxed_u64(d1 + tile, L_t >= 0 ? (LB_INC | (uint64_t)(tbase + L_t + RL_BIAS)) : LB_AGG);
            sbB = lookback_run_start(d1, tile, lane);
            if (L_t < 0 && lane == 0) st_relaxed_u64(d1 + tile, LB_INC | sbB);
        } else if (L_t >= 0 && lane == 0) {
            st_relaxed_u64(d1 + tile, LB_INC | (uint64_t)(tbase + L_t + RL_BIAS));
        } else if (lane == 0) {
            st_relaxed_u64(d1 + tile, LB_INC | sbB);
        }
        // boundaries of the entering run: sb + 255 j.  All quantities biased by RL_BIAS so they stay non-negative.
        const uint64_t tB = tbase + RL_BIAS;                    // biased tile base
        const uint64_t FB = tB + (uint64_t)F_t;                 // biased first run start (or tile end)
        const uint64_t q0 = (tB - 1 - sbB) / 255u;              // sbB < tB always (the entering run starts before the tile)
        const uint64_t BpB = sbB + 255u * q0;                   // last boundary before the tile
        const uint64_t limB = dmin(FB, tB + (uint64_t)rel_emit);  // emitted lead splits lie in [tB, limB)
        const uint32_t nl_all = (uint32_t)((FB - 1 - sbB) / 255u - q0);
        const uint32_t nl = limB > tB ? (uint32_t)((limB - 1 - sbB) / 255u - q0) : 0u;
        const uint32_t lagged = nl_all ? 255u : (uint32_t)(FB - BpB);  // count of the last pair before this tile
        bar_sync(BAR_ALL, RLE_THREADS);  // #2: chunk_cnt is posted
        const uint32_t staged = __reduce_add_sync(0xffffffffu, sh.chunk_cnt[lane] >> 16);
        const uint32_t tot = nl + staged;
        uint64_t g = 0;
        if (tile > 0) {
            if (lane == 0) st_relaxed_u64(d2 + tile, LB_AGG | tot);
            g = lookback_exclusive_sum(d2, tile, lane);
        }
        if (lane == 0) {
            st_relaxed_u64(d2 + tile, LB_INC | (g + tot));
            // the previous tile leaves its last pair's count to us when that pair is the last boundary before this
            // tile (it cannot know where the next boundary is); a boundary at or past n is not a pair
            if (tile > 0 && g >= 1 && g - 1 < pair_cap && BpB < n + RL_BIAS) counts[g - 1] = (uint8_t)lagged;
            sh.g0 = g;
            sh.nl = nl;
            sh.lead_first = (uint32_t)(BpB + 255u - tB);
            sh.first_start = (uint32_t)F_t;
            sh.nl_all = nl_all;
            if (last_tile) *d_pairs = g + tot;
            mbar_arrive(&base_ready);
        }
        return;
    }

    // ================================= compute warps =======================================
    uint32_t bm[RL_NSUB];     // phase 1: run-start mask; phase 2 on: all pair boundaries of the segment
    int sprev[RL_NSUB];       // phase 1: nearest earlier run start inside the warp (-1 none); phase 2 on: last boundary before the segment
    uint32_t slot[RL_NSUB];   // first staged slot of the segment
    const uint32_t lt_mask = (1u << lane) - 1u;

    // ---- phase 1: run-start masks, nearest earlier run start by ballot
#pragma unroll
    for (int s = 0; s < RL_NSUB; ++s) {
        const int off = s * RL_SUB + tid * RL_SEG;
        mbar_wait(&full[s], 0);
        uint32_t m = segment_mask(inbuf, off);
        if (ragged) {  // CTA-uniform: only the last tiles see the end of the data
            const int k = (int)rel_scan - off;
            if (k < 0) m = 0;
            else if (k < 32) m = (m & ((1u << k) - 1u)) | (1u << k);
        }
        const unsigned bal = __ballot_sync(0xffffffffu, m != 0);
        const int my_last = off + 31 - __clz(m);
        const unsigned lower = bal & lt_mask;
        const int S = __shfl_sync(0xffffffffu, my_last, 31 - __clz(lower | 1u));
        sprev[s] = lower ? S : -1;
        bm[s] = m;
        const int wc = s * 8 + warp;
        if (bal == 0) {
            if (lane == 0) {
                sh.chunk_first[wc] = RL_TILE;
                sh.chunk_last[wc] = -1;
            }
        } else {
            if (lane == __ffs(bal) - 1) sh.chunk_first[wc] = off + __ffs(m) - 1;
            if (lane == 31 - __clz(bal)) sh.chunk_last[wc] = my_last;
        }
    }
    bar_sync(BAR_ALL, RLE_THREADS);  // #1

    // ---- phase 2: the one possible 255-split of each segment, pair counts
    {
        const int cl = sh.chunk_last[lane];
        const unsigned b2 = __ballot_sync(0xffffffffu, cl >= 0);
#pragma unroll
        for (int s = 0; s < RL_NSUB; ++s) {
            const int off = s * RL_SUB + tid * RL_SEG;
            const int wc = s * 8 + warp;
            const unsigned lowerb = b2 & ((1u << wc) - 1u);
            const int Sin = __shfl_sync(0xffffffffu, cl, 31 - __clz(lowerb | 1u));
            int S = sprev[s];
            if (S < 0 && lowerb) S = Sin;
            uint32_t b = bm[s];
            if (S >= 0) {
                // boundaries of the entering run lie at S + 255 j; r = distance from the last one to off - 1
                const uint32_t r = (uint32_t)(off - 1 - S) % 255u;
                const uint32_t j = 254u - r;
                const uint32_t lead = b ? (uint32_t)(__ffs(b) - 1) : 32u;  // bytes that continue the entering run
                if (j < lead) b |= 1u << j;
                sprev[s] = off - 1 - (int)r;
            }
            bm[s] = b;
            uint32_t cn = __popc(b);
            if (ragged) {
                const int k = (int)rel_emit - off;
                const uint32_t em = k <= 0 ? 0u : (k >= 32 ? ~0u : ((1u << k) - 1u));
                cn |= __popc(b & em) << 16;
            } else {
                cn |= cn << 16;
            }
            uint32_t inc = cn;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d) inc += o;
            }
            if (lane == 31) sh.chunk_cnt[wc] = inc;
            slot[s] = (inc - cn) & 0xffffu;
        }
    }
    bar_sync(BAR_ALL, RLE_THREADS);  // #2

    // ---- phase 3: slots, compaction (values + lagged counts), stores
    uint32_t total_all, total_emit;
    {
        const uint32_t cc = sh.chunk_cnt[lane];
        uint32_t inc = cc;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += o;
        }
        const uint32_t tot = __shfl_sync(0xffffffffu, inc, 31);
        total_all = tot & 0xffffu;
        total_emit = tot >> 16;
        const uint32_t excl = inc - cc;
#pragma unroll
        for (int s = 0; s < RL_NSUB; ++s) slot[s] += __shfl_sync(0xffffffffu, excl, s * 8 + warp) & 0xffffu;
    }
    const bool simple = total_all == total_emit && total_all <= (uint32_t)RL_CAP;
    bool have_base = false;
    uint64_t g0 = 0;
    uint32_t nl = 0;
    for (uint32_t w0 = 0; w0 == 0 || w0 < total_emit; w0 += RL_CAP) {
        if (simple) {
#pragma unroll
            for (int s = 0; s < RL_NSUB; ++s) {
                const int off = s * RL_SUB + tid * RL_SEG;
                uint32_t bits = bm[s], kk = slot[s];
                int prevpos = sprev[s];
                while (bits) {
                    const int pos = off + __ffs(bits) - 1;
                    bits &= bits - 1;
                    val8[kk] = inbuf[16 + pos];
                    if (kk) cnt8[kk - 1] = (uint8_t)(pos - prevpos);
                    prevpos = pos;
                    ++kk;
                }
            }
        } else {
            if (w0) bar_sync(BAR_COMPUTE, RL_THREADS);  // the previous round's stores have read the staging buffers
#pragma unroll
            for (int s = 0; s < RL_NSUB; ++s) {
                const int off = s * RL_SUB + tid * RL_SEG;
                uint32_t bits = bm[s], kk = slot[s];
                int prevpos = sprev[s];
                while (bits) {
                    const int pos = off + __ffs(bits) - 1;
                    bits &= bits - 1;
                    if (kk >= w0 && kk < w0 + RL_CAP && kk < total_emit) val8[kk - w0] = inbuf[16 + pos];
                    if (kk > w0 && kk <= w0 + RL_CAP && kk <= total_emit) cnt8[kk - 1 - w0] = (uint8_t)(pos - prevpos);
                    prevpos = pos;
                    ++kk;
                }
            }
        }
        bar_sync(BAR_COMPUTE, RL_THREADS);
        if (!have_base) {
            mbar_wait(&base_ready, 0);
            g0 = sh.g0;
            nl = sh.nl;
            have_base = true;
            // pairs that start at 255-splits of the entering run, before the tile's first run start
            const uint32_t first_start = sh.first_start, lead_first = sh.lead_first;
            const bool lead_lag = sh.nl_all == nl && total_all == 0;  // the last of them is the tile's last boundary
            for (uint32_t j = tid; j < nl; j += RL_THREADS) {
                const uint32_t p = lead_first + 255u * j;
                const uint64_t k = g0 + j;
                if (k < pair_cap) {
                    values[k] = inbuf[16];
                    // a pair that starts at the tile's last boundary gets its count from the next tile
                    if (!(j == nl - 1 && lead_lag)) counts[k] = (uint8_t)min(255u, first_start - p);
                }
            }
        }
        if (total_emit > w0) {
            const uint32_t cw = min((uint32_t)RL_CAP, total_emit - w0);
            const bool tail = w0 + cw == total_emit && total_all == total_emit;  // that count is written by the next tile
            store_compacted(values, g0 + nl + w0, cw, pair_cap, val8, tid);
            store_compacted(counts, g0 + nl + w0, cw - (tail ? 1u : 0u), pair_cap, cnt8, tid);
        }
    }
}

// =============================================================================================
//                                        launcher
// =============================================================================================
// positions 0 .. n_scan are scanned, n_scan = min(n_avail, n + 255) being a forced boundary
size_t rl_encode_tiles(size_t n) { return (n + 255) / RL_TILE + 1; }
size_t rl_encode_workspace_bytes(size_t n) { return (2 * rl_encode_tiles(n) + 2) * sizeof(uint64_t); }

cudaError_t rl_encode_launch(const uint8_t* d_in, size_t n, size_t n_avail, uint64_t prev_run, uint8_t* d_counts,
                             uint8_t* d_values, size_t pair_cap, uint64_t* d_pairs, void* d_ws, cudaStream_t stream)
{
    if (n == 0) return cudaMemsetAsync(d_pairs, 0, sizeof(uint64_t), stream);
    const size_t n_scan = std::min(n_avail, n + 255);
    const size_t tiles = n_scan / RL_TILE + 1;
    uint64_t* d1 = reinterpret_cast<uint64_t*>(d_ws);
    uint64_t* d2 = d1 + rl_encode_tiles(n);
    cudaError_t e = cudaMemsetAsync(d1, 0, 2 * rl_encode_tiles(n) * sizeof(uint64_t), stream);
    if (e != cudaSuccess) return e;
    constexpr int smem = RL_INBUF + 2 * RL_STAGE;
    static bool configured[64];
    int dev = 0;
    e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev >= 64 || !configured[dev]) {
        e = cudaFuncSetAttribute((const void*)rl_encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        if (dev < 64) configured[dev] = true;
    }
    // tiles are taken in blockIdx order, so every tile a CTA waits for has already started
    rl_encode_kernel<<<(unsigned)tiles, RLE_THREADS, smem, stream>>>(d_in, n, n_avail, n_scan, (uint32_t)(prev_run % 255u), prev_run != 0,
                                                                      d_counts, d_values, pair_cap, d_pairs, d1, d2, (long long)tiles);
    return cudaGetLastError();
}

}  // namespace flrl
