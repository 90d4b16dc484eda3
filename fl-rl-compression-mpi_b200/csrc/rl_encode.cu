// rl_encode.cu — run-length encoder for sm_100a: one launch, input read once.
//
// The reference documents RL (IMPLEMENTATION-PLAN.md:83-152: mask -> scan -> start indices -> split of
// runs > 255 -> emit) but ships no code.  Structure here:
//
//   * small persistent CTAs (6 per SM) that take 16 KiB tiles from an atomic counter, so tile order is time
//     order: everything a tile depends on was started earlier, by a CTA that never waits for a later tile;
//   * 4 compute warps, thread = one 32-byte segment per 4 KiB sub-tile (four TMA bulk loads per tile, each with
//     its own mbarrier): run-start bit mask (XOR with the byte-shifted word, SWAR non-zero-byte test, IDP.4A
//     gathers the byte flags into mask bits) -> nearest earlier run start S by ballot (no shuffle scan) -> the
//     one possible 255-split of the segment, (i - S) % 255 == 0 (the plan's ":125" rule, replacing its iterative
//     rescan loop; skipped when the entering run is too short to reach 255) -> warp sum-scans of the pair
//     counts -> compaction of values and LAGGED counts (a pair's count is written when the NEXT boundary is
//     found: no look-ahead halo, no suffix scan) into a shared-memory RING of staged pairs;
//   * the compute warps never wait for another CTA: a tile's pairs stay in the ring until its first slot is
//     known and are written then, several tiles later (measured: the first slot arrives ~7 us after the
//     tile's pair count is posted, because it needs the slowest of some hundred earlier tiles); 16-byte
//     stores aligned in the GLOBAL address space (unaligned shared reads fixed up with PRMT);
//   * a service warp per CTA: tile ids, loads, look-back #1 (start of the run that enters the tile: gives the
//     splits that fall before the tile's first run start and the count of the PREVIOUS tile's last pair,
//     which this tile writes), posts the tile's pair count;
//   * one scanner warp for the whole grid turns the posted pair counts into first slots, in tile order (poll
//     traffic O(tiles), where per-tile decoupled look-backs read O(tiles x window)); a base warp per CTA polls
//     the slots of the CTA's waiting tiles, one lane per tile.
//
// A pair belongs to the tile (and the shard) in which it STARTS; its count looks ahead up to 255 bytes, so
// tiles and shards concatenate to the single-stream output byte for byte.
#include <stdlib.h>

#include <algorithm>

#include "flrl_bits.cuh"
#include "lookback.cuh"
#include "rl_kernels.cuh"
#include "rl_scan.cuh"

namespace flrl {

static constexpr int RL_WARPS = 4;                      // compute warps
static constexpr int RL_THREADS = RL_WARPS * 32;        // compute threads
static constexpr int RLE_THREADS = RL_THREADS + 64;     // + service warp + base warp
static constexpr int RL_SEG = 32;                       // input bytes per thread per sub-tile
static constexpr int RL_SUB = RL_THREADS * RL_SEG;      // 4 KiB
#ifndef RL_NSUB_N
#define RL_NSUB_N 4
#endif
static constexpr int RL_NSUB = RL_NSUB_N;               // 4 KiB sub-tiles per tile
static constexpr int RL_TILE = RL_SUB * RL_NSUB;        // 16 KiB
static constexpr int RL_NCHUNK = RL_NSUB * RL_WARPS;    // warp chunks (1 KiB) per tile
static constexpr int RL_INBUF = 16 + RL_TILE + 16;      // [16 left halo][tile][pad]
#ifndef RL_RING_PAIRS
#define RL_RING_PAIRS 8192
#endif
// pause between two looks of a polling loop (ns): the base warp with nothing to hand out / with answers pending
#ifndef RLE_SLEEP_A
#define RLE_SLEEP_A 100
#endif
#ifndef RLE_SLEEP_B
#define RLE_SLEEP_B 200
#endif
#ifndef RLE_PF_LEAD
#define RLE_PF_LEAD 1024  // tiles between a load and the L2 prefetch issued with it (0: none)
#endif
#ifndef RLE_CTAS
#define RLE_CTAS 6
#endif
static constexpr int RL_RING = RL_RING_PAIRS;           // staged pairs (power of two)
static constexpr int RL_RINGB = RL_RING + 32;           // + over-read slack of the unaligned 16-byte reads
static constexpr int RLE_Q = 32;                        // tiles of one CTA that may wait for their first slot
static constexpr int RLE_CTAS_PER_SM = RLE_CTAS;
static constexpr uint64_t RL_BIAS = 255;                // run starts are stored as S + 255 >= 0
static constexpr uint32_t BAR_COMPUTE = 1;              // named barrier of the compute warps
static_assert(RL_NCHUNK <= 32 && RL_NSUB % 2 == 0, "chunk arrays are handled by one warp; sub-tiles are scanned in pairs");
static_assert((RL_RING & (RL_RING - 1)) == 0 && RL_RING <= RL_TILE, "ring indices are masked");

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

// cnt staged bytes stage[soff .. soff + cnt) -> dst[g0 .. g0 + cnt): coalesced 16-byte stores on 16-byte aligned
// global addresses, the (at most 15 + 15) bytes of the two ragged ends as one byte store per thread; nothing at
// or past index `cap` is written.
__device__ __forceinline__ void store_compacted(uint8_t* __restrict__ dst, uint64_t g0, uint32_t cnt, uint64_t cap,
                                                const uint8_t* stage, uint32_t soff, int tid)
{
    if (cap <= g0) return;
    if (cap - g0 < cnt) cnt = (uint32_t)(cap - g0);
    const uint32_t a = (uint32_t)((reinterpret_cast<uintptr_t>(dst) + g0) & 15u);
    const uint32_t head = min(cnt, (16u - a) & 15u);        // bytes before the first aligned chunk
    const uint32_t nfull = (cnt - head) >> 4;               // aligned 16-byte chunks
    const uint32_t tail0 = head + 16u * nfull;              // staging index of the first byte after them
    for (uint32_t q = tid; q < nfull; q += RL_THREADS)
        *reinterpret_cast<uint4*>(dst + g0 + head + 16u * q) = lds_unaligned16(stage, soff + head + 16u * q);
    if ((uint32_t)tid < head) dst[g0 + tid] = stage[soff + tid];
    const uint32_t t2 = (uint32_t)tid - 16u;                // threads 16..31 take the tail
    if (t2 < cnt - tail0) dst[g0 + tail0 + t2] = stage[soff + tail0 + t2];
}
// the same from a ring of RL_RING bytes: at most two linear pieces
__device__ __forceinline__ void store_from_ring(uint8_t* __restrict__ dst, uint64_t g0, uint32_t cnt, uint64_t cap,
                                                const uint8_t* ring, uint32_t head, int tid)
{
    const uint32_t h = head & (RL_RING - 1);
    const uint32_t first = min(cnt, (uint32_t)RL_RING - h);
    store_compacted(dst, g0, first, cap, ring, h, tid);
    if (cnt > first) store_compacted(dst, g0 + first, cnt - first, cap, ring, 0, tid);
}

template <typename T>
__device__ __forceinline__ T dmin(T a, T b) { return a < b ? a : b; }
template <typename T>
__device__ __forceinline__ T dmax(T a, T b) { return a > b ? a : b; }

#ifdef FLRL_TRACE
// development aid (variant builds only): per-tile globaltimer stamps
__device__ uint64_t* g_rle_trace = nullptr;
__device__ __forceinline__ uint64_t gtime()
{
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define TRACE(slot)                                                                      \
    do {                                                                                 \
        if (g_rle_trace && lane == 0) g_rle_trace[(size_t)tile * 16 + (slot)] = gtime(); \
    } while (0)
#else
#define TRACE(slot) do { } while (0)
#endif

struct RlePending {               // a tile of this CTA, from "taken" to "stored"
    // written by the compute warps
    uint32_t head;                 // ring position of the tile's first staged pair
    uint32_t total;                // staged pairs that wait in the ring
    uint32_t total_all;            // pair boundaries of the tile after its first run start (emitted or not)
    uint32_t first_byte;           // in[tile base]: the value of the run that enters the tile
    // written by the service warp
    long long tile;
    uint32_t tot;                  // pairs of the tile (staged + nl)
    uint32_t nl;                   // emitted pairs that start at 255-splits before the tile's first run start
    uint32_t nl_all;               // the same splits, emitted or not
    uint32_t lead_first;           // tile-relative position of the first such split
    uint32_t first_start;          // tile-relative position of the tile's first run start (RL_TILE if none)
    uint32_t lagged;               // count of the last pair before the tile; 0 = not ours to write
    // written by the base warp
    uint64_t g0;                   // first pair slot of the tile
};

struct RleShared {
    int chunk_first[RL_NCHUNK];    // first run start in each 1 KiB warp chunk (tile relative), RL_TILE if none
    int chunk_last[RL_NCHUNK];     // last run start in each warp chunk (tile relative), -1 if none
    uint32_t chunk_cnt[RL_NCHUNK]; // pair boundaries in each warp chunk
    uint32_t emit_cnt[RL_WARPS];   // ragged tiles only: boundaries at positions < n, per warp
    long long cur_tile;            // service warp -> compute warps (-1: no more tiles)
    uint32_t n_meta;               // service warp -> base warp: tiles whose pair count is posted
    uint32_t n_ready;              // base warp -> compute warps: tiles whose first slot is known
    uint32_t store_upto;           // compute warps' common snapshot of n_ready
    uint32_t done;                 // service warp -> base warp: no more tiles
    RlePending q[RLE_Q];
};

__device__ __forceinline__ uint32_t ld_volatile_shared(const uint32_t* p) { return *reinterpret_cast<const volatile uint32_t*>(p); }
__device__ __forceinline__ void st_volatile_shared(uint32_t* p, uint32_t v) { *reinterpret_cast<volatile uint32_t*>(p) = v; }

__global__ void __launch_bounds__(RLE_THREADS, RLE_CTAS_PER_SM)
rl_encode_kernel(const uint8_t* __restrict__ in, size_t n, size_t n_avail, size_t n_scan, uint32_t prev_phase, int has_prev,
                 uint8_t* __restrict__ counts, uint8_t* __restrict__ values, size_t pair_cap, uint64_t* __restrict__ d_pairs,
                 uint64_t* __restrict__ d1, uint64_t* __restrict__ inc, uint32_t* __restrict__ agg, uint32_t* __restrict__ next_tile,
                 long long tiles)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* const inbuf = smem;                       // [16 left halo][tile][pad]
    uint8_t* const cnt8 = smem + RL_INBUF;             // ring of staged counts
    uint8_t* const val8 = cnt8 + RL_RINGB;             // ring of staged values
    __shared__ __align__(8) uint64_t full[RL_NSUB], freeb, p1_ready, p2_ready;
    __shared__ RleShared sh;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (blockIdx.x == 0) {  // the scanner CTA
        if (warp == 0) scan_pair_counts(agg, inc, tiles, lane);
        return;
    }
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < RL_NSUB; ++s) mbar_init(&full[s], 1);
        mbar_init(&freeb, RL_WARPS);
        mbar_init(&p1_ready, RL_WARPS);
        mbar_init(&p2_ready, RL_WARPS);
        fence_mbar_init();
        sh.n_meta = 0;
        sh.n_ready = 0;
        sh.done = 0;
    }
    __syncthreads();

    if (warp == RL_WARPS) {
        // ================================ service warp ===========================================
        long long next = 0;
        if (lane == 0) next = atomicAdd(next_tile, 1u);
        next = __shfl_sync(0xffffffffu, next, 0);
        for (uint32_t k = 0;; ++k) {
            const long long tile = next;
            const uint32_t par = k & 1u;
            if (k) mbar_wait(&freeb, (k - 1) & 1u);  // the compute warps are done with the input buffer
            if (tile >= tiles) {
                if (lane == 0) sh.cur_tile = -1;
                __syncwarp();
                if (lane < RL_NSUB) mbar_arrive(&full[lane]);
                break;
            }
            const size_t tbase = (size_t)tile * RL_TILE;
            if (lane == 0) sh.cur_tile = tile;
            {
                const size_t want = dmin(tbase + RL_TILE, n_avail);  // end of the bytes of the tile that exist
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
                    const size_t hi = dmin(tbase + (size_t)(lane + 1) * RL_SUB, hi16);
                    if (hi > lo) {
                        mbar_expect_tx(&full[lane], (uint32_t)(hi - lo));
                        tma_load_1d(inbuf + 16 + ((long long)lo - (long long)tbase), in + lo, (uint32_t)(hi - lo), &full[lane]);
                    } else {
                        mbar_arrive(&full[lane]);
                    }
#if RLE_PF_LEAD
                    // A CTA has one input buffer, so a tile pays its load latency in full: make that an L2 hit.  Whoever
                    // loads tile t asks L2 for tile t + RLE_PF_LEAD (somebody else's, about one grid of tiles later).
                    {
                        const size_t pb = tbase + (size_t)RLE_PF_LEAD * RL_TILE + (size_t)lane * RL_SUB;
                        const size_t pe = dmin(pb + RL_SUB, n_avail & ~(size_t)15);
                        if (pe > pb) tma_prefetch_l2(in + pb, (uint32_t)(pe - pb));
                    }
#endif
                }
            }
            TRACE(0);
            const uint32_t rel_emit = n > tbase ? (uint32_t)dmin((size_t)(n - tbase), (size_t)RL_TILE) : 0u;
            const uint32_t rel_scan = (uint32_t)dmin((size_t)(n_scan - tbase), (size_t)RL_TILE);
            const bool ragged = rel_emit < (uint32_t)RL_TILE || rel_scan < (uint32_t)RL_TILE;
            // look-back #1 starts at once: it does not depend on this tile (the compute warps publish d1[tile] themselves)
            uint64_t sbB = prev_phase ? RL_BIAS - prev_phase : 0;   // biased start of the run that enters the tile
            if (tile > 0) sbB = lookback_run_start(d1, tile, lane);
            TRACE(1);
            mbar_wait(&p1_ready, par);  // chunk_first / chunk_last are posted
            const int cf = lane < RL_NCHUNK ? sh.chunk_first[lane] : RL_TILE;
            const int cl = lane < RL_NCHUNK ? sh.chunk_last[lane] : -1;
            const int F_t = __reduce_min_sync(0xffffffffu, cf);   // first run start of the tile, RL_TILE if none
            const int L_t = __reduce_max_sync(0xffffffffu, cl);   // last run start, -1 if none
            if (L_t < 0 && lane == 0) st_relaxed_u64(d1 + tile, LB_INC | sbB);  // no run starts here: pass the entering one on
            // boundaries of the entering run: sb + 255 j.  All positions biased by RL_BIAS so they stay non-negative.
            const uint64_t tB = tbase + RL_BIAS;                    // biased tile base; sbB < tB always
            const uint64_t FB = tB + (uint64_t)F_t;                 // biased first run start (or tile end)
            const uint64_t q0 = (tB - 1 - sbB) / 255u;
            const uint64_t BpB = sbB + 255u * q0;                   // last boundary before the tile
            const uint64_t limB = dmin(FB, tB + (uint64_t)rel_emit);  // emitted lead splits lie in [tB, limB)
            const uint32_t nl_all = (uint32_t)((FB - 1 - sbB) / 255u - q0);
            const uint32_t nl = limB > tB ? (uint32_t)((limB - 1 - sbB) / 255u - q0) : 0u;
            const uint32_t lagged = nl_all ? 255u : (uint32_t)(FB - BpB);  // count of the last pair before this tile
            mbar_wait(&p2_ready, par);  // chunk_cnt (and emit_cnt) are posted
            TRACE(3);
            const uint32_t mine = ragged ? (lane < RL_WARPS ? sh.emit_cnt[lane] : 0u) : (lane < RL_NCHUNK ? sh.chunk_cnt[lane] : 0u);
            const uint32_t tot = nl + __reduce_add_sync(0xffffffffu, mine);
            if (lane == 0) {
                st_relaxed_u32(agg + tile, 0x80000000u | tot);  // to the scanner
                RlePending& e = sh.q[k % RLE_Q];
                e.tile = tile;
                e.tot = tot;
                e.nl = nl;
                e.nl_all = nl_all;
                e.lead_first = (uint32_t)(BpB + 255u - tB);
                e.first_start = (uint32_t)F_t;
                // the previous tile leaves its last pair's count to us when that pair is the last boundary before this
                // tile (it cannot know where the next boundary is); a boundary at or past n is not a pair
                e.lagged = (tile > 0 && BpB < n + RL_BIAS) ? lagged : 0u;
                __threadfence_block();
                st_volatile_shared(&sh.n_meta, k + 1);
                // the next tile id: taken late, so that tiles start (and publish) nearly in id order, yet it is here
                // before the compute warps release the input buffer
                next = atomicAdd(next_tile, 1u);
            }
            next = __shfl_sync(0xffffffffu, next, 0);
        }
        if (lane == 0) {
            __threadfence_block();
            st_volatile_shared(&sh.done, 1u);
        }
        return;
    }

    if (warp == RL_WARPS + 1) {
        // ================================== base warp ============================================
        // one lane per waiting tile polls the scanner's answer; first slots are handed over in tile order
        uint32_t n_ready = 0;
        while (true) {
            const uint32_t done = ld_volatile_shared(&sh.done);
            const uint32_t n_meta = ld_volatile_shared(&sh.n_meta);
            if (n_ready == n_meta) {
                if (done && n_ready == ld_volatile_shared(&sh.n_meta)) break;
                __nanosleep(RLE_SLEEP_A);
                continue;
            }
            __threadfence_block();
            const uint32_t idx = n_ready + lane;
            const bool have = idx < n_meta;
            RlePending& e = sh.q[idx % RLE_Q];
            uint64_t g = 0;
            if (have) g = e.tile > 0 ? ld_relaxed_u64(inc + e.tile) : (1ull << 63);
            const unsigned ok = __ballot_sync(0xffffffffu, have && (g >> 63));
            const uint32_t r = (uint32_t)__ffs(~ok) - 1u;  // leading lanes whose slot is known (ffs(0) = 0 -> all 32)
            const uint32_t rr = ok == 0xffffffffu ? 32u : r;
            if (lane < rr) {
                g &= ~(1ull << 63);
                e.g0 = g;
#ifdef FLRL_TRACE
                if (g_rle_trace) g_rle_trace[(size_t)e.tile * 16 + 5] = gtime();
#endif
                if (e.lagged && g >= 1 && g - 1 < pair_cap) counts[g - 1] = (uint8_t)e.lagged;
                if (e.tile == tiles - 1) *d_pairs = g + e.tot;
            }
            __syncwarp();
            if (rr) {
                n_ready += rr;
                if (lane == 0) {
                    __threadfence_block();
                    st_volatile_shared(&sh.n_ready, n_ready);
                }
            } else {
                __nanosleep(RLE_SLEEP_B);
            }
        }
        return;
    }

    // ================================= compute warps =======================================
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint32_t n_stored = 0;   // tiles of this CTA already written to global memory
    uint32_t ring_head = 0;  // ring position after the last staged pair (not masked)

    // writes a tile whose first slot is known: the pairs at 255-splits before its first run start (long runs only),
    // then its staged pairs.  The tile's last count is left to the next tile (it knows where the next boundary is).
    auto store_tile = [&](const RlePending& e) {
        const uint32_t nl = e.nl;
        const bool lead_lag = e.nl_all == nl && e.total_all == 0;  // the last split is the tile's last boundary
        for (uint32_t j = tid; j < nl; j += RL_THREADS) {
            const uint32_t p = e.lead_first + 255u * j;
            const uint64_t kk = e.g0 + j;
            if (kk < pair_cap) {
                values[kk] = (uint8_t)e.first_byte;
                if (!(j == nl - 1 && lead_lag)) counts[kk] = (uint8_t)min(255u, e.first_start - p);
            }
        }
        if (e.total) {
            store_from_ring(values, e.g0 + nl, e.total, pair_cap, val8, e.head, tid);
            store_from_ring(counts, e.g0 + nl, e.total - 1u, pair_cap, cnt8, e.head, tid);
        }
    };
    // blocks until the first slots of the CTA's first `need` tiles are known; returns how many are (>= need)
    auto wait_ready = [&](uint32_t need) -> uint32_t {
        bar_sync(BAR_COMPUTE, RL_THREADS);  // everybody has read the previous snapshot
        if (tid == 0) {
            uint32_t r;
            while ((r = ld_volatile_shared(&sh.n_ready)) < need) __nanosleep(RLE_SLEEP_A);
            __threadfence_block();
            sh.store_upto = r;
        }
        bar_sync(BAR_COMPUTE, RL_THREADS);
        return sh.store_upto;
    };

    for (uint32_t k = 0;; ++k) {
        const uint32_t par = k & 1u;
        if (k - n_stored >= (uint32_t)RLE_Q - 1u) {  // the queue of waiting tiles is full (rare)
            const uint32_t upto = wait_ready(n_stored + 1u);
            for (; n_stored < upto; ++n_stored) store_tile(sh.q[n_stored % RLE_Q]);
        }
#pragma unroll
        for (int s = 0; s < RL_NSUB; ++s) mbar_wait(&full[s], par);  // the whole tile: the four mask computations below interleave
        const long long tile = *reinterpret_cast<const volatile long long*>(&sh.cur_tile);
        if (tile < 0) break;
        const size_t tbase = (size_t)tile * RL_TILE;
        // tile-relative limits: positions >= rel_emit are not emitted; position rel_scan is a forced boundary (the end
        // of the data, or far enough into the look-ahead that every emitted pair has its count), nothing after it counts
        const uint32_t rel_emit = n > tbase ? (uint32_t)dmin((size_t)(n - tbase), (size_t)RL_TILE) : 0u;
        const uint32_t rel_scan = (uint32_t)dmin((size_t)(n_scan - tbase), (size_t)RL_TILE);  // == RL_TILE: not in this tile
        const bool ragged = rel_emit < (uint32_t)RL_TILE || rel_scan < (uint32_t)RL_TILE;

        uint32_t bm[RL_NSUB];     // phase 1: run-start mask; phase 2 on: all pair boundaries of the segment
        int sprev[RL_NSUB];       // phase 1: nearest earlier run start inside the warp (-1 none); phase 2 on: last boundary before the segment
        uint32_t slot[RL_NSUB];   // first staged slot of the segment (tile relative)

        // ---- phase 1: run-start masks, nearest earlier run start by ballot.  The LAST sub-tile goes first: the next
        // tile's look-back #1 only needs this tile's last run start, which is published the moment it is known
        if (warp == 0) TRACE(8);
#pragma unroll
        for (int i = 0; i < RL_NSUB; ++i) {
            const int s = (i + RL_NSUB - 1) % RL_NSUB;
            const int off = s * RL_SUB + tid * RL_SEG;
            uint32_t m = segment_mask(inbuf, off);
            if (ragged) {  // CTA-uniform: only the last tiles see the end of the data
                const int kk = (int)rel_scan - off;
                if (kk < 0) m = 0;
                else if (kk < 32) m = (m & ((1u << kk) - 1u)) | (1u << kk);
            }
            const unsigned bal = __ballot_sync(0xffffffffu, m != 0);
            const int my_last = off + 31 - __clz(m);
            const unsigned lower = bal & lt_mask;
            const int S = __shfl_sync(0xffffffffu, my_last, 31 - __clz(lower | 1u));
            sprev[s] = lower ? S : -1;
            bm[s] = m;
            const int wc = s * RL_WARPS + warp;
            if (m != 0 && lower == 0) sh.chunk_first[wc] = off + __ffs(m) - 1;
            if (m != 0 && (bal >> lane) == 1u) {
                sh.chunk_last[wc] = my_last;
                if (wc == RL_NCHUNK - 1) st_relaxed_u64(d1 + tile, LB_INC | (uint64_t)(tbase + my_last + RL_BIAS));
            }
            if (bal == 0 && lane == 0) {
                sh.chunk_first[wc] = RL_TILE;
                sh.chunk_last[wc] = -1;
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&p1_ready);
        if (warp == 0) TRACE(9);
        bar_sync(BAR_COMPUTE, RL_THREADS);  // #1

        // ---- phase 2: the one possible 255-split of each segment, pair counts
        {
            const int cl = lane < RL_NCHUNK ? sh.chunk_last[lane] : -1;
            const unsigned b2 = __ballot_sync(0xffffffffu, cl >= 0);
            if (warp == 0 && !(b2 >> (RL_NCHUNK - 1))) {
                // the tile's last 1 KiB held no run start, so d1[tile] is still empty: last run start of the tile, or "none"
                // (the service warp replaces "none" by the entering run's start once it knows it)
                const int L_t = __reduce_max_sync(0xffffffffu, cl);
                if (lane == 0) {
                    if (L_t >= 0) st_relaxed_u64(d1 + tile, LB_INC | (uint64_t)(tbase + L_t + RL_BIAS));
                    else if (tile > 0) atomicMax(reinterpret_cast<unsigned long long*>(d1 + tile), (unsigned long long)LB_AGG);  // never undoes the final value
                }
            }
            uint32_t cn[RL_NSUB / 2], cne = 0;
#pragma unroll
            for (int s = 0; s < RL_NSUB; ++s) {
                const int off = s * RL_SUB + tid * RL_SEG;
                const int wc = s * RL_WARPS + warp;
                const unsigned lowerb = b2 & ((1u << wc) - 1u);
                const int Sin = __shfl_sync(0xffffffffu, cl, 31 - __clz(lowerb | 1u));
                int S = sprev[s];
                if (S < 0 && lowerb) S = Sin;
                uint32_t bb = bm[s];
                if (S >= 0) {
                    const uint32_t lead = bb ? (uint32_t)(__ffs(bb) - 1) : 32u;  // bytes that continue the entering run
                    if ((uint32_t)(off - S) + lead > 255u) {
                        // boundaries of the entering run lie at S + 255 j; r = distance from the last one to off - 1
                        const uint32_t r = (uint32_t)(off - 1 - S) % 255u;
                        const uint32_t j = 254u - r;
                        if (j < lead) bb |= 1u << j;
                        S = off - 1 - (int)r;
                    }
                }
                sprev[s] = S;  // last boundary before the segment (or -1: none in this tile)
                bm[s] = bb;
                if (s & 1) cn[s >> 1] |= __popc(bb) << 16;
                else cn[s >> 1] = __popc(bb);
                if (ragged) {
                    const int kk = (int)rel_emit - off;
                    const uint32_t em = kk <= 0 ? 0u : (kk >= 32 ? ~0u : ((1u << kk) - 1u));
                    cne += __popc(bb & em);
                }
            }
#pragma unroll
            for (int h = 0; h < RL_NSUB / 2; ++h) {
                uint32_t inc2 = cn[h];
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const uint32_t o = __shfl_up_sync(0xffffffffu, inc2, d);
                    if (lane >= d) inc2 += o;
                }
                if (lane == 31) {
                    sh.chunk_cnt[(2 * h) * RL_WARPS + warp] = inc2 & 0xffffu;
                    sh.chunk_cnt[(2 * h + 1) * RL_WARPS + warp] = inc2 >> 16;
                }
                const uint32_t ex = inc2 - cn[h];
                slot[2 * h] = ex & 0xffffu;
                slot[2 * h + 1] = ex >> 16;
            }
            if (ragged) {
                cne = __reduce_add_sync(0xffffffffu, cne);
                if (lane == 0) sh.emit_cnt[warp] = cne;
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&p2_ready);
        if (tid == 0) {  // which of the waiting tiles can be written after this tile's compaction
            const uint32_t r = ld_volatile_shared(&sh.n_ready);
            __threadfence_block();
            sh.store_upto = r;
        }
        if (warp == 0) TRACE(10);
        bar_sync(BAR_COMPUTE, RL_THREADS);  // #2
        uint32_t upto = sh.store_upto;

        // ---- phase 3: slots, compaction (values + lagged counts) into the ring
        uint32_t total_all, total_emit;
        {
            const uint32_t cc = lane < RL_NCHUNK ? sh.chunk_cnt[lane] : 0u;
            uint32_t csum = cc;
#pragma unroll
            for (int d = 1; d < RL_NCHUNK; d <<= 1) {
                const uint32_t o = __shfl_up_sync(0xffffffffu, csum, d);
                if (lane >= d) csum += o;
            }
            total_all = __shfl_sync(0xffffffffu, csum, RL_NCHUNK - 1);
            total_emit = total_all;
            if (ragged) {
                const uint32_t e = lane < RL_WARPS ? sh.emit_cnt[lane] : 0u;
                total_emit = __reduce_add_sync(0xffffffffu, e);
            }
            const uint32_t excl = csum - cc;
#pragma unroll
            for (int s = 0; s < RL_NSUB; ++s) slot[s] += __shfl_sync(0xffffffffu, excl, s * RL_WARPS + warp);
        }
        RlePending& me = sh.q[k % RLE_Q];
        const bool deferred = total_all == total_emit && total_all <= (uint32_t)RL_RING;
        if (deferred) {
            // room in the ring?  (the oldest waiting tile's pairs start at its `head`)
            if (n_stored < k && ring_head + total_all - sh.q[n_stored % RLE_Q].head > (uint32_t)RL_RING) {
                do {
                    upto = wait_ready(n_stored + 1u);
                    for (; n_stored < upto && n_stored < k; ++n_stored) store_tile(sh.q[n_stored % RLE_Q]);
                } while (n_stored < k && ring_head + total_all - sh.q[n_stored % RLE_Q].head > (uint32_t)RL_RING);
                bar_sync(BAR_COMPUTE, RL_THREADS);  // every warp has read the ring space that is written next
            }
#pragma unroll
            for (int s = 0; s < RL_NSUB; ++s) {
                const int off = s * RL_SUB + tid * RL_SEG;
                const uint8_t* src = inbuf + 16 + off;
                uint32_t bits = bm[s], kk = slot[s];
                int prevpos = sprev[s];
                while (bits) {
                    const int bi = __ffs(bits) - 1;
                    bits &= bits - 1;
                    val8[(ring_head + kk) & (RL_RING - 1)] = src[bi];
                    if (kk) cnt8[(ring_head + kk - 1u) & (RL_RING - 1)] = (uint8_t)(off + bi - prevpos);
                    prevpos = off + bi;
                    ++kk;
                }
            }
            if (tid == 0) {
                me.head = ring_head;
                me.total = total_all;
                me.total_all = total_all;
                me.first_byte = inbuf[16];
            }
            ring_head += total_all;
            __syncwarp();
            if (lane == 0) mbar_arrive(&freeb);  // the input buffer may be refilled
            if (warp == 0) TRACE(12);
            for (; n_stored < upto && n_stored < k; ++n_stored) store_tile(sh.q[n_stored % RLE_Q]);
        } else {
            // rare: a ragged last tile, or more pairs than the ring holds -> written at once, in rounds, after
            // everything that waits
            if (tid == 0) {
                me.head = ring_head;
                me.total = 0;
                me.total_all = total_all;
                me.first_byte = inbuf[16];
            }
            upto = wait_ready(k + 1u);
            for (; n_stored < k; ++n_stored) store_tile(sh.q[n_stored % RLE_Q]);
            store_tile(me);  // its pairs at 255-splits before the first run start
            const uint64_t g0 = me.g0 + me.nl;
            for (uint32_t w0 = 0; w0 < total_emit; w0 += RL_RING) {
                bar_sync(BAR_COMPUTE, RL_THREADS);  // the ring is free (previous round / earlier tiles have been read)
#pragma unroll
                for (int s = 0; s < RL_NSUB; ++s) {
                    const int off = s * RL_SUB + tid * RL_SEG;
                    const uint8_t* src = inbuf + 16 + off;
                    uint32_t bits = bm[s], kk = slot[s];
                    int prevpos = sprev[s];
                    while (bits) {
                        const int bi = __ffs(bits) - 1;
                        bits &= bits - 1;
                        if (kk >= w0 && kk < w0 + RL_RING && kk < total_emit) val8[kk - w0] = src[bi];
                        if (kk > w0 && kk <= w0 + RL_RING && kk <= total_emit) cnt8[kk - 1 - w0] = (uint8_t)(off + bi - prevpos);
                        prevpos = off + bi;
                        ++kk;
                    }
                }
                bar_sync(BAR_COMPUTE, RL_THREADS);
                const uint32_t cw = min((uint32_t)RL_RING, total_emit - w0);
                const bool tail = w0 + cw == total_emit && total_all == total_emit;  // that count is written by the next tile
                store_compacted(values, g0 + w0, cw, pair_cap, val8, 0, tid);
                store_compacted(counts, g0 + w0, cw - (tail ? 1u : 0u), pair_cap, cnt8, 0, tid);
            }
            n_stored = k + 1u;
            bar_sync(BAR_COMPUTE, RL_THREADS);  // the ring has been read
            ring_head = 0;
            if (lane == 0) mbar_arrive(&freeb);
        }
        if (warp == 0) TRACE(14);
    }
    // ---- the tiles that still wait
    {
        const uint32_t k_end = ld_volatile_shared(&sh.n_meta);  // == tiles taken by this CTA (every one of them passed p2)
        if (n_stored < k_end) {
            wait_ready(k_end);
            for (; n_stored < k_end; ++n_stored) store_tile(sh.q[n_stored % RLE_Q]);
        }
    }
}

// =============================================================================================
//                                        launcher
// =============================================================================================
// positions 0 .. n_scan are scanned, n_scan = min(n_avail, n + 255) being a forced boundary
size_t rl_encode_tiles(size_t n) { return (((n + 255) / RL_TILE + 1) + 255) / 256 * 256; }  // padded: the scanner works on whole chunks
size_t rl_encode_workspace_bytes(size_t n) { return (3 * rl_encode_tiles(n) + 2) * sizeof(uint64_t); }

cudaError_t rl_encode_launch(const uint8_t* d_in, size_t n, size_t n_avail, uint64_t prev_run, uint8_t* d_counts,
                             uint8_t* d_values, size_t pair_cap, uint64_t* d_pairs, void* d_ws, cudaStream_t stream)
{
    if (n == 0) return cudaMemsetAsync(d_pairs, 0, sizeof(uint64_t), stream);
    const size_t n_scan = std::min(n_avail, n + 255);
    const size_t tiles = n_scan / RL_TILE + 1, tmax = rl_encode_tiles(n);
    uint64_t* d1 = reinterpret_cast<uint64_t*>(d_ws);   // look-back #1 descriptors
    uint64_t* inc = d1 + tmax;                          // first slots (scanner -> tiles)
    uint32_t* agg = reinterpret_cast<uint32_t*>(inc + tmax);  // pair counts (tiles -> scanner)
    uint32_t* next_tile = agg + tmax;                   // tile counter
    cudaError_t e = cudaMemsetAsync(d_ws, 0, rl_encode_workspace_bytes(n), stream);
    if (e != cudaSuccess) return e;
    constexpr int smem = RL_INBUF + 2 * RL_RINGB;
    // the grid must be fully co-resident: the tiles wait for the scanner CTA, the scanner for every tile taken
    int dev = 0;
    e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    static int resident[64];
    int cap = dev < 64 ? resident[dev] : 0;
    if (!cap) {
        e = cudaFuncSetAttribute((const void*)rl_encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        int sms = 0, per_sm = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, (const void*)rl_encode_kernel, RLE_THREADS, smem);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) return cudaErrorLaunchOutOfResources;
        static const int want = getenv("FLRL_RLE_CTAS") ? atoi(getenv("FLRL_RLE_CTAS")) : RLE_CTAS_PER_SM;
        cap = sms * std::max(1, std::min(per_sm, want));
        if (cap < 2) return cudaErrorLaunchOutOfResources;
        if (dev < 64) resident[dev] = cap;
    }
    const unsigned grid = (unsigned)std::min<size_t>(tiles, (size_t)cap - 1) + 1;  // the scanner CTA + the workers
    rl_encode_kernel<<<grid, RLE_THREADS, smem, stream>>>(d_in, n, n_avail, n_scan, (uint32_t)(prev_run % 255u), prev_run != 0, d_counts,
                                                          d_values, pair_cap, d_pairs, d1, inc, agg, next_tile, (long long)tiles);
    return cudaGetLastError();
}

}  // namespace flrl

#ifdef FLRL_TRACE
extern "C" int flrl_debug_set_rle_trace(void* d_buf)
{
    return (int)cudaMemcpyToSymbol(flrl::g_rle_trace, &d_buf, sizeof(d_buf));
}
#endif
