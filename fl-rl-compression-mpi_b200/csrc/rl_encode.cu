// rl_encode.cu — run-length encoder for sm_100a: one launch, input read once.
//
// The reference documents RL (IMPLEMENTATION-PLAN.md:83-152: mask -> scan -> start indices -> split of
// runs > 255 -> emit) but ships no code.  Structure here:
//
//   * persistent CTAs (6 per SM) take 16 KiB tiles from an atomic counter, so tile order is time order:
//     everything a tile depends on was started earlier, by a CTA that never waits for a later tile.  The CTA
//     that runs first becomes the grid's scanner (role counter), so nothing depends on CTA placement or on
//     the whole grid being resident;
//   * a tile is 128 ROWS of 128 bytes, loaded by TMA as a 2-D tensor with the 128-byte swizzle (four boxes of
//     32 rows, one mbarrier each; one tile buffer per CTA — measured: six CTAs with one buffer beat four with two):
//     a thread owns one row and reads it with eight conflict-free LDS.128 (16-byte piece j of row r sits at
//     piece j ^ (r & 7)); a warp owns 4 KiB of consecutive bytes;
//   * per row: run-start bit mask of its 128 bytes (XOR with the byte-shifted word, SWAR non-zero-byte test,
//     IDP.4A gathers the byte flags into mask bits) -> nearest earlier run start S inside the warp by ballot ->
//     the one possible 255-split of the row, (i - S) % 255 == 0 (the plan's ":125" rule, replacing its
//     iterative rescan loop) -> ONE warp scan of the pair counts.  The leading rows of a warp (no run start
//     before them inside the warp) get their splits in closed form once the previous warps' last run starts
//     are known (one named barrier per tile); the leading part of the TILE gets them from the service warp's
//     look-back, also in closed form;
//   * compaction into a shared-memory RING of staged pairs: the value and the POSITION (mod 256) of every pair
//     boundary.  A pair's count is the difference of two neighbouring positions (boundaries are never more
//     than 255 apart), taken when the tile is written: no look-ahead halo, no suffix scan, no per-pair count
//     arithmetic in the divergent loop;
//   * the compute warps never wait for another CTA: a tile's pairs stay in the ring until its first slot is
//     known and are written then, a few tiles later; 16-byte stores aligned in the GLOBAL address space
//     (unaligned shared reads fixed up with PRMT);
//   * a service warp per CTA: tile ids, loads, look-back #1 (start of the run that enters the tile: gives the
//     splits that fall before the tile's first run start and the count of the PREVIOUS tile's last pair,
//     which this tile writes), posts the tile's pair count;
//   * one scanner warp for the whole grid turns the posted pair counts into first slots, in tile order (poll
//     traffic O(tiles), where per-tile decoupled look-backs read O(tiles x window)); a base warp per CTA polls
//     the slots of the CTA's waiting tiles, one lane per tile.
//
// A pair belongs to the tile (and the shard) in which it STARTS; its count looks ahead up to 255 bytes, so
// tiles and shards concatenate to the single-stream output byte for byte.  tools/model_rle.py is the same
// arithmetic on the CPU, checked against the oracle.
#include <cuda.h>
#include <stdlib.h>

#include <algorithm>

#include "flrl_bits.cuh"
#include "lookback.cuh"
#include "rl_kernels.cuh"
#include "rl_scan.cuh"

namespace flrl {

#ifndef RL_SEG_N
#define RL_SEG_N 128
#endif
static constexpr int RL_ROW = 128;                      // bytes per row of the tile (the swizzle span)
static constexpr int RL_ROWS = 128;                     // rows per tile
static constexpr int RL_TILE = RL_ROW * RL_ROWS;        // 16 KiB
static constexpr int RL_BOXES = 4;                      // TMA boxes per tile
static constexpr int RL_BOX = RL_TILE / RL_BOXES;       // 4 KiB = 32 rows
static constexpr int RL_SEG = RL_SEG_N;                 // consecutive bytes per compute thread and tile: a row (128) or half a row (64)
static constexpr int RL_NW = RL_SEG / 32;               // 32-bit mask words per thread
static constexpr int RL_THREADS = RL_TILE / RL_SEG;     // compute threads
static constexpr int RL_WARPS = RL_THREADS / 32;        // compute warps
static constexpr int RLE_THREADS = RL_THREADS + 32;     // + service warp
static constexpr int RL_WSPAN = 32 * RL_SEG;            // consecutive bytes per warp
#ifndef RL_NBUF_N
#define RL_NBUF_N 1
#endif
static constexpr int RL_NBUF = RL_NBUF_N;               // tile buffers per CTA (2: the next tile lands while this one is worked on; 4 CTAs per SM then)
#ifndef RL_RING_PAIRS
#define RL_RING_PAIRS 8192
#endif
// pause between two looks of a polling loop (ns)
#ifndef RLE_SLEEP_A
#define RLE_SLEEP_A 200
#endif
#ifndef RLE_PF_LEAD
#define RLE_PF_LEAD 0  // tiles between a load and an L2 prefetch issued with it (0: none; measured: no gain with two buffers)
#endif
#ifndef RLE_CTAS
#define RLE_CTAS 6
#endif
#ifndef RLE_FLAT
#define RLE_FLAT 1  // skip the mask arithmetic of 32-byte pieces that only repeat the previous byte (warp vote)
#endif
static constexpr int RL_RING = RL_RING_PAIRS;           // staged pairs (power of two)
static constexpr int RL_RINGB = RL_RING + 32;           // + mirror of entry 0 + over-read slack of the unaligned 16-byte reads
static constexpr int RLE_Q = 32;                        // tiles of one CTA that may wait for their first slot
static constexpr int RLE_CTAS_PER_SM = RLE_CTAS;
static constexpr uint64_t RL_BIAS = 255;                // run starts are stored as S + 255 >= 0
static constexpr uint32_t BAR_COMPUTE = 1;              // named barrier of the compute warps
static_assert((RL_RING & (RL_RING - 1)) == 0 && RL_RING >= 512, "ring indices are masked");
static_assert(RL_ROW == 128 && (RL_SEG == 128 || RL_SEG == 64) && RL_BOX % RL_WSPAN == 0, "a thread owns a row or half a row; a warp stays inside one box");

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
            __nanosleep(RLE_SLEEP_A);
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
// 16 counts from 17 staged positions: pos[u + i + 1] - pos[u + i] (mod 256), i = 0 .. 15
__device__ __forceinline__ uint4 lds_diff16(const uint8_t* buf, uint32_t u)
{
    const uint32_t* w = reinterpret_cast<const uint32_t*>(buf) + (u >> 2);
    const uint32_t s0 = 0x3210u + 0x1111u * (u & 3u), s1 = s0 + 0x1111u;  // shift 4 selects the whole next word
    const uint32_t a = w[0], b = w[1], c = w[2], d = w[3], e = w[4];
    return make_uint4(__vsub4(__byte_perm(a, b, s1), __byte_perm(a, b, s0)), __vsub4(__byte_perm(b, c, s1), __byte_perm(b, c, s0)),
                      __vsub4(__byte_perm(c, d, s1), __byte_perm(c, d, s0)), __vsub4(__byte_perm(d, e, s1), __byte_perm(d, e, s0)));
}

// cnt output bytes derived from stage[soff ..] -> dst[g0 .. g0 + cnt): coalesced 16-byte stores on 16-byte aligned
// global addresses, the (at most 15 + 15) bytes of the two ragged ends as one byte store per thread; nothing at
// or past index `cap` is written.  DIFF: the output byte i is stage[soff + i + 1] - stage[soff + i] (a count from
// two positions), else stage[soff + i] (a value).
template <bool DIFF>
__device__ __forceinline__ void store_compacted(uint8_t* __restrict__ dst, uint64_t g0, uint32_t cnt, uint64_t cap,
                                                const uint8_t* stage, uint32_t soff, int tid, int nthr = RL_THREADS)
{
    if (cap <= g0) return;
    if (cap - g0 < cnt) cnt = (uint32_t)(cap - g0);
    const uint32_t a = (uint32_t)((reinterpret_cast<uintptr_t>(dst) + g0) & 15u);
    const uint32_t head = min(cnt, (16u - a) & 15u);        // bytes before the first aligned chunk
    const uint32_t nfull = (cnt - head) >> 4;               // aligned 16-byte chunks
    const uint32_t tail0 = head + 16u * nfull;              // staging index of the first byte after them
    for (uint32_t q = tid; q < nfull; q += nthr)
        *reinterpret_cast<uint4*>(dst + g0 + head + 16u * q) =
            DIFF ? lds_diff16(stage, soff + head + 16u * q) : lds_unaligned16(stage, soff + head + 16u * q);
    if ((uint32_t)tid < head) dst[g0 + tid] = DIFF ? (uint8_t)(stage[soff + tid + 1] - stage[soff + tid]) : stage[soff + tid];
    const uint32_t t2 = (uint32_t)tid - 16u;                // threads 16..31 take the tail
    if (t2 < cnt - tail0)
        dst[g0 + tail0 + t2] = DIFF ? (uint8_t)(stage[soff + tail0 + t2 + 1] - stage[soff + tail0 + t2]) : stage[soff + tail0 + t2];
}
// the same from a ring of RL_RING bytes: at most two linear pieces (entry RL_RING mirrors entry 0 for the differences)
template <bool DIFF>
__device__ __forceinline__ void store_from_ring(uint8_t* __restrict__ dst, uint64_t g0, uint32_t cnt, uint64_t cap,
                                                const uint8_t* ring, uint32_t head, int tid, int nthr)
{
    const uint32_t h = head & (RL_RING - 1);
    const uint32_t first = min(cnt, (uint32_t)RL_RING - h);
    store_compacted<DIFF>(dst, g0, first, cap, ring, h, tid, nthr);
    if (cnt > first) store_compacted<DIFF>(dst, g0 + first, cnt - first, cap, ring, 0, tid, nthr);
}

template <typename T>
__device__ __forceinline__ T dmin(T a, T b) { return a < b ? a : b; }
template <typename T>
__device__ __forceinline__ T dmax(T a, T b) { return a > b ? a : b; }

// 2-D tensor load of one box (128 bytes x 32 rows, 128-byte swizzle) starting at row `row`
__device__ __forceinline__ void tma_load_rows(void* smem_dst, const CUtensorMap* tmap, int row, uint64_t* bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(reinterpret_cast<uint64_t>(tmap)), "r"(0), "r"(row), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_only(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// byte i of row r of a swizzled tile buffer
__device__ __forceinline__ uint32_t swz(uint32_t r, uint32_t i) { return r * RL_ROW + ((((i >> 4) ^ r) & 7u) << 4) + (i & 15u); }

struct RlePending {               // a tile of this CTA, from "taken" to "stored"
    // written by the compute warps
    uint32_t head;                 // ring position of the tile's first staged pair
    uint32_t total;                // staged pairs that are emitted (start before n)
    uint32_t total_all;            // staged pair boundaries (after the tile's first run start), emitted or not
    uint32_t first_byte;           // in[tile base]: the value of the run that enters the tile
    long long tile;
    uint64_t g0;                   // first pair slot of the tile (the scanner's answer)
    uint32_t skip;                 // 1: more boundaries than the ring holds — the tile is left to the second pass
    // written by the service warp BEFORE it posts the tile's pair count: whoever sees the scanner's answer sees these
    uint32_t tot;                  // pairs of the tile (staged + nl)
    uint32_t nl;                   // emitted pairs that start at 255-splits before the tile's first run start
    uint32_t nl_all;               // the same splits, emitted or not
    uint32_t lead_first;           // tile-relative position of the first such split
    uint32_t first_start;          // tile-relative position of the tile's first run start (RL_TILE if none)
    uint32_t lagged;               // count of the last pair before the tile; 0 = not ours to write
    uint32_t posted;               // tile's sequence number + 1 in this CTA once the six words above are written
};

struct RleMeta {                   // compute warps -> service warp, per tile
    int first_start;               // first run start of the tile (tile relative), RL_TILE if none
    int last_start;                // last run start, -1 if none
    uint32_t total_all;            // staged boundaries
    uint32_t total_emit;           // those before n
};

struct RleShared {
    // per warp, double-buffered by tile parity (one named barrier per tile separates writers from readers)
    __align__(16) int wfirst[2][RL_WARPS];     // first run start in the warp's 4 KiB (tile relative), RL_TILE if none
    __align__(16) int wlast[2][RL_WARPS];      // last run start, -1 if none
    __align__(16) uint32_t wcnt[2][RL_WARPS];  // boundaries found by the warp's rows: all | those before n << 16
    __align__(16) uint8_t prev16[RL_NBUF][16]; // the 16 bytes before the tile (tile 0: unused)
    uint32_t store_upto[2];        // tiles of this CTA whose first slot is known (snapshot shared by the compute warps)
    RleMeta meta[2];
    long long cur_tile[RL_NBUF];   // service warp -> compute warps (-1: no more tiles)
    uint32_t role;                 // 0: this CTA is the grid's scanner
    RlePending q[RLE_Q];
};

// dynamic shared memory: [pad to 1 KiB][RL_NBUF swizzled tile buffers][ring of values][ring of positions][RleShared]
static constexpr int RLE_SMEM = 1024 + RL_NBUF * RL_TILE + 2 * RL_RINGB + (int)sizeof(RleShared);
__device__ __forceinline__ uint8_t* rle_smem()
{
    extern __shared__ uint8_t smem_raw[];
    return smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
}
__device__ __forceinline__ uint8_t* rle_val8() { return rle_smem() + RL_NBUF * RL_TILE; }
__device__ __forceinline__ uint8_t* rle_pos8() { return rle_val8() + RL_RINGB; }
__device__ __forceinline__ RleShared& rle_shared() { return *reinterpret_cast<RleShared*>(rle_pos8() + RL_RINGB); }

// Writes a tile whose first slot is known (compute threads, all of them): the count the previous tile left to this one,
// the pairs at 255-splits before its first run start (long runs only), then its staged pairs.  The tile's last count is
// left to the next tile (it knows where the next boundary is) unless a boundary past n is staged behind it.
// One copy of this code in the kernel: the tile loop has to stay small (instruction cache).  Called by `nthr` threads
// numbered tid = 0 .. nthr-1: all compute threads, or one warp for a tile with few pairs (the routine's fixed cost is
// then paid once, not four times).
__device__ __noinline__ void rle_store_tile(uint32_t qi, uint8_t* __restrict__ counts, uint8_t* __restrict__ values, size_t pair_cap,
                                            uint64_t* __restrict__ d_pairs, long long tiles, int tid, int nthr)
{
    const RlePending& e = rle_shared().q[qi];
    if (e.skip) return;  // nothing of it is written here: the second pass does the whole tile
    const uint32_t nl = e.nl;
    const uint64_t g0 = e.g0;
    if (tid == 0) {
        if (e.lagged && g0 >= 1 && g0 - 1 < pair_cap) counts[g0 - 1] = (uint8_t)e.lagged;
        if (e.tile == tiles - 1) *d_pairs = g0 + e.tot;
    }
    const bool lead_lag = e.nl_all == nl && e.total_all == 0;  // the last split is the tile's last boundary
    for (uint32_t j = tid; j < nl; j += nthr) {
        const uint32_t p = e.lead_first + 255u * j;
        const uint64_t kk = g0 + j;
        if (kk < pair_cap) {
            values[kk] = (uint8_t)e.first_byte;
            if (!(j == nl - 1 && lead_lag)) counts[kk] = (uint8_t)min(255u, e.first_start - p);
        }
    }
    if (e.total) {
        store_from_ring<false>(values, g0 + nl, e.total, pair_cap, rle_val8(), e.head, tid, nthr);
        store_from_ring<true>(counts, g0 + nl, e.total_all > e.total ? e.total : e.total - 1u, pair_cap, rle_pos8(), e.head, tid, nthr);
    }
}

__global__ void __launch_bounds__(RLE_THREADS, RLE_CTAS_PER_SM)
rl_encode_kernel(const __grid_constant__ CUtensorMap tmap, const uint8_t* __restrict__ in, size_t n, size_t n_avail, size_t n_scan,
                 uint32_t prev_phase, int has_prev, uint8_t* __restrict__ counts, uint8_t* __restrict__ values, size_t pair_cap,
                 uint64_t* __restrict__ d_pairs, uint64_t* __restrict__ d1, uint64_t* __restrict__ inc, uint32_t* __restrict__ agg,
                 uint32_t* __restrict__ next_tile, uint32_t* __restrict__ role, uint32_t* __restrict__ redo, long long tiles, int pass)
{
    // pass 0: every tile; a tile with more pair boundaries than the ring of staged pairs holds is only COUNTED (its pairs
    // take part in the scan) and put on the `redo` list.  pass 1 (a second launch, right behind): the tiles of that list,
    // with every first slot and every run start already final, written directly in rounds of one ring.  Nothing ever
    // waits for its own first slot while it holds a tile other tiles wait for: dense input (random bytes: 16384 pairs
    // per tile) costs a second read of those tiles, not a serialised grid (measured before: 170 ms per GiB).
    uint32_t* const redo_n = role + 1;     // pass 0: tiles on the list; pass 1: read only
    uint32_t* const redo_next = role + 2;  // pass 1: next list entry to take
    if (pass) {
        // launched with programmatic stream serialization: the launch itself overlaps the first pass; its results are
        // visible after this wait
        asm volatile("griddepcontrol.wait;" ::: "memory");
        if (*reinterpret_cast<volatile uint32_t*>(redo_n) == 0) return;  // the usual case: nothing was left over
    }
    uint8_t* const inb = rle_smem();                   // swizzled tile buffers
    uint8_t* const val8 = rle_val8();                  // ring of staged values
    uint8_t* const pos8 = rle_pos8();                  // ring of staged positions (mod 256)
    RleShared& sh = rle_shared();
    __shared__ __align__(8) uint64_t full[RL_NBUF][RL_BOXES], freeb[RL_NBUF], meta_bar[2], aux;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        sh.role = pass == 0 ? atomicAdd(role, 1u) : 1u;  // the second pass needs no scanner
#pragma unroll
        for (int b = 0; b < RL_NBUF; ++b) {
#pragma unroll
            for (int s = 0; s < RL_BOXES; ++s) mbar_init(&full[b][s], 1);
            mbar_init(&freeb[b], RL_WARPS);
        }
        mbar_init(&meta_bar[0], 1);
        mbar_init(&meta_bar[1], 1);
        mbar_init(&aux, 1);
        fence_mbar_init();
        for (int i = 0; i < RLE_Q; ++i) sh.q[i].posted = 0;
    }
    __syncthreads();
    if (sh.role == 0) {  // the first CTA that runs is the scanner
        if (warp == 0) scan_pair_counts(agg, inc, tiles, lane);
        return;
    }

    if (warp == RL_WARPS) {
        // ================================ service warp ===========================================
        const long long rows_full = (long long)(n_avail / RL_ROW);  // rows the tensor map covers
        const uint32_t tail_bytes = (uint32_t)(n_avail % RL_ROW);   // bytes of the partial row behind them (plain loads)
        uint32_t aux_par = 0;
        // starts the loads of tile `tile` into buffer b (tile >= tiles: the end marker).  Nothing here waits for
        // global memory: the 16 bytes before the tile come by bulk copy on the first box's mbarrier.
        auto start_tile = [&](long long tile, int b) {
            if (tile >= tiles) {
                if (lane == 0) sh.cur_tile[b] = -1;
                __syncwarp();
                if (lane < RL_BOXES) mbar_arrive(&full[b][lane]);
                return;
            }
            uint8_t* const buf = inb + b * RL_TILE;
            const size_t tbase = (size_t)tile * RL_TILE;
            const long long row0 = tile * RL_ROWS;
            if (lane == 0) sh.cur_tile[b] = tile;
            __syncwarp();
            // the box that holds the partial last row of the data (at most one per shard) goes through `aux`
            const bool tail_here = tail_bytes != 0 && rows_full >= row0 && rows_full < row0 + RL_ROWS;
            const int tail_part = tail_here ? (int)((rows_full - row0) >> 5) : -1;
            if (lane < RL_BOXES) {
                const bool have = rows_full > row0 + 32 * lane;  // at least one full row in this box (rows past the end read as 0)
                if (lane == tail_part) {
                    if (have) {
                        mbar_expect_tx(&aux, RL_BOX);
                        tma_load_rows(buf + lane * RL_BOX, &tmap, (int)(row0 + 32 * lane), &aux);
                    }
                } else if (have) {
                    if (lane == 0) mbar_expect_tx_only(&full[b][0], RL_BOX);  // lane 0 arrives below
                    else mbar_expect_tx(&full[b][lane], RL_BOX);
                    tma_load_rows(buf + lane * RL_BOX, &tmap, (int)(row0 + 32 * lane), &full[b][lane]);
                } else if (lane != 0) {
                    mbar_arrive(&full[b][lane]);
                }
                if (lane == 0 && tile > 0) {  // the 16 bytes before the tile (its last byte is what the first row needs)
                    mbar_expect_tx_only(&full[b][0], 16);
                    tma_load_1d(sh.prev16[b], in + tbase - 16, 16, &full[b][0]);
                }
#if RLE_PF_LEAD
                // make the load of the tile one grid of tiles ahead an L2 hit (somebody else's tile)
                {
                    const size_t pb = tbase + (size_t)RLE_PF_LEAD * RL_TILE + (size_t)lane * RL_BOX;
                    const size_t pe = dmin(pb + RL_BOX, n_avail & ~(size_t)15);
                    if (pe > pb) tma_prefetch_l2(in + pb, (uint32_t)(pe - pb));
                }
#endif
            }
            if (tail_here) {  // warp-uniform
                if (rows_full > row0 + 32 * tail_part) {
                    mbar_wait(&aux, aux_par);  // the box has landed (its rows past the end as zeros): now the partial row
                    aux_par ^= 1u;
                }
                const uint32_t r = (uint32_t)(rows_full - row0);
                for (uint32_t i = lane; i < tail_bytes; i += 32) buf[swz(r, i)] = in[(size_t)rows_full * RL_ROW + i];
                __syncwarp();
                if (lane == 0 && tail_part != 0) mbar_arrive(&full[b][tail_part]);
            }
            if (lane == 0) mbar_arrive(&full[b][0]);
        };
        // One id per atomic, also at the start: a CTA that held two CONSECUTIVE ids would make the next CTA's tile wait
        // (for its first slot) until this CTA has finished a whole tile and counted the second one, and so on down the
        // grid — measured: tiles with more pairs than the ring holds then run one after the other (170 ms per GiB).
        const uint32_t n_redo = pass ? *redo_n : 0u;
        auto take = [&]() -> long long {  // lane 0: the next tile id (>= tiles: none left)
            if (pass == 0) return (long long)atomicAdd(next_tile, 1u);
            const uint32_t i = atomicAdd(redo_next, 1u);
            return i < n_redo ? (long long)redo[i] : tiles;
        };
        long long cur = 0, nxt = 0, nxt2 = 0;
        if (lane == 0) cur = take();
        cur = __shfl_sync(0xffffffffu, cur, 0);
        start_tile(cur, 0);
        if (RL_NBUF > 1) {
            if (lane == 0) nxt = take();
            nxt = __shfl_sync(0xffffffffu, nxt, 0);
        }
        for (uint32_t k = 0;; ++k) {
            const long long tile = cur;
            if (tile >= tiles) break;
            const uint32_t kb = k & 1u;
            const size_t tbase = (size_t)tile * RL_TILE;
            const uint32_t rel_emit = n > tbase ? (uint32_t)dmin((size_t)(n - tbase), (size_t)RL_TILE) : 0u;
            if (RL_NBUF > 1) {
                // the next tile's loads go first: a look-back that has to wait must not hold them up (it would delay this
                // CTA's next tile, whose run start other look-backs wait for in turn)
                if (k) mbar_wait_sleep(&freeb[(k - 1) % RL_NBUF], ((k - 1) / RL_NBUF) & 1u);  // tile k-1's buffer takes tile k+1
                start_tile(nxt, (int)((k + 1) % RL_NBUF));
            }
            uint64_t sbB = prev_phase ? RL_BIAS - prev_phase : 0;   // biased start of the run that enters the tile
            if (tile > 0) sbB = lookback_run_start(d1, tile, lane);
            mbar_wait_sleep(&meta_bar[kb], (k >> 1) & 1u);  // the compute warps have counted the tile
            const int F_t = sh.meta[kb].first_start;
            const int L_t = sh.meta[kb].last_start;
            const uint32_t total_emit = sh.meta[kb].total_emit;
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
            const uint32_t tot = nl + total_emit;
            if (lane == 0) {
                RlePending& e = sh.q[k % RLE_Q];
                e.tot = tot;
                e.nl = nl;
                e.nl_all = nl_all;
                e.lead_first = (uint32_t)(BpB + 255u - tB);
                e.first_start = (uint32_t)F_t;
                // the previous tile leaves its last pair's count to us when that pair is the last boundary before this
                // tile (it cannot know where the next boundary is); a boundary at or past n is not a pair
                e.lagged = (tile > 0 && BpB < n + RL_BIAS) ? lagged : 0u;
                __threadfence_block();  // ... before the post: the scanner's answer implies the six words above
                *reinterpret_cast<volatile uint32_t*>(&e.posted) = k + 1u;
                if (pass == 0) st_relaxed_u32(agg + tile, 0x80000000u | tot);  // to the scanner
            }
            __syncwarp();
            // the id of this CTA's tile k+2, taken as late as possible (tiles should start in id order: a tile's first slot
            // waits for every earlier tile) yet early enough that the atomic's round trip hides behind the next wait
            if (lane == 0) nxt2 = take();
            if (RL_NBUF == 1) {  // one buffer: the id just taken is the next tile (never hold an id whose load cannot start)
                mbar_wait_sleep(&freeb[0], k & 1u);  // the compute warps are done with the tile: the buffer takes the next one
                nxt = __shfl_sync(0xffffffffu, nxt2, 0);
                start_tile(nxt, 0);
                cur = nxt;
                continue;
            }
            cur = nxt;
            nxt = __shfl_sync(0xffffffffu, nxt2, 0);
        }
        return;
    }

    // ================================= compute warps =======================================
    const uint32_t lt_mask = (1u << lane) - 1u;
    const int rowstart = tid * RL_SEG;            // tile-relative position of this thread's bytes (a row or half a row)
    const uint32_t sw = (uint32_t)(rowstart / RL_ROW) & 7u;          // swizzle of the row they are in
    const uint32_t chunk0 = (uint32_t)(rowstart % RL_ROW) >> 4;      // first 16-byte piece of the row that is ours
    const int mybox = rowstart / RL_BOX;
    uint32_t n_stored = 0;   // tiles of this CTA already written to global memory
    uint32_t ring_head = 0;  // ring position after the last staged boundary (not masked)

    auto store_tile = [&](uint32_t i) { rle_store_tile(i % RLE_Q, counts, values, pair_cap, d_pairs, tiles, tid, RL_THREADS); };
    // a tile with few pairs is written by ONE warp (they take turns); the others go on.  Safe: the ring space it reads is
    // not reused before the next tile's barrier, which that warp reaches only after the store.
    auto store_tile_light = [&](uint32_t i, uint32_t turn) {
        const RlePending& e = sh.q[i % RLE_Q];
        if (e.total + e.nl > 1024u) rle_store_tile(i % RLE_Q, counts, values, pair_cap, d_pairs, tiles, tid, RL_THREADS);
        else if ((uint32_t)warp == ((turn + i) & (RL_WARPS - 1))) rle_store_tile(i % RLE_Q, counts, values, pair_cap, d_pairs, tiles, lane, 32);
    };
    // blocks until the first slots of this CTA's tiles [n_stored, need) are known (the scanner has answered them)
    auto wait_slots = [&](uint32_t need) {
        if (tid == 0) {
            for (uint32_t i = n_stored; i < need; ++i) {
                uint64_t g;
                while (*reinterpret_cast<volatile uint32_t*>(&sh.q[i % RLE_Q].posted) != i + 1u) __nanosleep(50);
                while (!((g = ld_relaxed_u64(inc + sh.q[i % RLE_Q].tile)) >> 63)) __nanosleep(50);
                sh.q[i % RLE_Q].g0 = g & ~(1ull << 63);
            }
            __threadfence_block();  // the answer follows the service warp's post, which follows its part of the entry
        }
        bar_sync(BAR_COMPUTE, RL_THREADS);
    };

    [[maybe_unused]] bool try_flat = false;  // the previous tile of this CTA had few boundaries
    uint32_t k = 0;  // tiles of this CTA that have been staged
    for (;; ++k) {
        const uint32_t kb = k & 1u, buf_i = k % RL_NBUF, fpar = (k / RL_NBUF) & 1u;
        if (k - n_stored >= (uint32_t)RLE_Q - 1u) {  // the queue of waiting tiles is full (rare)
            wait_slots(n_stored + 1u);
            store_tile(n_stored++);
        }
        // ask for the first slots of the two oldest waiting tiles: the answers arrive while the masks are computed
        uint64_t ga = 0, gb = 0;
        if (tid == 0) {
            if (n_stored < k) ga = ld_relaxed_u64(inc + sh.q[n_stored % RLE_Q].tile);
            if (n_stored + 1u < k) gb = ld_relaxed_u64(inc + sh.q[(n_stored + 1u) % RLE_Q].tile);
        }
        if (mybox > 0) mbar_wait(&full[buf_i][mybox - 1], fpar);  // the byte before the warp's first one may be in the previous box
        mbar_wait(&full[buf_i][mybox], fpar);
        const long long tile = *reinterpret_cast<const volatile long long*>(&sh.cur_tile[buf_i]);
        if (tile < 0) break;
        const uint8_t* const buf = inb + buf_i * RL_TILE;
        const uint8_t* const rowp = buf + (rowstart / RL_ROW) * RL_ROW;  // the row (128 bytes, swizzled pieces)
        const size_t tbase = (size_t)tile * RL_TILE;
        // tile-relative limits: positions >= rel_emit are not emitted; position rel_scan is a forced boundary (the end
        // of the data, or far enough into the look-ahead that every emitted pair has its count), nothing after it counts
        const uint32_t rel_emit = n > tbase ? (uint32_t)dmin((size_t)(n - tbase), (size_t)RL_TILE) : 0u;
        const uint32_t rel_scan = (uint32_t)dmin((size_t)(n_scan - tbase), (size_t)RL_TILE);  // == RL_TILE: not in this tile
        const bool ragged = rel_emit < (uint32_t)RL_TILE || rel_scan < (uint32_t)RL_TILE;

        // ---- step A: the row's run-start mask, its 255-split, pair counts; nothing from other warps is needed
        uint32_t M[RL_NW];
        {
            // the word that ends with the byte before the row (only its top byte is used)
            // ("before" the stream: a byte that differs from in[0] unless a run is open)
            const uint32_t pp = (uint32_t)(rowstart - 16);  // the piece before ours (tid > 0): row pp / 128, piece (pp % 128) / 16
            uint32_t pw = tid > 0    ? *reinterpret_cast<const uint32_t*>(buf + (pp / RL_ROW) * RL_ROW + (((((pp % RL_ROW) >> 4) ^ (pp / RL_ROW)) & 7u) << 4) + 12)
                          : tile > 0 ? *reinterpret_cast<const volatile uint32_t*>(&sh.prev16[buf_i][12])
                                     : (uint32_t)(has_prev ? buf[0] : (uint8_t)~buf[0]) << 24;
#pragma unroll
            for (int g = 0; g < RL_NW; ++g) {
                const uint4 a = *reinterpret_cast<const uint4*>(rowp + (((chunk0 + 2u * g) ^ sw) << 4));
                const uint4 c = *reinterpret_cast<const uint4*>(rowp + (((chunk0 + 2u * g + 1u) ^ sw) << 4));
                const uint32_t w[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
#if RLE_FLAT
                // long runs: most 32-byte pieces only repeat the byte before them.  When that holds for the piece of
                // every lane (a warp vote: no divergence) the mask is 0 and its ~50 instructions are skipped for 12.
                // Tried only after a tile with few boundaries (CTA-uniform), so short-run data never pays for the test.
                if (try_flat) {
                    const uint32_t pb = __byte_perm(pw, 0, 0x3333);  // the previous byte in all four lanes of a word
                    const uint32_t diff = ((a.x ^ pb) | (a.y ^ pb)) | ((a.z ^ pb) | (a.w ^ pb)) | ((c.x ^ pb) | (c.y ^ pb)) | ((c.z ^ pb) | (c.w ^ pb));
                    if (__all_sync(0xffffffffu, diff == 0)) {
                        M[g] = 0;
                        continue;  // pw stays: c.w ends with the same byte
                    }
                }
#endif
                M[g] = start_mask32(w, pw);
                pw = c.w;
            }
        }
        if (ragged) {  // CTA-uniform: only the last tiles see the end of the data
#pragma unroll
            for (int g = 0; g < RL_NW; ++g) {
                const int kk = (int)rel_scan - (rowstart + 32 * g);
                if (kk < 0) M[g] = 0;
                else if (kk < 32) M[g] = (M[g] & ((1u << kk) - 1u)) | (1u << kk);
            }
        }
        // bytes of the row that continue the entering run / last run start of the row (tile relative)
        const unsigned long long mlo = (unsigned long long)M[0] | ((unsigned long long)M[1] << 32),
                                 mhi = RL_NW == 4 ? (unsigned long long)M[RL_NW - 2] | ((unsigned long long)M[RL_NW - 1] << 32) : 0ull;
        const bool any = (mlo | mhi) != 0;
        const int lead = mlo ? __ffsll((long long)mlo) - 1 : mhi ? 63 + __ffsll((long long)mhi) : RL_SEG;
        const unsigned bal = __ballot_sync(0xffffffffu, any);
        const int my_last = rowstart + (mhi ? 127 - __clzll((long long)mhi) : 63 - __clzll((long long)mlo));
        const unsigned lower = bal & lt_mask;
        const int S = __shfl_sync(0xffffffffu, my_last, 31 - __clz(lower | 1u));  // nearest run start before the row, if lower != 0
        if (lower != 0 && (rowstart - S) + lead > 255 && !(ragged && rowstart > (int)rel_scan)) {
            // boundaries of the entering run lie at S + 255 j; rr = distance from the last one to rowstart - 1
            const uint32_t rr = (uint32_t)(rowstart - 1 - S) % 255u;
            const uint32_t j = 254u - rr;
            if (j < (uint32_t)lead) {
#pragma unroll
                for (int g = 0; g < RL_NW; ++g)
                    if ((j >> 5) == (uint32_t)g) M[g] |= 1u << (j & 31u);
            }
        }
        uint32_t packed = __popc(M[0]) + __popc(M[1]) + (RL_NW == 4 ? __popc(M[RL_NW - 2]) + __popc(M[RL_NW - 1]) : 0);  // <= 128; a warp's sum <= 4096
        if (ragged) {
            uint32_t ce = 0;
#pragma unroll
            for (int g = 0; g < RL_NW; ++g) {
                const int kk = (int)rel_emit - (rowstart + 32 * g);
                const uint32_t em = kk <= 0 ? 0u : (kk >= 32 ? ~0u : ((1u << kk) - 1u));
                ce += __popc(M[g] & em);
            }
            packed |= ce << 16;
        } else {
            packed |= packed << 16;
        }
        uint32_t incl = packed;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += o;
        }
        const uint32_t excl = (incl - packed) & 0xffffu;  // boundaries of the warp's earlier rows
        if (lane == 31) sh.wcnt[kb][warp] = incl;
        if (bal) {
            if (lane == __ffs(bal) - 1) sh.wfirst[kb][warp] = rowstart + lead;  // this row has no split: lead is its first run start
            if (lane == 31 - __clz(bal)) {
                sh.wlast[kb][warp] = my_last;
                // the next tile's look-back #1 only needs this tile's last run start: published the moment it is known
                if (warp == RL_WARPS - 1) st_relaxed_u64(d1 + tile, LB_INC | (uint64_t)(tbase + my_last + RL_BIAS));
            }
        } else if (lane == 0) {
            sh.wfirst[kb][warp] = RL_TILE;
            sh.wlast[kb][warp] = -1;
        }
        if (tid == 0) {  // which of the waiting tiles can be written after this tile's compaction
            uint32_t r = n_stored;
            if (ga >> 63) {
                sh.q[r++ % RLE_Q].g0 = ga & ~(1ull << 63);
                if (gb >> 63) sh.q[r++ % RLE_Q].g0 = gb & ~(1ull << 63);
                __threadfence_block();  // the answer follows the service warp's post, which follows its part of the entry
            }
            sh.store_upto[kb] = r;
        }
        bar_sync(BAR_COMPUTE, RL_THREADS);
        uint32_t upto = sh.store_upto[kb];

        // ---- step B: every warp derives the same tile layout from the four warps' (first, last, count):
        // the splits in a warp's leading rows (closed form from the previous warps' last run start), bases, totals
        uint32_t total_all = 0, total_emit = 0, mybase = 0, mynl = 0, myfirst = 0;
        int F_t, L_t, L_lastwarp;
        {
            int Fa[RL_WARPS], La[RL_WARPS];
            uint32_t Ca[RL_WARPS];
#pragma unroll
            for (int i = 0; i < RL_WARPS; i += 4) {
                const int4 Fv = *reinterpret_cast<const int4*>(&sh.wfirst[kb][i]);
                const int4 Lv = *reinterpret_cast<const int4*>(&sh.wlast[kb][i]);
                const uint4 Cv = *reinterpret_cast<const uint4*>(&sh.wcnt[kb][i]);
                Fa[i] = Fv.x, Fa[i + 1] = Fv.y, Fa[i + 2] = Fv.z, Fa[i + 3] = Fv.w;
                La[i] = Lv.x, La[i + 1] = Lv.y, La[i + 2] = Lv.z, La[i + 3] = Lv.w;
                Ca[i] = Cv.x, Ca[i + 1] = Cv.y, Ca[i + 2] = Cv.z, Ca[i + 3] = Cv.w;
            }
            int Srun = -1;
#pragma unroll
            for (int i = 0; i < RL_WARPS; ++i) {
                const int ws = RL_WSPAN * i;
                uint32_t nl_i = 0, nle_i = 0, first_i = 0;
                if (i > 0 && Srun >= 0 && !(ragged && ws > (int)rel_scan)) {
                    const int end_i = min(Fa[i], ws + RL_WSPAN);
                    const uint32_t q0 = (uint32_t)(ws - 1 - Srun) / 255u;
                    nl_i = (uint32_t)(end_i - 1 - Srun) / 255u - q0;
                    first_i = (uint32_t)Srun + 255u * (q0 + 1u);
                    nle_i = nl_i;
                    if (ragged) {
                        const int end_e = min(end_i, (int)rel_emit);
                        nle_i = end_e > ws ? (uint32_t)(end_e - 1 - Srun) / 255u - q0 : 0u;
                    }
                }
                if (i == warp) {
                    mybase = total_all;
                    mynl = nl_i;
                    myfirst = first_i;
                }
                total_all += nl_i + (Ca[i] & 0xffffu);
                total_emit += nle_i + (Ca[i] >> 16);
                if (La[i] >= 0) Srun = La[i];
            }
            F_t = Fa[0], L_t = La[0];
#pragma unroll
            for (int i = 1; i < RL_WARPS; ++i) F_t = min(F_t, Fa[i]), L_t = max(L_t, La[i]);
            L_lastwarp = La[RL_WARPS - 1];
        }
        try_flat = total_all < 192u;  // the same in every thread of the CTA
        RlePending& me = sh.q[k % RLE_Q];
        if (tid == 0) {
            // the tile's last 4 KiB held no run start, so d1[tile] is still empty: last run start of the tile, or "none"
            // (the service warp replaces "none" by the entering run's start once it knows it)
            if (L_lastwarp < 0) {
                if (L_t >= 0) st_relaxed_u64(d1 + tile, LB_INC | (uint64_t)(tbase + L_t + RL_BIAS));
                else if (tile > 0) atomicMax(reinterpret_cast<unsigned long long*>(d1 + tile), (unsigned long long)LB_AGG);  // never undoes the final value
            }
            sh.meta[kb].first_start = F_t;
            sh.meta[kb].last_start = L_t;
            sh.meta[kb].total_all = total_all;
            sh.meta[kb].total_emit = total_emit;
            mbar_arrive(&meta_bar[kb]);  // release: the service warp reads the four words above
            me.total_all = total_all;
            me.first_byte = buf[0];
            me.tile = tile;
        }

        // ---- compaction: value and position of every boundary into the ring
        const uint32_t wbyte = buf[warp * RL_WSPAN];  // value of the run that enters the warp's rows
        if (pass == 0 && total_all <= (uint32_t)RL_RING) {
            // room in the ring?  (the oldest waiting tile's boundaries start at its `head`)
            if (n_stored < k && ring_head + total_all - sh.q[n_stored % RLE_Q].head > (uint32_t)RL_RING) {
                do {
                    if (n_stored >= upto) wait_slots(n_stored + 1u);
                    store_tile(n_stored++);
                } while (n_stored < k && ring_head + total_all - sh.q[n_stored % RLE_Q].head > (uint32_t)RL_RING);
                if (upto < n_stored) upto = n_stored;
                bar_sync(BAR_COMPUTE, RL_THREADS);  // every warp has read the ring space that is written next
            }
            const uint32_t wb = ring_head + mybase;
            for (uint32_t j = lane; j < mynl; j += 32) {  // splits in the warp's leading rows
                const uint32_t idx = (wb + j) & (RL_RING - 1);
                const uint8_t p8 = (uint8_t)(myfirst + 255u * j);
                val8[idx] = (uint8_t)wbyte;
                pos8[idx] = p8;
                if (idx == 0) pos8[RL_RING] = p8;
            }
            const uint32_t kk0 = wb + mynl + excl;
            uint32_t kk = kk0;
#pragma unroll
            for (int g = 0; g < RL_NW; ++g) {
                uint32_t bits = M[g];
                while (bits) {
                    const uint32_t p = 32u * g + (uint32_t)__ffs(bits) - 1u;
                    bits &= bits - 1;
                    const uint32_t idx = kk & (RL_RING - 1);
                    val8[idx] = rowp[((((chunk0 + (p >> 4)) ^ sw) & 7u) << 4) | (p & 15u)];
                    pos8[idx] = (uint8_t)(rowstart + p);
                    ++kk;
                }
            }
            // entry RL_RING mirrors entry 0 (a count is the difference of neighbouring positions)
            if (kk != kk0 && (((kk - 1u) ^ (kk0 - 1u)) & ~(uint32_t)(RL_RING - 1))) pos8[RL_RING] = pos8[0];
            if (tid == 0) {
                me.head = ring_head;
                me.total = total_emit;
                me.skip = 0;
            }
            ring_head += total_all;
            __syncwarp();
            if (lane == 0) mbar_arrive(&freeb[buf_i]);  // the tile buffer may be refilled
            while (n_stored < upto) store_tile_light(n_stored++, k);
        } else if (pass == 0) {
            // more boundaries than the ring holds: counted (the service warp posts it like any tile), left to the second pass
            if (tid == 0) {
                me.head = ring_head;
                me.total = 0;
                me.skip = 1;
                redo[atomicAdd(redo_n, 1u)] = (uint32_t)tile;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&freeb[buf_i]);
            while (n_stored < upto) store_tile_light(n_stored++, k);
        } else {
            // second pass: the tile's first slot is final; its pairs are written at once, in rounds of one ring
            if (tid == 0) {
                me.head = 0;
                me.total = 0;
                me.skip = 0;
            }
            bar_sync(BAR_COMPUTE, RL_THREADS);  // me.* is written
            wait_slots(k + 1u);
            while (n_stored < k) store_tile(n_stored++);
            store_tile(k);  // its pairs at 255-splits before the first run start
            const uint64_t g0 = me.g0 + me.nl;
            const uint32_t rel0 = mybase + mynl + excl;  // tile-relative index of the row's first boundary
            for (uint32_t w0 = 0; w0 < total_emit; w0 += RL_RING) {
                bar_sync(BAR_COMPUTE, RL_THREADS);  // the ring is free (previous round / earlier tiles have been read)
                // boundaries w0 .. w0 + RL_RING (the last one only as the end of the count before it)
                for (uint32_t j = lane; j < mynl; j += 32) {
                    const uint32_t q = mybase + j - w0;
                    if (q <= (uint32_t)RL_RING) {
                        val8[q] = (uint8_t)wbyte;
                        pos8[q] = (uint8_t)(myfirst + 255u * j);
                    }
                }
                uint32_t q = rel0 - w0;
                // rows whose boundaries all lie outside this round's window have nothing to stage
                const bool mine = rel0 <= w0 + (uint32_t)RL_RING && rel0 + (packed & 0xffffu) > w0;
#pragma unroll
                for (int g = 0; g < RL_NW; ++g) {
                    uint32_t bits = mine ? M[g] : 0u;
                    while (bits) {
                        const uint32_t p = 32u * g + (uint32_t)__ffs(bits) - 1u;
                        bits &= bits - 1;
                        if (q <= (uint32_t)RL_RING) {
                            val8[q] = rowp[((((chunk0 + (p >> 4)) ^ sw) & 7u) << 4) | (p & 15u)];
                            pos8[q] = (uint8_t)(rowstart + p);
                        }
                        ++q;
                    }
                }
                bar_sync(BAR_COMPUTE, RL_THREADS);
                const uint32_t cw = min((uint32_t)RL_RING, total_emit - w0);
                const bool has_next = w0 + cw < total_all;  // else that count is written by the next tile
                store_compacted<false>(values, g0 + w0, cw, pair_cap, val8, 0, tid);
                store_compacted<true>(counts, g0 + w0, cw - (has_next ? 0u : 1u), pair_cap, pos8, 0, tid);
            }
            n_stored = k + 1u;
            bar_sync(BAR_COMPUTE, RL_THREADS);  // the ring has been read
            ring_head = 0;
            if (lane == 0) mbar_arrive(&freeb[buf_i]);
        }
    }
    // ---- the tiles that still wait
    {
        // the service warp may still be posting the last of them (it starts the next load, here the end marker, first)
        const uint32_t k_end = k;
        if (n_stored < k_end) {
            bar_sync(BAR_COMPUTE, RL_THREADS);
            wait_slots(k_end);
            while (n_stored < k_end) store_tile(n_stored++);
        }
    }
}

// =============================================================================================
//                                        launcher
// =============================================================================================
// positions 0 .. n_scan are scanned, n_scan = min(n_avail, n + 255) being a forced boundary
size_t rl_encode_tiles(size_t n) { return (((n + 255) / RL_TILE + 1) + 255) / 256 * 256; }  // padded: the scanner works on whole chunks
size_t rl_encode_workspace_bytes(size_t n) { return (3 * rl_encode_tiles(n) + 2) * sizeof(uint64_t); }

namespace {
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_fn()
{
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}
}  // namespace

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
    uint32_t* role = next_tile + 1;                     // role counter: the first CTA that runs is the scanner
    uint32_t* redo = next_tile + 4;                     // [+2] tiles left to the second pass, [+3] its counter, then the list
    cudaError_t e = cudaMemsetAsync(d_ws, 0, rl_encode_workspace_bytes(n), stream);
    if (e != cudaSuccess) return e;
    // the input as rows of 128 bytes; the partial last row is loaded with plain loads
    CUtensorMap tmap;
    memset(&tmap, 0, sizeof tmap);
    const size_t rows_full = n_avail / RL_ROW;
    if (rows_full) {
        EncodeTiledFn enc = encode_tiled_fn();
        if (!enc) return cudaErrorNotSupported;
        if (rows_full > 0x7fffffffull) return cudaErrorInvalidValue;  // 256 GiB per call
        const cuuint64_t gdim[2] = {RL_ROW, rows_full};
        const cuuint64_t gstride[1] = {RL_ROW};
        const cuuint32_t box[2] = {RL_ROW, 32};
        const cuuint32_t estr[2] = {1, 1};
        if (enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<uint8_t*>(d_in), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
            return cudaErrorInvalidValue;
    }
    int dev = 0;
    e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    static int resident[64];
    int cap = dev < 64 ? resident[dev] : 0;
    if (!cap) {
        e = cudaFuncSetAttribute((const void*)rl_encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, RLE_SMEM);
        if (e != cudaSuccess) return e;
        int sms = 0, per_sm = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, (const void*)rl_encode_kernel, RLE_THREADS, RLE_SMEM);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) return cudaErrorLaunchOutOfResources;
        cap = sms * std::max(1, std::min(per_sm, RLE_CTAS_PER_SM));
        if (cap < 2) return cudaErrorLaunchOutOfResources;
        if (dev < 64) resident[dev] = cap;
    }
    // one CTA becomes the scanner, the others take tiles; CTAs that are not resident at once simply start later
    const unsigned grid = (unsigned)std::min<size_t>(tiles, (size_t)cap - 1) + 1;
    rl_encode_kernel<<<grid, RLE_THREADS, RLE_SMEM, stream>>>(tmap, d_in, n, n_avail, n_scan, (uint32_t)(prev_run % 255u), prev_run != 0, d_counts,
                                                              d_values, pair_cap, d_pairs, d1, inc, agg, next_tile, role, redo, (long long)tiles, 0);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    // the second pass (it finds its list empty unless tiles had more pairs than the ring holds): launched early
    // (programmatic dependent launch) so that only its griddepcontrol.wait, not its launch, follows the first pass
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(RLE_THREADS);
    cfg.dynamicSmemBytes = RLE_SMEM;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, rl_encode_kernel, tmap, d_in, n, n_avail, n_scan, (uint32_t)(prev_run % 255u), (int)(prev_run != 0), d_counts, d_values,
                           pair_cap, d_pairs, d1, inc, agg, next_tile, role, redo, (long long)tiles, 1);
    if (e != cudaSuccess) return e;
    return cudaGetLastError();
}

}  // namespace flrl
