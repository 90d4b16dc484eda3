// rl_encode.cu — run-length encoder for sm_100a: persistent CTAs, TMA input ring, deferred output.
//
// The reference documents RL (IMPLEMENTATION-PLAN.md:83-152: mask -> scan -> start indices -> split of
// runs > 255 -> emit) but ships no code.  One launch, input read once:
//
//   * grid = RLE_CTAS_PER_SM persistent CTAs per SM; CTA c takes tiles c, c + G, c + 2G, ... (8 KiB each,
//     +16 B left halo, +256 B look-ahead), so the tiles in flight at any moment are neighbours in memory and
//     a look-back never reaches further than one "row" of G tiles;
//   * a control warp keeps RLE_STAGES TMA bulk loads in flight (full / free mbarriers) and runs the
//     pair-slot look-back (#2) of tile k WHILE the 8 compute warps already work on tile k + 1;
//   * compute warps, per tile: run-start bit mask of each 32-byte segment (XOR with the byte-shifted word,
//     SWAR non-zero-byte test, one multiply gathers 8 flags) -> block max-scan for the start S of the run
//     entering each segment (+ look-back #1 across tiles, depth 1 unless a whole tile is one run) -> the one
//     possible 255-split of a segment, (i - S) % 255 == 0 (the plan's ":125" rule, replacing its iterative
//     rescan loop) -> block sum-scan of pair counts + suffix-min of "next boundary" -> compaction of
//     (count, value) bytes into one of two shared staging buffers;
//   * the staged pairs of tile k are written to global memory one iteration later, when the control warp
//     has their first slot: 16-byte stores aligned in the GLOBAL address space (unaligned shared reads are
//     fixed up with PRMT), byte stores on the two ragged ends.
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
static constexpr int RLE_THREADS = RL_THREADS + 32;     // + control warp
static constexpr int RL_SEG = 32;                       // input bytes per thread
static constexpr int RL_TILE = RL_THREADS * RL_SEG;     // 8 KiB
static constexpr int RL_HALO = 256;                     // look-ahead: a count never exceeds 255
static constexpr int RL_INBUF = 16 + RL_TILE + RL_HALO + 16;
static constexpr int RL_STAGE = RL_TILE + 32;           // compacted counts / values (+ over-read slack)
static constexpr int RLE_STAGES = 2;                    // input ring
static constexpr int RLE_DEFER = 2;                     // a tile's pairs are stored this many iterations later
static constexpr int RLE_NB = RLE_DEFER + 1;            // staging buffers / per-tile scratch slots
static constexpr int RLE_CTAS_PER_SM = 3;
static constexpr uint64_t RL_BIAS = 255;                // run starts are stored as S + 255 >= 0
static constexpr uint32_t BAR_COMPUTE = 1;              // named barrier of the 256 compute threads

// ---- run-start flags of 8 consecutive words -------------------------------------------------
// bit (4k + j) of the result: byte j of word k differs from the byte before it.
// nz: high bit of each byte set iff that byte is non-zero; the multiply gathers the eight byte flags of a
// word pair into adjacent bits (no two partial products share a bit, so there are no carries).
__device__ __forceinline__ uint32_t nz_bytes_e(uint32_t x)
{
    return (((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x) & 0x80808080u;
}
__device__ __forceinline__ uint32_t start_mask32(const uint32_t (&w)[8], uint32_t prev_word)
{
    uint32_t m = 0, prev = prev_word;
#pragma unroll
    for (int k = 0; k < 8; k += 2) {
        const uint32_t a = nz_bytes_e(w[k] ^ __byte_perm(prev, w[k], 0x6543));
        const uint32_t b = nz_bytes_e(w[k + 1] ^ __byte_perm(w[k], w[k + 1], 0x6543));
        const uint32_t g = (a >> 7) | (b >> 3);  // bit 0 / bit 4 of every byte
        m |= ((g * 0x00204081u) >> 21 & 0xffu) << (4 * k);
        prev = w[k + 1];
    }
    return m;
}
__device__ __forceinline__ uint32_t segment_mask(const uint8_t* inbuf, int off)
{
    const uint4 a = *reinterpret_cast<const uint4*>(inbuf + 16 + off);
    const uint4 b = *reinterpret_cast<const uint4*>(inbuf + 32 + off);
    const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    return start_mask32(w, *reinterpret_cast<const uint32_t*>(inbuf + 12 + off));
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

struct RlEncShared {    // per-tile scratch, double-buffered by tile parity
    int wmax[8];        // per-warp last run start (tile relative), -1 none
    uint32_t wsum[8];   // per-warp pair count
    int wmin[8];        // per-warp first boundary position (tile relative)
    int sentinel;       // first run start in the look-ahead halo (tile relative), or RL_TILE + RL_HALO
    uint32_t d0;        // (tile base - start of the run entering the tile) % 255
    uint32_t total;     // pairs that start in the tile (compute -> control warp)
    uint64_t gbase;     // first pair slot of the tile (control warp -> compute)
};

__global__ void __launch_bounds__(RLE_THREADS, RLE_CTAS_PER_SM)
rl_encode_kernel(const uint8_t* __restrict__ in, size_t n, size_t n_avail, uint32_t prev_phase, int has_prev,
                 uint8_t* __restrict__ counts, uint8_t* __restrict__ values, size_t pair_cap, uint64_t* __restrict__ d_pairs,
                 uint64_t* __restrict__ d1, uint64_t* __restrict__ d2, long long tiles, int dbg)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* const inbufs = smem;                                // RLE_STAGES x [16 left halo][tile][halo][pad]
    uint8_t* const stages = smem + RLE_STAGES * RL_INBUF;        // 2 x (counts, values) staging
    __shared__ __align__(8) uint64_t full[RLE_STAGES], freeb[RLE_STAGES], agg_ready[RLE_NB], base_ready[RLE_NB];
    __shared__ RlEncShared sh[RLE_NB];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long G = gridDim.x, c = blockIdx.x;
    const long long K = (tiles - c + G - 1) / G;  // tiles of this CTA: c, c + G, ...

    if (tid == 0) {
        for (int s = 0; s < RLE_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&freeb[s], RL_THREADS / 32);
        }
        for (int b = 0; b < RLE_NB; ++b) {
            mbar_init(&agg_ready[b], 1);
            mbar_init(&base_ready[b], 1);
        }
        fence_mbar_init();
    }
    __syncthreads();

    if (warp == RL_THREADS / 32) {
        // =============================== control warp ======================================
        auto issue_load = [&](long long k) {
            const long long tile = k * G + c;
            const size_t tbase = (size_t)tile * RL_TILE;
            const int s = (int)(k % RLE_STAGES);
            uint8_t* buf = inbufs + s * RL_INBUF;
            const size_t lo = tile > 0 ? tbase - 16 : tbase;
            const size_t want = tbase + RL_TILE + RL_HALO < n_avail ? tbase + RL_TILE + RL_HALO : n_avail;
            const size_t hi16 = want & ~(size_t)15;
            // the ragged end (< 16 bytes) and the byte "before" the stream: plain loads
            if (lane < 15) {
                const size_t g = (hi16 > lo ? hi16 : lo) + lane;
                if (g < want) buf[16 + (g - tbase)] = in[g];
            }
            if (tile == 0 && lane == 15) buf[15] = has_prev ? in[0] : (uint8_t)~in[0];
            __syncwarp();
            if (lane == 0) {
                if (hi16 > lo) {
                    mbar_expect_tx(&full[s], (uint32_t)(hi16 - lo));
                    tma_load_1d(buf + (tile > 0 ? 0 : 16), in + lo, (uint32_t)(hi16 - lo), &full[s]);
                } else {
                    mbar_arrive(&full[s]);
                }
            }
        };
        for (long long k = 0; k < K && k < RLE_STAGES - 1; ++k) issue_load(k);
        for (long long k = 0; k < K; ++k) {
            const long long kn = k + RLE_STAGES - 1;
            if (kn < K) {
                if (kn >= RLE_STAGES) mbar_wait(&freeb[kn % RLE_STAGES], (uint32_t)(kn / RLE_STAGES - 1) & 1u);
                issue_load(kn);
            }
            // look-back #2 of tile k, while the compute warps are already on tile k + 1
            const int b = (int)(k % RLE_NB);
            const long long tile = k * G + c;
            mbar_wait(&agg_ready[b], (uint32_t)(k / RLE_NB) & 1u);
            const uint32_t total = sh[b].total;
            uint64_t g = 0;
            if (tile > 0 && !(dbg & 2)) {
                if (lane == 0) st_relaxed_u64(d2 + tile, LB_AGG | total);
                g = lookback_exclusive_sum(d2, tile, lane, (unsigned)dbg >> 4);
            } else if (dbg & 2) {
                g = (uint64_t)tile * 100;  // TIMING EXPERIMENT ONLY: wrong slots, no look-back #2
            }
            if (lane == 0) {
                st_relaxed_u64(d2 + tile, LB_INC | (g + total));
                sh[b].gbase = g;
                if (tile == tiles - 1) *d_pairs = g + total;
                mbar_arrive(&base_ready[b]);
            }
        }
        return;
    }

    // ================================= compute warps =======================================
    auto store_tile = [&](long long j) {
        const int pb = (int)(j % RLE_NB);
        mbar_wait(&base_ready[pb], (uint32_t)(j / RLE_NB) & 1u);
        const uint64_t g0 = sh[pb].gbase;
        const uint32_t cnt = sh[pb].total;
        const uint8_t* pc = stages + pb * 2 * RL_STAGE;
        store_compacted(counts, g0, cnt, pair_cap, pc, tid);
        store_compacted(values, g0, cnt, pair_cap, pc + RL_STAGE, tid);
    };
    for (long long k = 0; k < K; ++k) {
        const int b = (int)(k % RLE_NB), s = (int)(k % RLE_STAGES);
        const long long tile = k * G + c;
        const size_t tbase = (size_t)tile * RL_TILE;
        const uint8_t* inbuf = inbufs + s * RL_INBUF;
        uint8_t* cnt8 = stages + b * 2 * RL_STAGE;
        uint8_t* val8 = cnt8 + RL_STAGE;
        RlEncShared& sc = sh[b];
        mbar_wait(&full[s], (uint32_t)(k / RLE_STAGES) & 1u);

        // ---- run-start mask of this thread's 32 bytes
        const int off = tid * RL_SEG;
        uint32_t m = segment_mask(inbuf, off);
        const size_t gpos = tbase + off;
        // positions at or past the end of the available data close the last run; positions >= n are not emitted
        uint32_t emit = ~0u;
        const bool ragged = tbase + RL_TILE + RL_HALO > n;  // CTA-uniform: only the last tiles see the end of the data
        if (ragged) {
            const uint32_t k_av = n_avail > gpos ? (n_avail - gpos < 32 ? (uint32_t)(n_avail - gpos) : 32u) : 0u;
            const uint32_t k_em = n > gpos ? (n - gpos < 32 ? (uint32_t)(n - gpos) : 32u) : 0u;
            if (k_av < 32) m |= ~0u << k_av;
            emit = k_em < 32 ? ((1u << k_em) - 1u) : ~0u;
        }

        // warp 7 also finds the first run start inside the look-ahead halo
        if (warp == 7) {
            int first = RL_HALO;
            if (lane < RL_HALO / RL_SEG) {
                const int hoff = RL_TILE + lane * RL_SEG;
                uint32_t hm = segment_mask(inbuf, hoff);
                if (ragged) {
                    const size_t hg = tbase + hoff;
                    const uint32_t hk = n_avail > hg ? (n_avail - hg < 32 ? (uint32_t)(n_avail - hg) : 32u) : 0u;
                    if (hk < 32) hm |= ~0u << hk;
                }
                if (hm) first = lane * RL_SEG + (__ffs(hm) - 1);
            }
            first = __reduce_min_sync(0xffffffffu, first);
            if (lane == 0) sc.sentinel = RL_TILE + first;
        }

        // ---- S: last run start before this segment (block-wide exclusive max-scan)
        const int mine = m ? off + 31 - __clz(m) : -1;
        int inc = mine;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc = max(inc, o);
        }
        int s_rel = __shfl_up_sync(0xffffffffu, inc, 1);
        if (lane == 0) s_rel = -1;
        if (lane == 31) sc.wmax[warp] = inc;
        bar_sync(BAR_COMPUTE, RL_THREADS);
        {
            int pm = -1;
#pragma unroll
            for (int q = 0; q < 7; ++q)
                if (q < warp) pm = max(pm, sc.wmax[q]);
            s_rel = max(s_rel, pm);
        }
        if (warp == 0) {
            // look-back #1: only tiles whose first byte continues an earlier run need it
            int tl = -1;
#pragma unroll
            for (int q = 0; q < 8; ++q) tl = max(tl, sc.wmax[q]);
            const bool need = (__shfl_sync(0xffffffffu, m, 0) & 1u) == 0;  // thread 0's first byte is not a run start
            if (tl >= 0 && lane == 0) st_relaxed_u64(d1 + tile, LB_INC | (uint64_t)(tbase + tl + RL_BIAS));
            if (need) {
                uint64_t sb;  // biased start of the entering run
                if (tile == 0) {
                    sb = RL_BIAS - prev_phase;
                } else if (dbg & 1) {
                    sb = RL_BIAS;  // TIMING EXPERIMENT ONLY: wrong phase, no look-back #1
                } else {
                    if (tl < 0 && lane == 0) st_relaxed_u64(d1 + tile, LB_AGG);
                    sb = lookback_run_start(d1, tile, lane);
                }
                if (lane == 0) {
                    if (tl < 0) st_relaxed_u64(d1 + tile, LB_INC | sb);
                    sc.d0 = (uint32_t)((tbase + RL_BIAS - sb) % 255u);
                }
            }
        }
        bar_sync(BAR_COMPUTE, RL_THREADS);

        // ---- the one possible 255-split inside the leading part of the segment
        const uint32_t lead = m ? (uint32_t)(__ffs(m) - 1) : 32u;  // bytes that continue the entering run
        uint32_t bm = m;                                           // pair boundaries: run starts + the split
        if (lead) {
            const uint32_t phase = s_rel >= 0 ? (uint32_t)(off - s_rel) % 255u : (sc.d0 + (uint32_t)off) % 255u;
            const uint32_t j = phase ? 255u - phase : 0u;
            if (j < lead) bm |= 1u << j;
        }
        const uint32_t e = bm & emit;
        const uint32_t cn = __popc(e);

        // ---- pair slots (exclusive sum-scan) and the next boundary after this segment (suffix min)
        uint32_t isum = cn;
        int nxt = bm ? off + (__ffs(bm) - 1) : 0x7fffffff;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, isum, d);
            const int q = __shfl_down_sync(0xffffffffu, nxt, d);
            if (lane >= d) isum += o;
            if (lane + d < 32) nxt = min(nxt, q);
        }
        if (lane == 31) sc.wsum[warp] = isum;
        if (lane == 0) sc.wmin[warp] = nxt;
        int nxt_after = __shfl_down_sync(0xffffffffu, nxt, 1);
        if (lane == 31) nxt_after = 0x7fffffff;
        bar_sync(BAR_COMPUTE, RL_THREADS);
        uint32_t slot = isum - cn, total = 0;
        {
            nxt_after = min(nxt_after, sc.sentinel);
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const uint32_t ws = sc.wsum[q];
                if (q < warp) slot += ws;
                if (q > warp) nxt_after = min(nxt_after, sc.wmin[q]);
                total += ws;
            }
        }
        if (tid == 0) {  // hand the tile's pair count to the control warp (look-back #2 runs behind our back)
            sc.total = total;
            mbar_arrive(&agg_ready[b]);
        }

        // ---- compaction: (count, value) of every pair that starts in this segment
        {
            uint32_t bits = e, kk = slot;
            int prevb = -1;
            while (bits) {
                const int bi = __ffs(bits) - 1;
                bits &= bits - 1;
                val8[kk] = inbuf[16 + off + bi];
                if (prevb >= 0) cnt8[kk - 1] = (uint8_t)(bi - prevb);
                prevb = bi;
                ++kk;
            }
            if (prevb >= 0) {
                // the segment's last pair ends at the next boundary (this segment's non-emitted ones included)
                const uint32_t later = bm & ~emit & ~((2u << prevb) - 1u);
                const int nb = later ? off + (__ffs(later) - 1) : nxt_after;
                const int len = nb - (off + prevb);
                cnt8[kk - 1] = (uint8_t)(len < 255 ? len : 255);
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&freeb[s]);  // this warp is done with the input stage

        // ---- deferred output of tile k - RLE_DEFER: its first slot has been resolved meanwhile
        if (k >= RLE_DEFER) store_tile(k - RLE_DEFER);
    }
    // ---- output of the CTA's last tiles
    bar_sync(BAR_COMPUTE, RL_THREADS);
    for (long long j = K > RLE_DEFER ? K - RLE_DEFER : 0; j < K; ++j) store_tile(j);
}

// =============================================================================================
//                                        launcher
// =============================================================================================
size_t rl_encode_tiles(size_t n) { return (n + RL_TILE - 1) / RL_TILE; }
size_t rl_encode_workspace_bytes(size_t n) { return (2 * rl_encode_tiles(n) + 2) * sizeof(uint64_t); }

cudaError_t rl_encode_launch(const uint8_t* d_in, size_t n, size_t n_avail, uint64_t prev_run, uint8_t* d_counts,
                             uint8_t* d_values, size_t pair_cap, uint64_t* d_pairs, void* d_ws, cudaStream_t stream)
{
    const size_t tiles = rl_encode_tiles(n);
    uint64_t* d1 = reinterpret_cast<uint64_t*>(d_ws);
    uint64_t* d2 = d1 + tiles;
    cudaError_t e = cudaMemsetAsync(d1, 0, 2 * tiles * sizeof(uint64_t), stream);
    if (e != cudaSuccess) return e;
    constexpr int smem = RLE_STAGES * RL_INBUF + 2 * RLE_NB * RL_STAGE;
    // the grid must be fully co-resident: CTAs wait on tiles that other CTAs of the same launch own
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
        if (dev < 64) resident[dev] = cap;
    }
    const unsigned grid = (unsigned)std::min<size_t>(tiles, (size_t)cap);
    rl_encode_kernel<<<grid, RLE_THREADS, smem, stream>>>(d_in, n, n_avail, (uint32_t)(prev_run % 255u), prev_run != 0, d_counts,
                                                          d_values, pair_cap, d_pairs, d1, d2, (long long)tiles,
                                                          getenv("FLRL_RLE_DBG") ? atoi(getenv("FLRL_RLE_DBG")) : 0);
    return cudaGetLastError();
}

}  // namespace flrl
