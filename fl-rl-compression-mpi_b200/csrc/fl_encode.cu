// fl_encode.cu — fixed-length encoder for sm_100a: one launch, input read from DRAM once, output written once.
//
// Replaces the reference's compress kernels + thrust scan + byte-atomic CAS loops
// (/root/reference/src/fl/fl_gpu.cu:648-726, :758-808).
//
// The value offset of a tile needs the widths of every earlier tile.  Two touches per 128-frame tile (16 KiB),
// FLE_LEAD iterations apart, by PERSISTENT CTAs (4 pack warps + 1 service warp; tiles from an atomic counter):
//   * width touch of tile t (first touch: DRAM -> L2 -> shared memory, L2 evict-last): per-frame width from the OR
//     of the frame (fl_cpu.cu:37-50) — one thread per frame, its eight 16-byte pieces read in a rotated order so
//     that a quarter-warp covers all banks, no shuffles.  Warp 0 writes the 128 widths to `bits`, scans them once,
//     keeps (offset, width) of every frame in a small shared-memory ring for the pack touch, and posts the tile's
//     packed size (16-byte units: a full frame of width b packs to exactly 16 b bytes) to the grid's scanner warp
//     (rl_scan.cuh), which turns sizes into offsets in tile order;
//   * pack touch of tile t - FLE_LEAD * G, in the same loop iteration (second touch: an L2 hit, evict-first):
//     (offset, width) from the ring, the tile's offset in `values` (posted thousands of tiles ago: one load, asked
//     for at the top of the iteration), lane-local multiply/shift packing (flrl_bits.cuh; fl_cpu.cu:64-84; four
//     lanes per frame) IN PLACE over the staged tile once every pack warp holds its part in registers, one TMA bulk
//     store (every tile starts on a 16-byte boundary of `values`).
// FLE_LEAD * G tiles (~19 MiB) exceed what is in flight and stay inside the 126 MB L2 (a longer lead loses more of
// the second touches to L2 misses).  The pack warps never wait for each other: everything that joins them — loads,
// the scan of a tile's widths, the bulk store — goes through the service warp (bar.arrive from the pack warps,
// bar.sync in the service warp); the pack warps wait on mbarriers for loaded slots and meet once per tile, before the
// tile is overwritten in place.  Two 16 KiB slots per CTA leave room for 6 CTAs per SM, which is what hides the
// latencies (a separate output slot, 4 CTAs per SM and no barrier at all: 0.322 ms).
// The kernel is bound by instruction issue and latency in the pack warps (about 1000 instructions per warp and tile,
// ~450 of them the packing), not by memory: without the packing arithmetic it runs at 86 % of the HBM bandwidth.
// (History: one CTA per tile with a look-back, 0.417 ms / GiB; persistent two-touch with widths by shuffles and
// recomputed in the pack touch, 0.42 ms; widths by one thread per frame + the (offset, width) ring, 0.36 ms; joins
// moved to the service warp and a lead of 2 iterations, 0.33 ms; tiles from an atomic counter, 0.32 ms; packing in
// place and 6 CTAs per SM, 0.307 ms.)
//
// Tiles are handed out by an atomic counter (as in the run-length kernels): only running CTAs own tiles, so nothing
// depends on the whole grid being resident; the launcher still sizes the grid to what fits, plus the scanner CTA.
//
// Only the final partial frame (n % 128 != 0) goes through a small cooperative scalar path in the CTA that
// owns the last tile, which also publishes the total value size.
#include <stdlib.h>

#include <algorithm>

#include "fl_kernels.cuh"
#include "flrl_bits.cuh"
#include "rl_scan.cuh"

namespace flrl {

static constexpr int FLE_TILE_FRAMES = 128;
static constexpr int FLE_TILE_BYTES = FLE_TILE_FRAMES * 128;  // 16 KiB
#ifndef FLE_PACK
#define FLE_PACK 128  // pack threads; four of them share a frame
#endif
static constexpr int FLE_THREADS = FLE_PACK + 32;             // + service warp
static constexpr int FLE_ITEMS = 4 * FLE_TILE_FRAMES / FLE_PACK;  // 32-byte items per thread per tile
// 16 KiB slots per CTA: width touch, pack touch (packed in place), and with !FLE_INPLACE a separate output slot.
// One of each measured best (more slots and fewer CTAs / SM: slower every time; 256 pack threads: slower).
#ifndef FLE_WSLOTS
#define FLE_WSLOTS 1
#endif
#ifndef FLE_PSLOTS
#define FLE_PSLOTS 1
#endif
#ifndef FLE_OSLOTS
#define FLE_OSLOTS 1
#endif
#ifndef FLE_INPLACE
#define FLE_INPLACE 1  // 1: pack in place over the staged tile (no output slot; one barrier of the pack warps per tile)
#endif
#ifndef FLE_CTAS
#define FLE_CTAS 6  // per SM
#endif
#ifndef FLE_LEAD
#define FLE_LEAD 2  // iterations between the two touches of a tile (1: 0.370 ms, 2: 0.328, 3: 0.337, 5: 0.356)
#endif
#ifndef FLE_EXP
#define FLE_EXP 0  // timing experiments only (wrong output): 1 = no packing arithmetic, 2 = no width arithmetic
#endif
static constexpr int FLE_IDS = FLE_LEAD + 3;
static constexpr uint32_t BAR_P = 1;  // named barriers: the pack warps among themselves (tail frame only),
static constexpr uint32_t BAR_W = 2;  //   widths of a tile ready: pack warps arrive, the service warp waits,
static constexpr uint32_t BAR_O = 3;  //   a tile is packed: pack warps arrive, the service warp waits

struct FleShared {
    uint8_t bw[2][FLE_TILE_FRAMES];  // width touch: widths (two copies: warp 0 reads one while the others fill the next)
    // kept from the width touch to the pack touch FLE_LEAD iterations later: per frame (offset in 16-byte units) << 8 | width
    __align__(16) uint32_t meta[FLE_LEAD + 1][FLE_TILE_FRAMES];
    uint32_t total[FLE_LEAD + 1];    // packed size of the tile (16-byte units)
    int tile[FLE_LEAD + 3];          // tile of iteration k at [k % FLE_IDS], from its first load to its store
};

__device__ __forceinline__ uint32_t fle_byte_sum(uint32_t v) { return __dp4a(v, 0x01010101u, 0u); }

// final partial frame (len = n % 128 in 1..127), the 128 pack threads participate
// fl_cpu.cu:37-50 (width), :64-84 (packing), :53 (size = ceil(bits/8)).
__device__ __noinline__ uint32_t fle_tail_frame(const uint8_t* in, uint32_t len, uint8_t* bits_out, uint8_t* vout,
                                                uint32_t* scratch /* >= 35 words smem */)
{
    const int tid = threadIdx.x;
    const uint32_t v = (tid < (int)len) ? in[tid] : 0u;
    if (tid < 35) scratch[tid] = 0;
    bar_sync(BAR_P, FLE_PACK);
    const uint32_t wo = __reduce_or_sync(0xffffffffu, v);
    if ((tid & 31) == 0 && wo) atomicOr(&scratch[34], wo);
    bar_sync(BAR_P, FLE_PACK);
    const uint32_t b = width_of_or(scratch[34]);
    if (tid < (int)len) {
        const uint32_t bit = tid * b;
        const uint32_t w = bit >> 5, sh = bit & 31;
        atomicOr(&scratch[w], v << sh);
        if (sh + b > 32) atomicOr(&scratch[w + 1], v >> (32 - sh));
    }
    bar_sync(BAR_P, FLE_PACK);
    const uint32_t nbytes = (len * b + 7) >> 3;
    if (tid < (int)nbytes) vout[tid] = (uint8_t)(scratch[tid >> 2] >> (8 * (tid & 3)));
    if (tid == 0) *bits_out = (uint8_t)b;
    return nbytes;
}

__device__ __forceinline__ int fle_full_frames_of(size_t full_frames, long long tile)
{
    const size_t f0 = (size_t)tile * FLE_TILE_FRAMES;
    return (int)(full_frames > f0 ? (full_frames - f0 < (size_t)FLE_TILE_FRAMES ? full_frames - f0 : FLE_TILE_FRAMES) : 0);
}

__global__ void __launch_bounds__(FLE_THREADS, FLE_CTAS)
fl_encode_kernel(const uint8_t* __restrict__ in, size_t n, uint8_t* __restrict__ bits, uint8_t* __restrict__ values,
                 uint64_t* __restrict__ values_size, uint64_t* __restrict__ inc, uint32_t* __restrict__ agg,
                 unsigned int* __restrict__ next_tile, long long tiles)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* const wbuf = smem;                                 // FLE_WSLOTS x 16 KiB: width touch
    uint8_t* const pbuf = smem + FLE_WSLOTS * FLE_TILE_BYTES;   // FLE_PSLOTS x 16 KiB: pack touch
    [[maybe_unused]] uint8_t* const obuf = pbuf + FLE_PSLOTS * FLE_TILE_BYTES;  // !FLE_INPLACE: FLE_OSLOTS x 16 KiB, the packed tile
    __shared__ __align__(8) uint64_t w_full[FLE_WSLOTS], w_free[FLE_WSLOTS], p_full[FLE_PSLOTS], p_free[FLE_PSLOTS],
        o_free[FLE_OSLOTS > 0 ? FLE_OSLOTS : 1];
    __shared__ FleShared sh;
    __shared__ __align__(16) uint4 ktab[9];  // PackK{k1,k2,k3,ra} per width
    __shared__ uint32_t tail_scratch[36];

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
    // Tiles are handed out by an atomic counter, in the order the CTAs ask: a tile is only ever waited for by CTAs that
    // run, and its owner runs too (no co-residency needed), and the scanner sees sizes posted roughly in tile order.
    // sh.tile[k % FLE_IDS] is the tile of this CTA's iteration k (-1: there was none left).
    int Kw = 1 << 30;  // number of tiles this CTA got: known when the counter runs out
    const size_t full_frames = n / FRAME;
    if (tid == 0) {
        for (int s = 0; s < FLE_WSLOTS; ++s) {
            mbar_init(&w_full[s], 1);
            mbar_init(&w_free[s], FLE_PACK / 32);
        }
        for (int s = 0; s < FLE_PSLOTS; ++s) {
            mbar_init(&p_full[s], 1);
            mbar_init(&p_free[s], FLE_PACK / 32);
        }
        for (int s = 0; s < (FLE_OSLOTS > 0 ? FLE_OSLOTS : 1); ++s) mbar_init(&o_free[s], 1);
        fence_mbar_init();
    }
    if (tid <= 8) {
        const PackK k = make_pack_k(tid ? tid : 8);
        ktab[tid] = tid ? make_uint4(k.k1, k.k2, k.k3, k.ra) : make_uint4(0, 0, 0, 0);
    }
    __syncthreads();

    if (warp == FLE_PACK / 32) {
        // ================================ service warp ===========================================
        // Everything that joins the four pack warps goes through this warp, so that they never wait for each other:
        //   loads (both touches), the per-tile scan of the widths (bits, (offset, width) ring, size to the scanner),
        //   the bulk store of a packed tile.  Order per iteration k:  P load(k - LEAD) | widths(k) | W load(k + 1) | store(k - LEAD)
        const uint64_t keep = l2_policy_evict_last(), drop = l2_policy_evict_first();
        unsigned int idn = 0;  // lane 0: the tile of the next width load, asked for one iteration ahead
        if (lane == 0) idn = atomicAdd(next_tile, 1u);
        auto load_w = [&](int k) {  // refill a width slot with this CTA's k-th tile: first touch, stays in L2 for the second one
            const int sw = k % FLE_WSLOTS;
            if (k >= FLE_WSLOTS) mbar_wait(&w_free[sw], (uint32_t)(k / FLE_WSLOTS - 1) & 1u);
            const unsigned int id = __shfl_sync(0xffffffffu, idn, 0);
            const bool none = (long long)id >= tiles;
            if (lane == 0) {
                sh.tile[k % FLE_IDS] = none ? -1 : (int)id;  // (published by the arrive / expect_tx below)
                const uint32_t nfr = none ? 0u : (uint32_t)fle_full_frames_of(full_frames, id);
                if (nfr) {
                    mbar_expect_tx(&w_full[sw], nfr * FRAME);
                    tma_load_1d_hint(wbuf + sw * FLE_TILE_BYTES, in + (size_t)id * FLE_TILE_BYTES, nfr * FRAME, &w_full[sw], keep);
                } else {
                    mbar_arrive(&w_full[sw]);
                }
                if (!none) idn = atomicAdd(next_tile, 1u);
            }
            __syncwarp();
            if (none) Kw = k;
        };
        load_w(0);
        for (int k = 0; k < Kw + FLE_LEAD; ++k) {
            const bool doW = k < Kw;
            const int kp = k - FLE_LEAD;
            const bool doP = kp >= 0;
            const long long tileW = doW ? sh.tile[k % FLE_IDS] : -1, tileP = doP ? sh.tile[kp % FLE_IDS] : -1;
            uint64_t g = 1ull << 63;
            if (doP) {  // second touch: an L2 hit, not needed afterwards
                const int sp = kp % FLE_PSLOTS;
#if !FLE_INPLACE
                if (kp >= FLE_PSLOTS) mbar_wait(&p_free[sp], (uint32_t)(kp / FLE_PSLOTS - 1) & 1u);
#endif  // (in place: this warp waited for the slot's last bulk store itself, right after issuing it)
                if (lane == 0) {
                    const uint32_t nfr = (uint32_t)fle_full_frames_of(full_frames, tileP);
                    if (nfr) {
                        mbar_expect_tx(&p_full[sp], nfr * FRAME);
                        tma_load_1d_hint(pbuf + sp * FLE_TILE_BYTES, in + (size_t)tileP * FLE_TILE_BYTES, nfr * FRAME, &p_full[sp], drop);
                    } else {
                        mbar_arrive(&p_full[sp]);
                    }
                    // the tile's first unit in `values`: posted FLE_LEAD * G tiles ago, asked for now, used at the store
                    if (tileP > 0) g = ld_relaxed_u64(inc + tileP);
                }
                __syncwarp();
            }
            if (doW) {
                // widths of tile W (written by the pack warps) -> `bits`, the tile's size to the scanner, (offset, width)
                // of every frame kept for the pack touch
                const int mw = k % (FLE_LEAD + 1);
                bar_sync(BAR_W, FLE_THREADS);
                const int nfrW = fle_full_frames_of(full_frames, tileW);
                const uint32_t v = reinterpret_cast<const uint32_t*>(sh.bw[k & 1])[lane];
                const uint32_t b0 = v & 0xff, b1 = (v >> 8) & 0xff, b2 = (v >> 16) & 0xff, b3 = v >> 24;
                const uint32_t s = b0 + b1 + b2 + b3;
                uint32_t inc_s = s;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const uint32_t o = __shfl_up_sync(0xffffffffu, inc_s, d);
                    if (lane >= d) inc_s += o;
                }
                const uint32_t ex = inc_s - s;
                reinterpret_cast<uint4*>(sh.meta[mw])[lane] =
                    make_uint4(ex << 8 | b0, (ex + b0) << 8 | b1, (ex + b0 + b1) << 8 | b2, (ex + b0 + b1 + b2) << 8 | b3);
                const size_t f0 = (size_t)tileW * FLE_TILE_FRAMES;
                if (nfrW == FLE_TILE_FRAMES && (reinterpret_cast<uintptr_t>(bits) & 3) == 0) {
                    reinterpret_cast<uint32_t*>(bits + f0)[lane] = v;
                } else {
                    for (int f = lane; f < nfrW; f += 32) bits[f0 + f] = sh.bw[k & 1][f];
                }
                if (lane == 31) {
                    sh.total[mw] = inc_s;
                    st_relaxed_u32(agg + tileW, 0x80000000u | inc_s);  // to the scanner
                }
                __syncwarp();  // the ring row is complete before lane 0 hands out the pack touch of this tile
                load_w(k + 1);  // its slot was released before the pack warps arrived at the barrier above
            }
            if (doP) {
#if FLE_INPLACE
                const int so = kp % FLE_PSLOTS;
                uint8_t* const obuf = pbuf;
#else
                const int so = kp % FLE_OSLOTS;
#endif
                const int mp = kp % (FLE_LEAD + 1);
                bar_sync(BAR_O, FLE_THREADS);  // tile P is packed (and fenced for the bulk store)
                if (lane == 0) {
                    while ((g >> 63) == 0) {
                        __nanosleep(64);
                        g = ld_relaxed_u64(inc + tileP);
                    }
                    const uint64_t base = g & ~(1ull << 63);
                    const uint32_t total = sh.total[mp];
                    if (total) tma_store_1d_hint(values + base * 16, obuf + so * FLE_TILE_BYTES, total * 16, drop);
                    tma_store_commit();
                    tma_store_wait_read<0>();  // (in place: the slot may be loaded again — this warp does that itself)
#if !FLE_INPLACE
                    mbar_arrive(&o_free[so]);  // the output slot may be written again
#endif
                }
                __syncwarp();
            }
        }
        return;
    }

    // ================================== pack warps =========================================
    // They wait only for data: a loaded slot, a free output slot.  What they produce for others they hand over with
    // bar.arrive, never waiting for each other.
    const int t = tid & 3;
    for (int k = 0; k < Kw + FLE_LEAD; ++k) {
        const int kp = k - FLE_LEAD;
        const int sw = k % FLE_WSLOTS;
        long long tileW = -1;
        if (k < Kw) {
            mbar_wait(&w_full[sw], (uint32_t)(k / FLE_WSLOTS) & 1u);
            tileW = sh.tile[k % FLE_IDS];
            if (tileW < 0) Kw = k;  // the counter ran out: FLE_LEAD pack touches are left
        }

        // ---- width touch of tile W
        if (tileW >= 0) {
            const int nfrW = fle_full_frames_of(full_frames, tileW);
            const uint8_t* wt = wbuf + sw * FLE_TILE_BYTES;
            // one thread per frame, no shuffles: the eight 16-byte pieces are read in an order rotated by the frame
            // number, so the eight lanes of a quarter-warp hit eight different bank groups (frames are 128 bytes apart)
            for (int fr = tid; fr < FLE_TILE_FRAMES; fr += FLE_PACK) {
                const uint4* p = reinterpret_cast<const uint4*>(wt + fr * 128);
                uint32_t o = 0;
#if !(FLE_EXP & 2)
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const uint4 a = p[(j + fr) & 7];  // frames past nfr read stale shared memory and get width 0
                    o |= (a.x | a.y) | (a.z | a.w);
                }
                o |= o >> 16;
                o |= o >> 8;
#else
                o = 1u << (fr & 7);
#endif
                sh.bw[k & 1][fr] = (uint8_t)((fr < nfrW) ? width_of_or(o) : 0u);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&w_free[sw]);  // the slot may be refilled (its values were used above)
            bar_arrive(BAR_W, FLE_THREADS);           // the widths are in shared memory: over to the service warp
        }
        if (kp < 0) continue;

        // ---- pack touch of tile P
        const long long tileP = sh.tile[kp % FLE_IDS];
        const int sp = kp % FLE_PSLOTS, mp = kp % (FLE_LEAD + 1);
        const uint8_t* ptile = pbuf + sp * FLE_TILE_BYTES;
        mbar_wait(&p_full[sp], (uint32_t)(kp / FLE_PSLOTS) & 1u);  // (also: the tile's ring row is complete)
        uint4 A[FLE_ITEMS], C[FLE_ITEMS];
        uint32_t M[FLE_ITEMS];
#pragma unroll
        for (int i = 0; i < FLE_ITEMS; ++i) {
            const int fr = i * (FLE_PACK / 4) + (tid >> 2);
            const uint4* p = reinterpret_cast<const uint4*>(ptile + fr * 128 + t * 32);
            A[i] = p[0];
            C[i] = p[1];
            M[i] = sh.meta[mp][fr];  // (offset in 16-byte units) << 8 | width
        }
#if FLE_INPLACE
        bar_sync(BAR_P, FLE_PACK);  // every pack warp has the tile in registers: it is overwritten with the packed words
        {
            const uint32_t tw = smem_u32(ptile);
#else
        const int so = kp % FLE_OSLOTS;
        if (kp >= FLE_OSLOTS) mbar_wait(&o_free[so], (uint32_t)(kp / FLE_OSLOTS - 1) & 1u);  // its last bulk store has read it
        {
            const uint32_t tw = smem_u32(obuf + so * FLE_TILE_BYTES);
#endif
#pragma unroll
            for (int i = 0; i < FLE_ITEMS; ++i) {
                const uint32_t b = M[i] & 0xffu;  // 0 for frames past the end: every store below is predicated off
                const uint4 kk = ktab[b];
                PackK pk;
                pk.k1 = kk.x; pk.k2 = kk.y; pk.k3 = kk.z; pk.ra = kk.w; pk.s8 = 8 * b;
                const uint32_t w = tw + ((M[i] >> 8) * 4u + b * t) * 4u;
                const uint32_t a[4] = {A[i].x, A[i].y, A[i].z, A[i].w};
                const uint32_t c4[4] = {C[i].x, C[i].y, C[i].z, C[i].w};
#if (FLE_EXP & 1)
                if (b == 77) sts32_pred(w, a[0] ^ c4[0] ^ a[1] ^ c4[1] ^ a[2] ^ c4[2] ^ a[3] ^ c4[3] ^ pk.k1, true);
#else
                pack32_store(a, c4, b, pk, [&](bool p, uint32_t j, uint32_t v) { sts32_pred(w + 4u * j, v, p); });
#endif
            }
        }
        // every shared-memory read of the staged tile has returned (its values were used above): the slot may be refilled.
        // (Not earlier: an arrive right behind the loads can overtake them.)
#if !FLE_INPLACE
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_free[sp]);
#endif
        fence_proxy_async_smem();          // the packed tile is read by a bulk store
        bar_arrive(BAR_O, FLE_THREADS);    // over to the service warp, which stores it

        if (tileP == tiles - 1) {
            // ---- the partial frame and the total size: this CTA owns the last tile
            uint64_t g = 1ull << 63;
            if (tileP > 0)
                while (((g = ld_relaxed_u64(inc + tileP)) >> 63) == 0) __nanosleep(64);
            const uint32_t tail = (uint32_t)(n % FRAME);
            uint64_t vsize = ((g & ~(1ull << 63)) + sh.total[mp]) * 16;
            if (tail) vsize += fle_tail_frame(in + full_frames * FRAME, tail, bits + full_frames, values + vsize, tail_scratch);
            if (tid == 0) *values_size = vsize;
        }
    }
}

// =============================================================================================
//                                        launcher
// =============================================================================================
static size_t fle_tiles_padded(size_t n) { return (fl_tiles(n) + 255) / 256 * 256; }  // whole scanner chunks
// workspace: inc[tiles] u64 | agg[tiles] u32 | tile counter
size_t fl_encode_workspace_bytes(size_t n) { return fle_tiles_padded(n) * 12 + 16; }

cudaError_t fl_encode_launch(const uint8_t* d_in, size_t n, uint8_t* d_bits, uint8_t* d_values, uint64_t* d_values_size,
                             void* d_ws, cudaStream_t stream)
{
    const size_t tiles = fl_tiles(n), tmax = fle_tiles_padded(n);
    uint64_t* inc = reinterpret_cast<uint64_t*>(d_ws);
    uint32_t* agg = reinterpret_cast<uint32_t*>(inc + tmax);
    unsigned int* next_tile = agg + tmax;
    cudaError_t e = cudaMemsetAsync(d_ws, 0, fl_encode_workspace_bytes(n), stream);
    if (e != cudaSuccess) return e;
    constexpr int smem = (FLE_WSLOTS + FLE_PSLOTS + (FLE_INPLACE ? 0 : FLE_OSLOTS)) * FLE_TILE_BYTES;
    // persistent CTAs: as many as fit (more would only queue behind them), plus the scanner CTA (block 0, placed first)
    int dev = 0;
    e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    static int resident[64];
    int cap = dev < 64 ? resident[dev] : 0;
    if (!cap) {
        e = cudaFuncSetAttribute((const void*)fl_encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        int sms = 0, per_sm = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, (const void*)fl_encode_kernel, FLE_THREADS, smem);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) return cudaErrorLaunchOutOfResources;
        cap = sms * std::max(1, std::min(per_sm, FLE_CTAS));
        if (cap < 2) return cudaErrorLaunchOutOfResources;
        if (dev < 64) resident[dev] = cap;
    }
    const unsigned grid = (unsigned)std::min<size_t>(tiles, (size_t)cap - 1) + 1;  // the scanner CTA + the workers
    fl_encode_kernel<<<grid, FLE_THREADS, smem, stream>>>(d_in, n, d_bits, d_values, d_values_size, inc, agg, next_tile, (long long)tiles);
    return cudaGetLastError();
}

}  // namespace flrl
