// fl_kernels.cu — fixed-length codec for sm_100a.
//
// Replaces the reference's four kernels + thrust scan + byte-atomic CAS loops
// (/root/reference/src/fl/fl_gpu.cu:648-808) with ONE pass per direction:
//
//   encode: TMA bulk load of a 128-frame tile (16 KiB) -> per-frame width from the OR of the
//           frame (4 lanes x 32 bytes, two shuffles) -> in-tile scan of widths -> lane-local
//           multiply/shift packing (flrl_bits.cuh) written IN PLACE over the staged tile ->
//           decoupled look-back for the tile's output offset -> TMA bulk store.
//           Input is read once; no atomics on data; no frame-start array.
//   decode: widths of the tile -> scan + look-back -> TMA bulk load of exactly the tile's
//           16*sum(b) value bytes -> lane-local unpack -> smem -> TMA bulk store.
//
// Every full frame packs to 16*b bytes, so every tile's value range starts and ends on a
// 16-byte boundary: all bulk copies are legal (size % 16 == 0, 16-byte aligned).  Only the
// final partial frame (n % 128 != 0) goes through a small cooperative scalar path in the
// last tile, which also publishes the total value size.
#include "fl_kernels.cuh"

#include <stdlib.h>

#include "flrl_bits.cuh"
#include "lookback.cuh"

namespace flrl {

static constexpr int TILE_FRAMES = 128;
static constexpr int TILE_BYTES = TILE_FRAMES * 128;  // 16 KiB
static constexpr int THREADS = 128;
static constexpr int ITEMS = 4;  // 32-byte items per thread per tile
static constexpr long long FL_LEAD_TILES = 2048;  // distance between the two touches of a tile (32 MiB)

// Development trace (FLRL_TRACE builds only): per-tile globaltimer stamps.
#ifdef FLRL_TRACE
__device__ uint64_t* g_trace = nullptr;
__device__ __forceinline__ void trace(long long tile, int slot)
{
    if (g_trace) {
        uint64_t t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        g_trace[tile * 12 + slot] = t;
    }
}
#define TRACE(tile, slot, cond) do { if (cond) trace(tile, slot); } while (0)
#else
#define TRACE(tile, slot, cond) do { } while (0)
#endif

struct __align__(16) TileScan {
    uint16_t off[TILE_FRAMES];  // exclusive prefix of widths inside the tile (16-byte units)
    uint8_t b[TILE_FRAMES];
    uint32_t total;
    uint64_t base;  // exclusive prefix over all earlier tiles (16-byte units)
};

// warp 0: exclusive scan of sc.b[0..128) -> sc.off, sc.total.  Each lane owns 4 frames.
__device__ __forceinline__ uint32_t tile_scan_widths(TileScan& sc, int lane)
{
    const uint32_t v = reinterpret_cast<const uint32_t*>(sc.b)[lane];
    const uint32_t b0 = v & 0xff, b1 = (v >> 8) & 0xff, b2 = (v >> 16) & 0xff, b3 = v >> 24;
    const uint32_t s = b0 + b1 + b2 + b3;
    uint32_t inc = s;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    const uint32_t ex = inc - s;
    ushort4 o4 = make_ushort4((unsigned short)ex, (unsigned short)(ex + b0), (unsigned short)(ex + b0 + b1),
                              (unsigned short)(ex + b0 + b1 + b2));
    reinterpret_cast<ushort4*>(sc.off)[lane] = o4;
    return __shfl_sync(0xffffffffu, inc, 31);
}

// ---- final partial frame (len = n % 128 in 1..127), all THREADS threads participate --------
// fl_cpu.cu:37-50 (width), :64-84 (packing), :53 (size = ceil(bits/8)).
__device__ __noinline__ uint32_t encode_tail_frame(const uint8_t* in, uint32_t len, uint8_t* bits_out, uint8_t* vout,
                                                   uint32_t* scratch /* >= 35 words smem */)
{
    const int tid = threadIdx.x;
    const uint32_t v = (tid < (int)len) ? in[tid] : 0u;
    if (tid < 35) scratch[tid] = 0;
    __syncthreads();
    const uint32_t wo = __reduce_or_sync(0xffffffffu, v);
    if ((tid & 31) == 0 && wo) atomicOr(&scratch[34], wo);
    __syncthreads();
    const uint32_t b = width_of_or(scratch[34]);
    if (tid < (int)len) {
        const uint32_t bit = tid * b;
        const uint32_t w = bit >> 5, sh = bit & 31;
        atomicOr(&scratch[w], v << sh);
        if (sh + b > 32) atomicOr(&scratch[w + 1], v >> (32 - sh));
    }
    __syncthreads();
    const uint32_t nbytes = (len * b + 7) >> 3;
    if (tid < (int)nbytes) vout[tid] = (uint8_t)(scratch[tid >> 2] >> (8 * (tid & 3)));
    if (tid == 0) *bits_out = (uint8_t)b;
    return nbytes;
}

// =============================================================================================
//                                         ENCODE
// =============================================================================================
// Two touches per tile, `lead` tiles apart, in ONE kernel:
//   CTA c  = width pass of tile c            (first touch: DRAM -> L2 -> smem; posts the tile's
//                                             aggregate and writes its 128 widths)
//          + pack  pass of tile c - lead     (second touch: an L2 hit; its own aggregate and those
//                                             of all predecessors were posted `lead` CTAs ago, so
//                                             the look-back never waits on a straggling load)
// Why: the value offset of a tile needs the widths of every earlier tile.  In a classic
// one-touch decoupled look-back the packed tile sits in shared memory until the slowest
// earlier load has landed; measured on B200 that wait was 6.5 us of a 10 us CTA lifetime
// and capped 1 GiB encode at 0.45 ms no matter how wide the look-back was.  `lead` (2048
// tiles = 32 MiB) exceeds the ~1000 resident CTAs and stays far inside the 126 MB L2, so the
// input is still read from DRAM once.
//
// CTA = 4 pack warps + 1 control warp.  Control: both TMA loads, look-back + inclusive
// prefix at CTA start, TMA store.  Pack warps: pack tile c-lead in place, then widths of tile c.
static constexpr int ENC_THREADS = THREADS + 32;
static constexpr uint32_t BAR_OFFSETS = 2, BAR_PACKED = 3, BAR_PACKW_A = 4, BAR_PACKW_C = 6;

// 16 width bytes -> 16 nibbles; frames at or past `valid` get the filler 0xf (never 0)
__device__ __forceinline__ uint64_t nibbles_of_bytes(uint4 q, int valid)
{
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
    uint64_t r = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const uint32_t b = (w[i >> 2] >> (8 * (i & 3))) & 0xf;
        r |= (uint64_t)(i < valid ? b : 0xfu) << (4 * i);
    }
    return r;
}
__device__ __forceinline__ uint4 bytes_of_nibbles(uint64_t x, int valid)
{
    uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const uint32_t b = (uint32_t)(x >> (4 * i)) & 0xf;
        w[i >> 2] |= (i < valid ? b : 0u) << (8 * (i & 3));
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

__device__ __forceinline__ int tile_full_frames(size_t full_frames, size_t f0)
{
    return (int)(full_frames > f0 ? (full_frames - f0 < (size_t)TILE_FRAMES ? full_frames - f0 : TILE_FRAMES) : 0);
}

// widths of the 4 frame-quarters this lane owns in a staged tile; optionally keeps the data
template <bool KEEP>
__device__ __forceinline__ void lane_widths(const uint8_t* tile, int tid, int nfr, uint4 (&A)[ITEMS], uint4 (&C)[ITEMS],
                                            uint32_t (&bw)[ITEMS], uint8_t* sb)
{
    const int t = tid & 3;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const int fr = i * 32 + (tid >> 2);
        // unconditional: frames past nfr read stale shared memory and get width 0 (no stores)
        const uint4* p = reinterpret_cast<const uint4*>(tile + fr * 128 + t * 32);
        const uint4 a = p[0], c = p[1];
        if (KEEP) {
            A[i] = a;
            C[i] = c;
        }
        uint32_t o = (a.x | a.y | a.z) | (a.w | c.x | c.y) | (c.z | c.w);
        o |= __shfl_xor_sync(0xffffffffu, o, 1);
        o |= __shfl_xor_sync(0xffffffffu, o, 2);
        o |= o >> 16;
        o |= o >> 8;
        const uint32_t b = (fr < nfr) ? width_of_or(o) : 0u;
        bw[i] = b;
        if (t == 0) sb[fr] = (uint8_t)b;
    }
}

__global__ void __launch_bounds__(ENC_THREADS, 6)
fl_encode_kernel(const uint8_t* __restrict__ in, size_t n, uint8_t* __restrict__ bits, uint8_t* __restrict__ values,
                 uint64_t* __restrict__ values_size, uint64_t* __restrict__ desc, uint64_t* __restrict__ wdesc,
                 long long tiles, long long lead)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* tileP = smem;               // pack pass: input tile, then packed output in place
    uint8_t* tileW = smem + TILE_BYTES;  // width pass
    __shared__ __align__(8) uint64_t barP, barW;
    __shared__ TileScan sc;
    __shared__ __align__(16) uint8_t wb[TILE_FRAMES];
    __shared__ __align__(16) uint4 ktab[9];  // PackK{k1,k2,k3,ra} per width
    __shared__ uint32_t tail_scratch[36];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t full_frames = n / FRAME;
    const long long cw = blockIdx.x;         // tile of the width pass (valid if < tiles)
    const long long cp = cw - lead;          // tile of the pack pass  (valid if >= 0)
    const bool doW = cw < tiles, doP = cp >= 0;
    const size_t fW = (size_t)cw * TILE_FRAMES, fP = doP ? (size_t)cp * TILE_FRAMES : 0;
    const int nfrW = doW ? tile_full_frames(full_frames, fW) : 0;
    const int nfrP = doP ? tile_full_frames(full_frames, fP) : 0;
    const bool lastP = doP && cp == tiles - 1;

    if (tid == THREADS) {
        TRACE(cw, 0, true);
        mbar_init(&barP, 1);
        mbar_init(&barW, 1);
        fence_mbar_init();
        if (nfrP) {  // second touch: L2 hit, not needed afterwards
            mbar_expect_tx(&barP, (uint32_t)nfrP * FRAME);
            tma_load_1d_hint(tileP, in + fP * FRAME, (uint32_t)nfrP * FRAME, &barP, l2_policy_evict_first());
        }
        if (nfrW) {  // first touch: from DRAM, keep in L2 for the pack pass `lead` CTAs later
            mbar_expect_tx(&barW, (uint32_t)nfrW * FRAME);
            tma_load_1d_hint(tileW, in + fW * FRAME, (uint32_t)nfrW * FRAME, &barW, l2_policy_evict_last());
        }
    }
    if (tid <= 8) {
        const PackK k = make_pack_k(tid ? tid : 8);
        ktab[tid] = tid ? make_uint4(k.k1, k.k2, k.k3, k.ra) : make_uint4(0, 0, 0, 0);
    }
    __syncthreads();

    if (warp < 4) {
        // =============================== pack warps ========================================
        uint4 A[ITEMS], C[ITEMS];
        uint32_t bw[ITEMS];
        if (doP) {
            const int t = tid & 3;
            bar_sync(BAR_OFFSETS, ENC_THREADS);  // control warp: sc.b (widths from the width pass) and sc.off
            if (nfrP) mbar_wait(&barP, 0);
            TRACE(cw, 1, tid == 0);
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) {
                const int fr = i * 32 + (tid >> 2);
                const uint4* p = reinterpret_cast<const uint4*>(tileP + fr * 128 + t * 32);
                A[i] = p[0];
                C[i] = p[1];
                bw[i] = sc.b[fr];
            }
            bar_sync(BAR_PACKW_A, THREADS);  // every read of the staged tile is done: pack in place
            TRACE(cw, 2, tid == 0);
            const uint32_t tw = smem_u32(tileP);
#pragma unroll
            for (int i = 0; i < ITEMS; ++i) {
                const int fr = i * 32 + (tid >> 2);
                const uint32_t b = bw[i];  // 0 for frames past nfr: every store below is predicated off
                const uint4 kk = ktab[b];
                PackK k;
                k.k1 = kk.x; k.k2 = kk.y; k.k3 = kk.z; k.ra = kk.w; k.s8 = 8 * b;
                const uint32_t w = tw + ((uint32_t)sc.off[fr] * 4u + b * t) * 4u;
                const uint32_t a[4] = {A[i].x, A[i].y, A[i].z, A[i].w};
                const uint32_t c[4] = {C[i].x, C[i].y, C[i].z, C[i].w};
                pack32_store(a, c, b, k, [&](bool p, uint32_t j, uint32_t v) { sts32_pred(w + 4u * j, v, p); });
            }
            fence_proxy_async_smem();
            TRACE(cw, 5, tid == 0);
            bar_arrive(BAR_PACKED, ENC_THREADS);
        }
        if (doW) {
            if (nfrW) mbar_wait(&barW, 0);
            TRACE(cw, 3, tid == 0);
            lane_widths<false>(tileW, tid, nfrW, A, C, bw, wb);
            bar_sync(BAR_PACKW_C, THREADS);
            if (warp == 0) {
                const uint32_t v = reinterpret_cast<const uint32_t*>(wb)[lane];
                uint32_t s = (v & 0xff) + ((v >> 8) & 0xff) + ((v >> 16) & 0xff) + (v >> 24);
                s = __reduce_add_sync(0xffffffffu, s);
                if (nfrW == TILE_FRAMES && (reinterpret_cast<uintptr_t>(bits) & 3) == 0) {
                    reinterpret_cast<uint32_t*>(bits + fW)[lane] = v;
                } else {
                    for (int f = lane; f < nfrW; f += 32) bits[fW + f] = wb[f];
                }
                // hand the widths to the pack pass of this tile: 16 nibbles per word, each word
                // self-validating (a width is never 0), so relaxed stores/loads suffice — no fences
                if (lane < 8) {
                    const uint4 q = reinterpret_cast<const uint4*>(wb)[lane];
                    st_relaxed_u64(wdesc + cw * 8 + lane, nibbles_of_bytes(q, nfrW - 16 * lane));
                }
                if (lane == 0) st_relaxed_u64(desc + cw, LB_AGG | s);
            }
            TRACE(cw, 4, tid == 0);
        }
        if (!lastP) return;
    } else {
        // =============================== control warp ======================================
        if (!doP) return;
        // widths of tile cp were published by CTA cp's width pass `lead` CTAs ago
        if (lane < 8) {
            uint64_t w;
            do { w = ld_relaxed_u64(wdesc + cp * 8 + lane); } while (w == 0);
            reinterpret_cast<uint4*>(sc.b)[lane] = bytes_of_nibbles(w, nfrP - 16 * lane);
        }
        __syncwarp();
        const uint32_t total = tile_scan_widths(sc, lane);
        if (lane == 0) sc.total = total;
        __syncwarp();
        bar_arrive(BAR_OFFSETS, ENC_THREADS);
        const uint64_t excl = lookback_exclusive_sum(desc, cp, lane);
        if (lane == 0) st_relaxed_u64(desc + cp, LB_INC | (excl + total));
        TRACE(cw, 6, lane == 0);
        bar_sync(BAR_PACKED, ENC_THREADS);
        if (lane == 0) {
            if (total) {
                tma_store_1d_hint(values + excl * 16, tileP, total * 16, l2_policy_evict_first());
                tma_store_commit();
            }
            sc.base = excl;
        }
        if (!lastP) {
            if (lane == 0) tma_store_wait_read<0>();  // smem must outlive the bulk store's reads
            TRACE(cw, 7, lane == 0);
            return;
        }
    }
    // ---- CTA of the last tile only: the partial frame and the total size ----------------------
    __syncthreads();
    {
        const uint32_t tail = (uint32_t)(n % FRAME);
        uint64_t vsize = (sc.base + sc.total) * 16;
        if (tail) {
            vsize += encode_tail_frame(in + full_frames * FRAME, tail, bits + full_frames, values + vsize, tail_scratch);
        }
        if (tid == 0) *values_size = vsize;
    }
    if (tid == THREADS) tma_store_wait_read<0>();
}

// =============================================================================================
//                                        launchers
// =============================================================================================
size_t fl_tiles(size_t n)
{
    const size_t frames = (n + FRAME - 1) / FRAME;
    return (frames + TILE_FRAMES - 1) / TILE_FRAMES;
}
// look-back descriptors (1 word/tile) + published widths (8 words/tile, encode) + status
size_t fl_workspace_bytes(size_t n) { return (9 * fl_tiles(n) + 4) * sizeof(uint64_t) + 16; }

#ifdef FLRL_TRACE
extern "C" int flrl_debug_set_trace(void* d_buf)
{
    uint64_t* p = static_cast<uint64_t*>(d_buf);
    return (int)cudaMemcpyToSymbol(g_trace, &p, sizeof p);
}
#endif

// the opt-in shared-memory size is a per-device attribute: set it once per device
static cudaError_t set_smem_once(const void* fn, int bytes, bool* done /*[64]*/)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 64 && done[dev]) return cudaSuccess;
    e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess && dev < 64) done[dev] = true;
    return e;
}

cudaError_t fl_encode_launch(const uint8_t* d_in, size_t n, uint8_t* d_bits, uint8_t* d_values, uint64_t* d_values_size,
                             void* d_ws, cudaStream_t stream)
{
    const size_t tiles = fl_tiles(n);
    uint64_t* desc = reinterpret_cast<uint64_t*>(d_ws);
    uint64_t* wdesc = desc + ((tiles + 1) & ~(size_t)1);
    cudaError_t e = cudaMemsetAsync(desc, 0, (((tiles + 1) & ~(size_t)1) + 8 * tiles) * sizeof(uint64_t), stream);
    if (e != cudaSuccess) return e;
    static const long long lead_cfg = getenv("FLRL_FL_LEAD") ? atoll(getenv("FLRL_FL_LEAD")) : FL_LEAD_TILES;
    const long long lead = lead_cfg < (long long)tiles ? lead_cfg : (long long)tiles;
    constexpr int smem = 2 * TILE_BYTES;
    static bool done[64];
    e = set_smem_once((const void*)fl_encode_kernel, smem, done);
    if (e != cudaSuccess) return e;
    fl_encode_kernel<<<(unsigned)(tiles + lead), ENC_THREADS, smem, stream>>>(d_in, n, d_bits, d_values, d_values_size, desc,
                                                                              wdesc, (long long)tiles, lead);
    return cudaGetLastError();
}

}  // namespace flrl
