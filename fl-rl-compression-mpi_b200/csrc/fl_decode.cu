// fl_decode.cu — fixed-length decoder for sm_100a: offsets pre-pass, then persistent CTAs over a moving window of tiles.
//
// Replaces compressInitializeFrameStartIndiciesBits + thrust::exclusive_scan + decompressCalculateOutput
// (/root/reference/src/fl/fl_gpu.cu:687-698, :805-808, :728-755) with one launch.
//
// Where a frame's values start depends only on the widths array (N/128 bytes), not on the payload.  So:
//   * a small PRE-PASS (fl_tile_offsets_kernel, one thread per 128-frame tile, block scan + one decoupled look-back per
//     block of 256 tiles) turns the widths into the first value unit of every tile and checks that every width is
//     1..8; it reads N/128 bytes — 0.5 % of the decoder's traffic — and removes every dependency between tiles;
//   * the decoder proper is a grid of persistent CTAs (FLD_CTAS per SM) that take tiles c, c + G, c + 2G, ...: all
//     CTAs together read and write ONE moving window of G consecutive tiles (measured with a plain streaming kernel:
//     one window beats one contiguous range per CTA, and more, smaller CTAs beat two big ones);
//   * a producer warp per CTA walks the CTA's tiles: one coalesced load brings the widths of 4 tiles, a segmented
//     warp scan gives every frame's offset inside its tile, and for each tile one TMA bulk copy of exactly its
//     16*sum(b) value bytes is issued into a ring of FLD_STAGES shared-memory stages (full / free mbarriers);
//   * the unpack warps wait on the stage's mbarrier, pull their words into registers, unpack lane-locally
//     (flrl_bits.cuh), store 2 x 16 bytes per lane (a warp writes 1 KiB of contiguous output per item, no staging
//     buffer, no store-side synchronisation) and release the stage.
#include <stdlib.h>

#include <algorithm>

#include "fl_kernels.cuh"
#include "flrl_bits.cuh"
#include "lookback.cuh"

namespace flrl {

static constexpr int FLD_TILE_FRAMES = 128;
static constexpr int FLD_TILE_BYTES = FLD_TILE_FRAMES * 128;  // 16 KiB of output per tile
#ifndef FLD_EXP
#define FLD_EXP 0  // timing experiments only: 1 = no unpack arithmetic, 2 = no stores (both: wrong output), 4 = 256-bit stores,
                   // 8 = streaming stores (st.cs), 16 = write-through stores (st.wt): none of 1, 4, 8, 16 changes the time
#endif
#ifndef FLD_STAGES_N
#define FLD_STAGES_N 4
#endif
static constexpr int FLD_STAGES = FLD_STAGES_N;
#ifndef FLD_UNPACK
#define FLD_UNPACK 256
#endif
static constexpr int FLD_UNPACK_THREADS = FLD_UNPACK;
static constexpr int FLD_THREADS = FLD_UNPACK_THREADS + 32;   // + producer warp
static constexpr int FLD_ITEMS = FLD_TILE_BYTES / 32 / FLD_UNPACK_THREADS;  // 32-byte items per thread per tile = 2
static constexpr int FLD_STAGE_BYTES = FLD_TILE_BYTES + 128;  // slack for the unpacker's over-reads
#ifndef FLD_CTAS
#define FLD_CTAS 3
#endif
static constexpr int FLD_CTAS_PER_SM = FLD_CTAS;

__device__ __forceinline__ size_t umin64(size_t a, size_t b) { return a < b ? a : b; }

struct __align__(16) FldMeta {
    uint16_t off[FLD_TILE_FRAMES];  // exclusive prefix of widths inside the tile (16-byte units)
    uint8_t b[FLD_TILE_FRAMES];     // width, 0 = no frame here
};

// widths of frames [f, f + 16) as 16 bytes; frames at or past `fend` read as 0
__device__ __forceinline__ uint4 ld_widths16(const uint8_t* __restrict__ bits, size_t f, size_t fend, bool aligned)
{
    if (aligned && f + 16 <= fend) return *reinterpret_cast<const uint4*>(bits + f);
    uint32_t w[4] = {0, 0, 0, 0};
    for (int i = 0; i < 16; ++i)
        if (f + i < fend) w[i >> 2] |= (uint32_t)bits[f + i] << (8 * (i & 3));
    return make_uint4(w[0], w[1], w[2], w[3]);
}
// clamp every width byte to 8 and report whether any of the `valid` leading bytes was 0 or > 8
__device__ __forceinline__ uint4 clamp_widths(uint4 q, int valid, bool& bad)
{
    uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int v = valid - 4 * k;
        const uint32_t keep = v >= 4 ? ~0u : (v <= 0 ? 0u : ((1u << (8 * v)) - 1u));
        const uint32_t c = __vminu4(w[k], 0x08080808u);
        const uint32_t nz = (((c & 0x7f7f7f7fu) + 0x7f7f7f7fu) | c) & 0x80808080u;  // high bit: byte non-zero
        bad = bad || (((c ^ w[k]) | (~nz & 0x80808080u)) & keep) != 0;
        w[k] = c & keep;
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}
__device__ __forceinline__ uint32_t sum_bytes16(uint4 q)
{
    const uint32_t a = (q.x & 0x00ff00ffu) + ((q.x >> 8) & 0x00ff00ffu) + (q.y & 0x00ff00ffu) + ((q.y >> 8) & 0x00ff00ffu) +
                       (q.z & 0x00ff00ffu) + ((q.z >> 8) & 0x00ff00ffu) + (q.w & 0x00ff00ffu) + ((q.w >> 8) & 0x00ff00ffu);
    return (a & 0xffffu) + (a >> 16);
}

// one look-back per CTA over 64-bit sums (called by one full warp)
__device__ __forceinline__ uint64_t lookback_sum64(const uint64_t* desc, int idx, int lane)
{
    uint64_t excl = 0;
    int j = idx - 1 - lane;
    while (true) {
        uint64_t v;
        do {
            v = j >= 0 ? ld_relaxed_u64(desc + j) : LB_INC;
        } while (__any_sync(0xffffffffu, (v >> 62) == 0));
        const unsigned incm = __ballot_sync(0xffffffffu, (v >> 62) == 2);
        if (incm) {
            const int first = __ffs(incm) - 1;
            return excl + warp_sum_u64(lane <= first ? (v & LB_MASK) : 0ull);
        }
        excl += warp_sum_u64(v & LB_MASK);
        j -= 32;
    }
}

// fl_cpu.cu:117-141 for the final partial frame (len = n % 128), FLD_THREADS threads
__device__ __noinline__ void decode_tail_frame(const uint8_t* vin, uint32_t avail, uint32_t len, uint32_t b, uint8_t* out)
{
    const int tid = threadIdx.x;
    if (tid < (int)len && b >= 1 && b <= 8) {
        const uint32_t bit = tid * b;
        const uint32_t byte = bit >> 3, sh = bit & 7;
        uint32_t v = byte < avail ? vin[byte] : 0u;
        if (byte + 1 < avail) v |= (uint32_t)vin[byte + 1] << 8;
        out[tid] = (uint8_t)((v >> sh) & ((1u << b) - 1u));
    }
}

// ---- pre-pass: first value unit (16 bytes) of every tile ------------------------------------------
// tile_base[t] = sum of the widths of all full frames before tile t, t = 0 .. tiles (the last entry is the total);
// *err = 1 if a width of a full frame is outside 1..8 (widths are clamped to 8 for the sums, as the decoder does).
static constexpr int FLO_THREADS = 256;
__global__ void __launch_bounds__(FLO_THREADS) fl_tile_offsets_kernel(const uint8_t* __restrict__ bits, size_t full_frames, long long tiles,
                                                                      uint64_t* __restrict__ tile_base, uint64_t* __restrict__ desc,
                                                                      int* __restrict__ err)
{
    __shared__ uint32_t wsum[FLO_THREADS / 32];
    __shared__ uint64_t block_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long t = (long long)blockIdx.x * FLO_THREADS + tid;
    const bool aligned = (reinterpret_cast<uintptr_t>(bits) & 15) == 0;
    // a warp reads the widths of its 32 tiles (4 KiB) with coalesced 16-byte loads: in step i lane l holds 16 widths of
    // tile 4 i + l / 8; the eight partial sums of a tile are added across lanes, and tile j's sum ends up in lane j
    __shared__ uint32_t tsum[FLO_THREADS];
    bool bad = false;
    {
        const long long wt0 = (long long)blockIdx.x * FLO_THREADS + warp * 32;  // the warp's first tile
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const size_t f = (size_t)wt0 * FLD_TILE_FRAMES + (size_t)(i * 32 + lane) * 16;
            uint32_t part = 0;
            if (f < full_frames) part = sum_bytes16(clamp_widths(ld_widths16(bits, f, full_frames, aligned), (int)umin64(16, full_frames - f), bad));
            part += __shfl_xor_sync(0xffffffffu, part, 1);
            part += __shfl_xor_sync(0xffffffffu, part, 2);
            part += __shfl_xor_sync(0xffffffffu, part, 4);
            if ((lane & 7) == 0) tsum[warp * 32 + 4 * i + (lane >> 3)] = part;
        }
        __syncwarp();
    }
    const uint32_t sum = t < tiles ? tsum[tid] : 0u;
    if (bad) atomicExch(err, 1);  // malformed container
    uint32_t inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    uint32_t wbase = 0, total = 0;
#pragma unroll
    for (int k = 0; k < FLO_THREADS / 32; ++k) {
        if (k < warp) wbase += wsum[k];
        total += wsum[k];
    }
    if (warp == 0) {
        uint64_t base = 0;
        if (blockIdx.x > 0) {
            if (lane == 0) st_relaxed_u64(desc + blockIdx.x, LB_AGG | total);
            base = lookback_sum64(desc, (int)blockIdx.x, lane);
        }
        if (lane == 0) {
            st_relaxed_u64(desc + blockIdx.x, LB_INC | (base + total));
            block_base = base;
        }
    }
    __syncthreads();
    if (t < tiles) {
        const uint64_t excl = block_base + wbase + (inc - sum);
        tile_base[t] = excl;
        if (t == tiles - 1) tile_base[tiles] = excl + sum;
    }
}

__global__ void __launch_bounds__(FLD_THREADS, FLD_CTAS_PER_SM)
fl_decode_kernel(const uint8_t* __restrict__ bits, size_t nframes, const uint8_t* __restrict__ values, size_t nvalues,
                 uint8_t* __restrict__ out, size_t n, const uint64_t* __restrict__ tile_base, int* __restrict__ err, long long tiles)
{
    extern __shared__ __align__(128) uint8_t smem[];  // FLD_STAGES x FLD_STAGE_BYTES
    __shared__ FldMeta meta[FLD_STAGES];
    __shared__ __align__(8) uint64_t full[FLD_STAGES], freeb[FLD_STAGES];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long G = gridDim.x, c = blockIdx.x;
    const long long T = (tiles - c + G - 1) / G;  // this CTA's tiles: c, c + G, c + 2G, ...
    const size_t full_frames = n / FRAME;
    const bool aligned = (reinterpret_cast<uintptr_t>(bits) & 15) == 0;

    if (tid == 0) {
        for (int s = 0; s < FLD_STAGES; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&freeb[s], FLD_UNPACK_THREADS / 32);
        }
        fence_mbar_init();
    }
    __syncthreads();

    if (warp == FLD_UNPACK_THREADS / 32) {
        // =============================== producer warp =====================================
        const uint64_t nv16 = nvalues / 16;
        const int sub = lane >> 3, part = lane & 7;
        // launched with programmatic stream serialization: everything above overlapped the pre-pass, whose tile_base is
        // needed (and visible) from here on
        asm volatile("griddepcontrol.wait;" ::: "memory");
        // widths of the 16 frames this lane scans in batch kb (tile 4 kb + sub of the CTA's list) and the tile's first value unit
        auto load_batch = [&](long long kb, uint64_t* tb) -> uint4 {
            const long long k = 4 * kb + sub;
            if (k >= T) {
                *tb = 0;
                return make_uint4(0, 0, 0, 0);
            }
            const long long tile = c + k * G;
            *tb = tile_base[tile];
            return ld_widths16(bits, (size_t)tile * FLD_TILE_FRAMES + part * 16, full_frames, aligned);
        };
        uint64_t tbn = 0;
        uint4 qn = load_batch(0, &tbn);
        for (long long kb = 0; 4 * kb < T; ++kb) {
            uint4 q = qn;
            const uint64_t tb = tbn;
            qn = load_batch(kb + 1, &tbn);  // widths of the next 4 tiles are in flight while these are issued
            const long long ktile = 4 * kb + sub < T ? c + (4 * kb + sub) * G : tiles;
            const size_t f = (size_t)ktile * FLD_TILE_FRAMES + part * 16;
            bool bad = false;  // reported by the pre-pass; here the widths are only clamped
            q = clamp_widths(q, f < full_frames ? (int)umin64(16, full_frames - f) : 0, bad);
            const uint32_t s16 = sum_bytes16(q);
            uint32_t inc = s16;
#pragma unroll
            for (int d = 1; d < 8; d <<= 1) {
                const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
                if (part >= d) inc += o;
            }
            uint32_t run = inc - s16;  // exclusive offset of this lane's 16 frames inside its tile
            const uint32_t w4[4] = {q.x, q.y, q.z, q.w};
            uint32_t o16[8];
#pragma unroll
            for (int i = 0; i < 16; i += 2) {
                const uint32_t b0 = (w4[i >> 2] >> (8 * (i & 3))) & 0xffu, b1 = (w4[i >> 2] >> (8 * (i & 3) + 8)) & 0xffu;
                o16[i >> 1] = run | ((run + b0) << 16);
                run += b0 + b1;
            }
#pragma unroll
            for (int s2 = 0; s2 < 4; ++s2) {
                const long long k = 4 * kb + s2;
                if (k >= T) break;
                const int s = (int)(k % FLD_STAGES);
                const long long it = k / FLD_STAGES;
                if (it > 0) mbar_wait(&freeb[s], (uint32_t)(it - 1) & 1u);  // the unpack warps have drained tile k - STAGES
                if (sub == s2) {
                    reinterpret_cast<uint4*>(meta[s].b)[part] = q;
                    reinterpret_cast<uint4*>(meta[s].off)[2 * part] = make_uint4(o16[0], o16[1], o16[2], o16[3]);
                    reinterpret_cast<uint4*>(meta[s].off)[2 * part + 1] = make_uint4(o16[4], o16[5], o16[6], o16[7]);
                }
                const uint32_t tot = __shfl_sync(0xffffffffu, inc, 8 * s2 + 7);
                const uint64_t base = __shfl_sync(0xffffffffu, tb, 8 * s2);
                __syncwarp();
                if (lane == 0) {
                    uint64_t units = tot;
                    if (base + units > nv16) {  // values shorter than the widths imply: never read past the array
                        atomicExch(err, 2);
                        units = base < nv16 ? nv16 - base : 0;
                    }
                    if (units) {
                        mbar_expect_tx(&full[s], (uint32_t)units * 16);
                        tma_load_1d(smem + (size_t)s * FLD_STAGE_BYTES, values + base * 16, (uint32_t)units * 16, &full[s]);
                    } else {
                        mbar_arrive(&full[s]);
                    }
                }
            }
        }
    } else {
        // =============================== unpack warps ======================================
        const int t = tid & 3;
        for (long long k = 0; k < T; ++k) {
            const int s = (int)(k % FLD_STAGES);
            mbar_wait(&full[s], (uint32_t)(k / FLD_STAGES) & 1u);
            const uint32_t* vw = reinterpret_cast<const uint32_t*>(smem + (size_t)s * FLD_STAGE_BYTES);
            uint32_t bw[FLD_ITEMS], r[FLD_ITEMS][4], g[FLD_ITEMS][5];
#pragma unroll
            for (int i = 0; i < FLD_ITEMS; ++i) {
                const int fr = i * (FLD_UNPACK_THREADS / 4) + (tid >> 2);
                const uint32_t b = meta[s].b[fr];
                const uint32_t* w = vw + (uint32_t)meta[s].off[fr] * 4u + b * t;
                const uint32_t q = b >> 1;
                bw[i] = b;
#pragma unroll
                for (int j = 0; j < 4; ++j) r[i][j] = w[j];
#pragma unroll
                for (int j = 0; j < 5; ++j) g[i][j] = w[q + j];
            }
            uint8_t* otile = out + (size_t)(c + k * G) * FLD_TILE_BYTES;
#pragma unroll
            for (int i = 0; i < FLD_ITEMS; ++i) {
                const int fr = i * (FLD_UNPACK_THREADS / 4) + (tid >> 2);
                if (bw[i]) {
                    const UnpackK kk = make_unpack_k(bw[i]);
                    uint32_t a[4], cc[4];
#if (FLD_EXP & 1)
                    (void)kk;
                    for (int j = 0; j < 4; ++j) a[j] = r[i][j], cc[j] = g[i][j] ^ g[i][4];
#else
                    unpack32_regs(bw[i], kk, r[i], g[i], a, cc);
#endif
                    uint4* o = reinterpret_cast<uint4*>(otile + fr * 128 + t * 32);
#if (FLD_EXP & 2)
                    if (a[0] == 0x12345679u && cc[3] == 0x9abcdef1u) o[0] = make_uint4(a[0] ^ a[1], a[2] ^ a[3], cc[0] ^ cc[1], cc[2]);
#elif (FLD_EXP & 4)
                    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(o), "r"(a[0]), "r"(a[1]), "r"(a[2]),
                                 "r"(a[3]), "r"(cc[0]), "r"(cc[1]), "r"(cc[2]), "r"(cc[3])
                                 : "memory");
#elif (FLD_EXP & 8)
                    __stcs(o, make_uint4(a[0], a[1], a[2], a[3]));
                    __stcs(o + 1, make_uint4(cc[0], cc[1], cc[2], cc[3]));
#elif (FLD_EXP & 16)
                    __stwt(o, make_uint4(a[0], a[1], a[2], a[3]));
                    __stwt(o + 1, make_uint4(cc[0], cc[1], cc[2], cc[3]));
#else
                    o[0] = make_uint4(a[0], a[1], a[2], a[3]);
                    o[1] = make_uint4(cc[0], cc[1], cc[2], cc[3]);
#endif
                }
            }
            // Release the stage only now, after its words were used: an arrive right behind the loads (everything is "in
            // registers") can overtake them (seen in fl_encode); releasing early measured no faster anyway.
            __syncwarp();
            if (lane == 0) mbar_arrive(&freeb[s]);
        }
    }
    // ---- the first CTA: the final partial frame ----------------------------------------------------
    if (c != 0) return;
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const uint32_t tail = (uint32_t)(n % FRAME);
    if (tail && nframes == full_frames + 1) {
        const uint64_t off = tile_base[tiles] * 16;  // value bytes of all full frames
        const uint32_t b = bits[full_frames];
        const uint64_t avail = nvalues > off ? nvalues - off : 0;
        if (b == 0 || b > 8 || ((uint64_t)tail * b + 7) / 8 > avail) {
            if (tid == 0) atomicExch(err, 3);
        } else {
            decode_tail_frame(values + off, (uint32_t)(avail > 128 ? 128 : avail), tail, b, out + full_frames * FRAME);
        }
    }
}

// =============================================================================================
//                                        launcher
// =============================================================================================
const int* fl_decode_status(const void* d_ws, size_t n)
{
    return reinterpret_cast<const int*>(reinterpret_cast<const uint64_t*>(d_ws) + fl_tiles(n));
}

cudaError_t fl_decode_launch(const uint8_t* d_bits, size_t nframes, const uint8_t* d_values, size_t nvalues, uint8_t* d_out,
                             size_t n, void* d_ws, cudaStream_t stream)
{
    const size_t tiles = fl_tiles(n);
    int dev = 0, sms = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    static int sm_count[64];
    if (dev < 64 && sm_count[dev]) {
        sms = sm_count[dev];
    } else {
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return e;
        if (dev < 64) sm_count[dev] = sms;
    }
    // workspace: [tiles words unused][status word][look-back descriptors of the pre-pass: one per 256 tiles][tile_base: tiles + 1 words]
    uint64_t* ws = reinterpret_cast<uint64_t*>(d_ws);
    int* err = reinterpret_cast<int*>(ws + tiles);
    uint64_t* desc = ws + tiles + 1;
    const unsigned oblocks = (unsigned)((tiles + FLO_THREADS - 1) / FLO_THREADS);
    uint64_t* tile_base = desc + oblocks;
    e = cudaMemsetAsync(err, 0, (1 + oblocks) * sizeof(uint64_t), stream);  // status word + descriptors, adjacent
    if (e != cudaSuccess) return e;
    fl_tile_offsets_kernel<<<oblocks, FLO_THREADS, 0, stream>>>(d_bits, n / FRAME, (long long)tiles, tile_base, desc, err);
    const unsigned grid = (unsigned)std::min<size_t>(tiles, (size_t)sms * FLD_CTAS_PER_SM);
    constexpr int smem = FLD_STAGES * FLD_STAGE_BYTES;
    static bool done[64];
    if (!(dev < 64 && done[dev])) {
        e = cudaFuncSetAttribute((const void*)fl_decode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        if (dev < 64) done[dev] = true;
    }
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(FLD_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // the decoder's set-up overlaps the pre-pass
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, fl_decode_kernel, d_bits, nframes, d_values, nvalues, d_out, n, (const uint64_t*)tile_base, err, (long long)tiles);
}

}  // namespace flrl

// =============================================================================================
//            flrl_fl_decode_static — the decoder with NO allocation anywhere on its path
// =============================================================================================
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
// No counterpart in the reference.  The decoder's scratch (status word,
// look-back descriptors of the pre-pass, first value unit of every tile: 2 words per 16 KiB tile) lives in the
// file-scope device array below, sized for FLRL_STATIC_MAX_BYTES = 1 GiB of output per call: 1.0 MiB, held by every
// GPU that loads the library until the process exits.  In host code the array's NAME is a host-side stand-in, not a
// device address: the address comes from cudaGetSymbolAddress, once per device, with that device current.
// Every call on a device shares the array, which keeps the last call's contents: fl_decode_launch clears what must start
// at zero (status word + descriptors) on the caller's stream, everything else is written before it is read.  Calls that
// OVERLAP on one GPU (two streams) race on it: serialise them, or use flrl_fl_decode with a workspace per call.
#include <mutex>

#include "abi_common.cuh"

namespace flrl {
static constexpr size_t FLD_STATIC_TILES = (size_t)FLRL_STATIC_MAX_BYTES / FLD_TILE_BYTES;               // 65536
static constexpr size_t FLD_STATIC_WORDS = 2 * FLD_STATIC_TILES + FLD_STATIC_TILES / FLO_THREADS + 16;  // what fl_decode_launch lays out
static __device__ __align__(256) uint64_t g_fld_static_scratch[FLD_STATIC_WORDS];
}  // namespace flrl

extern "C" int flrl_fl_decode_static(const uint8_t* d_bits, size_t nbits, const uint8_t* d_values, size_t nvalues, uint8_t* d_out, size_t n,
                                     int* d_status, flrl_stream_t stream)
{
    using namespace flrl;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (n > (size_t)FLRL_STATIC_MAX_BYTES)
        return fail(FLRL_E_CAPACITY, "flrl_fl_decode_static: n = %zu exceeds FLRL_STATIC_MAX_BYTES = %zu (the static scratch is sized for that)", n,
                    (size_t)FLRL_STATIC_MAX_BYTES);
    if (n == 0) {
        if (d_status) {
            cudaError_t e = cudaMemsetAsync(d_status, 0, sizeof(int), s);
            if (e != cudaSuccess) return fail_cuda(e, "flrl_fl_decode_static/memset");
        }
        return FLRL_OK;
    }
    if (!d_bits || !d_values || !d_out) return fail(FLRL_E_ARG, "flrl_fl_decode_static: NULL buffer");
    if (nbits != (n + FRAME - 1) / FRAME) return fail(FLRL_E_FORMAT, "flrl_fl_decode_static: bitsSize != ceil(outputSize/128)");
    if (!aligned16(d_values) || !aligned16(d_out)) return fail(FLRL_E_ALIGN, "flrl_fl_decode_static: d_values and d_out must be 16-byte aligned");
    static_assert(FLD_STATIC_WORDS * sizeof(uint64_t) >= (2 * FLD_STATIC_TILES + FLD_STATIC_TILES / 256 + 8) * sizeof(uint64_t) + 16,
                  "the scratch holds the decoder's part of fl_workspace_bytes(FLRL_STATIC_MAX_BYTES)");
    // the array's device address on the CURRENT device, looked up once per device
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail_cuda(e, "flrl_fl_decode_static/cudaGetDevice");
    static std::mutex mu;
    static void* addr[64];
    void* ws = nullptr;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (dev < 64 && addr[dev]) {
            ws = addr[dev];
        } else {
            e = cudaGetSymbolAddress(&ws, g_fld_static_scratch);
            if (e != cudaSuccess) return fail_cuda(e, "flrl_fl_decode_static/cudaGetSymbolAddress");
            if (dev < 64) addr[dev] = ws;
        }
    }
    e = fl_decode_launch(d_bits, nbits, d_values, nvalues, d_out, n, ws, s);  // memsets + two launches on `s`; allocates nothing
    if (e != cudaSuccess) return fail_cuda(e, "flrl_fl_decode_static");
    if (d_status) {
        e = cudaMemcpyAsync(d_status, fl_decode_status(ws, n), sizeof(int), cudaMemcpyDeviceToDevice, s);
        if (e != cudaSuccess) return fail_cuda(e, "flrl_fl_decode_static/status");
    }
    return FLRL_OK;
}
