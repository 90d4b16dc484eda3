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
#include "flrl_bits.cuh"
#include "lookback.cuh"

namespace flrl {

static constexpr int TILE_FRAMES = 128;
static constexpr int TILE_BYTES = TILE_FRAMES * 128;  // 16 KiB
static constexpr int THREADS = 128;
static constexpr int ITEMS = 4;  // 32-byte items per thread per tile

struct TileScan {
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

// =============================================================================================
//                                         ENCODE
// =============================================================================================
__global__ void __launch_bounds__(THREADS, 8)
fl_encode_kernel(const uint8_t* __restrict__ in, size_t n, uint8_t* __restrict__ bits, uint8_t* __restrict__ values,
                 uint64_t* __restrict__ values_size, uint64_t* __restrict__ desc)
{
    extern __shared__ __align__(128) uint8_t tile[];  // TILE_BYTES: input tile, then packed output in place
    __shared__ __align__(8) uint64_t bar;
    __shared__ TileScan sc;
    __shared__ uint32_t tail_scratch[36];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int t = tid & 3;  // quarter of the frame owned by this lane
    const size_t full_frames = n / FRAME;
    const long long tileno = blockIdx.x;
    const size_t f0 = (size_t)tileno * TILE_FRAMES;
    const int nfr = (int)(full_frames > f0 ? (full_frames - f0 < (size_t)TILE_FRAMES ? full_frames - f0 : TILE_FRAMES) : 0);

    if (tid == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
        if (nfr) {
            mbar_expect_tx(&bar, (uint32_t)nfr * FRAME);
            tma_load_1d(tile, in + f0 * FRAME, (uint32_t)nfr * FRAME, &bar);
        }
    }
    __syncthreads();
    if (nfr) mbar_wait(&bar, 0);

    // ---- stage 1: items to registers, width per frame -------------------------------------
    uint4 A[ITEMS], C[ITEMS];
    uint32_t bw[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const int fr = i * 32 + (tid >> 2);
        uint32_t o = 0;
        if (fr < nfr) {
            const uint4* p = reinterpret_cast<const uint4*>(tile + fr * 128 + t * 32);
            A[i] = p[0];
            C[i] = p[1];
            o = (A[i].x | A[i].y | A[i].z) | (A[i].w | C[i].x | C[i].y) | (C[i].z | C[i].w);
        } else {
            A[i] = make_uint4(0, 0, 0, 0);
            C[i] = make_uint4(0, 0, 0, 0);
        }
        o |= __shfl_xor_sync(0xffffffffu, o, 1);
        o |= __shfl_xor_sync(0xffffffffu, o, 2);
        o |= o >> 16;
        o |= o >> 8;
        const uint32_t b = (fr < nfr) ? width_of_or(o) : 0u;
        bw[i] = b;
        if (t == 0) sc.b[fr] = (uint8_t)b;
    }
    __syncthreads();

    // ---- stage 2: widths out, in-tile scan, publish aggregate -----------------------------
    if (tid < nfr) bits[f0 + tid] = sc.b[tid];
    if (warp == 0) {
        const uint32_t total = tile_scan_widths(sc, lane);
        if (lane == 0) {
            sc.total = total;
            st_relaxed_u64(desc + tileno, (tileno == 0 ? LB_INC : LB_AGG) | total);
        }
    }
    __syncthreads();  // also: every thread's reads of the staged tile are complete -> in-place writes are safe

    // ---- stage 3: pack in place ------------------------------------------------------------
    uint32_t* tw = reinterpret_cast<uint32_t*>(tile);
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const int fr = i * 32 + (tid >> 2);
        if (fr < nfr) {
            const uint32_t b = bw[i];
            const PackK k = make_pack_k(b);
            uint32_t* w = tw + (uint32_t)sc.off[fr] * 4u + b * t;
            const uint32_t a[4] = {A[i].x, A[i].y, A[i].z, A[i].w};
            const uint32_t c[4] = {C[i].x, C[i].y, C[i].z, C[i].w};
            pack32_store(a, c, b, k, [&](uint32_t j, uint32_t v) { w[j] = v; });
        }
    }
    fence_proxy_async_smem();

    // ---- stage 4: look-back, then one bulk store -------------------------------------------
    if (warp == 0) {
        uint64_t excl = 0;
        if (tileno > 0) {
            excl = lookback_exclusive_sum(desc, tileno, lane);
            if (lane == 0) st_relaxed_u64(desc + tileno, LB_INC | (excl + sc.total));
        }
        if (lane == 0) sc.base = excl;
    }
    __syncthreads();
    const uint32_t total = sc.total;
    const uint64_t base = sc.base;
    if (tid == 0 && total) {
        tma_store_1d(values + base * 16, tile, total * 16);
        tma_store_commit();
    }
    if (tileno == (long long)gridDim.x - 1) {
        const uint32_t tail = (uint32_t)(n % FRAME);
        uint64_t vsize = (base + total) * 16;
        if (tail) {
            vsize += encode_tail_frame(in + full_frames * FRAME, tail, bits + full_frames, values + vsize, tail_scratch);
        }
        if (tid == 0) *values_size = vsize;
    }
    if (tid == 0) tma_store_wait_read<0>();  // smem must outlive the bulk store's reads
}

// =============================================================================================
//                                         DECODE
// =============================================================================================
__global__ void __launch_bounds__(THREADS, 6)
fl_decode_kernel(const uint8_t* __restrict__ bits, size_t nframes, const uint8_t* __restrict__ values, size_t nvalues,
                 uint8_t* __restrict__ out, size_t n, uint64_t* __restrict__ desc, int* __restrict__ err)
{
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* vin = smem;                     // TILE_BYTES + 64 (slack for the unpacker's over-reads)
    uint8_t* vout = smem + TILE_BYTES + 128;  // TILE_BYTES
    __shared__ __align__(8) uint64_t bar;
    __shared__ TileScan sc;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int t = tid & 3;
    const size_t full_frames = n / FRAME;
    const long long tileno = blockIdx.x;
    const size_t f0 = (size_t)tileno * TILE_FRAMES;
    const int nfr = (int)(full_frames > f0 ? (full_frames - f0 < (size_t)TILE_FRAMES ? full_frames - f0 : TILE_FRAMES) : 0);

    if (tid == 0) {
        mbar_init(&bar, 1);
        fence_mbar_init();
    }
    {
        uint32_t b = 0;
        if (tid < nfr) {
            b = bits[f0 + tid];
            if (b > 8u || b == 0u) {  // malformed container: never index out of range
                atomicExch(err, 1);
                b = b ? 8u : 0u;
            }
        }
        sc.b[tid] = (uint8_t)b;
    }
    __syncthreads();

    if (warp == 0) {
        uint32_t total = tile_scan_widths(sc, lane);
        uint64_t excl = 0;
        if (tileno > 0) {
            if (lane == 0) st_relaxed_u64(desc + tileno, LB_AGG | total);
            excl = lookback_exclusive_sum(desc, tileno, lane);
        }
        if (lane == 0) {
            st_relaxed_u64(desc + tileno, LB_INC | (excl + total));
            if ((excl + total) * 16 > nvalues) {  // truncated values array
                atomicExch(err, 2);
                total = 0;
            }
            sc.total = total;
            sc.base = excl;
            if (total) {
                mbar_expect_tx(&bar, total * 16);
                tma_load_1d(vin, values + excl * 16, total * 16, &bar);
            }
        }
    }
    __syncthreads();
    const uint32_t total = sc.total;
    if (total) {
        mbar_wait(&bar, 0);
        const uint32_t* vw = reinterpret_cast<const uint32_t*>(vin);
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const int fr = i * 32 + (tid >> 2);
            if (fr < nfr) {
                const uint32_t b = sc.b[fr];
                const UnpackK k = make_unpack_k(b ? b : 1u);
                const uint32_t* w = vw + (uint32_t)sc.off[fr] * 4u + b * t;
                uint32_t a[4], c[4];
                unpack32_load(b, k, [&](uint32_t j) { return w[j]; }, a, c);
                if (b == 0) { a[0] = a[1] = a[2] = a[3] = 0; c[0] = c[1] = c[2] = c[3] = 0; }
                uint4* o = reinterpret_cast<uint4*>(vout + fr * 128 + t * 32);
                o[0] = make_uint4(a[0], a[1], a[2], a[3]);
                o[1] = make_uint4(c[0], c[1], c[2], c[3]);
            }
        }
        fence_proxy_async_smem();
    }
    __syncthreads();
    if (tid == 0 && total && nfr) {
        tma_store_1d(out + f0 * FRAME, vout, (uint32_t)nfr * FRAME);
        tma_store_commit();
    }
    if (tileno == (long long)gridDim.x - 1) {
        const uint32_t tail = (uint32_t)(n % FRAME);
        if (tail && nframes == full_frames + 1) {
            const uint64_t off = (sc.base + sc.total) * 16;
            const uint32_t b = bits[full_frames];
            const uint64_t avail = nvalues > off ? nvalues - off : 0;
            if (b == 0 || b > 8 || ((uint64_t)tail * b + 7) / 8 > avail) {
                if (tid == 0) atomicExch(err, 3);
            } else {
                decode_tail_frame(values + off, (uint32_t)(avail > 128 ? 128 : avail), tail, b, out + full_frames * FRAME);
            }
        }
    }
    if (tid == 0) tma_store_wait_read<0>();
}

// =============================================================================================
//                                        launchers
// =============================================================================================
size_t fl_tiles(size_t n)
{
    const size_t frames = (n + FRAME - 1) / FRAME;
    return (frames + TILE_FRAMES - 1) / TILE_FRAMES;
}
size_t fl_workspace_bytes(size_t n) { return (fl_tiles(n) + 1) * sizeof(uint64_t) + 16; }

const int* fl_decode_status(const void* d_ws, size_t n)
{
    return reinterpret_cast<const int*>(reinterpret_cast<const uint64_t*>(d_ws) + fl_tiles(n));
}

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
    cudaError_t e = cudaMemsetAsync(desc, 0, tiles * sizeof(uint64_t), stream);
    if (e != cudaSuccess) return e;
    fl_encode_kernel<<<(unsigned)tiles, THREADS, TILE_BYTES, stream>>>(d_in, n, d_bits, d_values, d_values_size, desc);
    return cudaGetLastError();
}

cudaError_t fl_decode_launch(const uint8_t* d_bits, size_t nframes, const uint8_t* d_values, size_t nvalues, uint8_t* d_out,
                             size_t n, void* d_ws, cudaStream_t stream)
{
    const size_t tiles = fl_tiles(n);
    uint64_t* desc = reinterpret_cast<uint64_t*>(d_ws);
    int* err = reinterpret_cast<int*>(desc + tiles);
    cudaError_t e = cudaMemsetAsync(desc, 0, (tiles + 1) * sizeof(uint64_t), stream);
    if (e != cudaSuccess) return e;
    constexpr int smem = 2 * TILE_BYTES + 128;
    static bool done[64];
    e = set_smem_once((const void*)fl_decode_kernel, smem, done);
    if (e != cudaSuccess) return e;
    fl_decode_kernel<<<(unsigned)tiles, THREADS, smem, stream>>>(d_bits, nframes, d_values, nvalues, d_out, n, desc, err);
    return cudaGetLastError();
}

}  // namespace flrl
