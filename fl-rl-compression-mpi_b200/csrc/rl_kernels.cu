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

// high bit of each byte of the result is set iff that byte of x is non-zero
__device__ __forceinline__ uint32_t nz_bytes(uint32_t x)
{
    return (((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x) & 0x80808080u;
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
size_t rl_decode_tiles(size_t pairs) { return (pairs + RLD_PAIRS - 1) / RLD_PAIRS; }
size_t rl_decode_workspace_bytes(size_t pairs) { return (rl_decode_tiles(pairs) + 2) * sizeof(uint64_t); }
const int* rl_decode_status(const void* d_ws, size_t pairs)
{
    return reinterpret_cast<const int*>(reinterpret_cast<const uint64_t*>(d_ws) + rl_decode_tiles(pairs));
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
