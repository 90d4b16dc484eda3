// rl_kernels.cu — small reductions the host layer needs before it can cut a compressed stream into shards and
// chunks: the sum of a byte array (shard totals) and the sums of its consecutive segments (chunk totals), i.e. 16 x the
// value bytes behind FL widths, or the output bytes behind RL counts.  The codec kernels themselves are in
// rl_encode.cu / rl_decode.cu / fl_encode.cu / fl_decode.cu.
#include "rl_kernels.cuh"

#include <algorithm>

#include "flrl_bits.cuh"

namespace flrl {

// high bit of each byte of the result is set iff that byte of x is non-zero
__device__ __forceinline__ uint32_t nz_bytes(uint32_t x)
{
    return (((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x) & 0x80808080u;
}

static __device__ __forceinline__ uint32_t byte_sum4(uint32_t w)
{
    const uint32_t h = (w & 0x00ff00ffu) + ((w >> 8) & 0x00ff00ffu);
    return (h & 0xffffu) + (h >> 16);
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

// ---- sums of consecutive segments of a byte array + what the containers promise about those bytes ------------
// sums[i] = sum of p[i*seg .. min((i+1)*seg, n)); sums[nseg] |= 1 if a byte is 0, |= 2 if a byte is > 8.
// blockIdx.y = segment, blockIdx.x = part of the segment.
__global__ void __launch_bounds__(256) segment_sums_kernel(const uint8_t* __restrict__ p, size_t n, size_t seg, unsigned long long* __restrict__ sums,
                                                           unsigned nseg)
{
    const size_t s0 = (size_t)blockIdx.y * seg, s1 = s0 + seg < n ? s0 + seg : n;
    const size_t n16 = (s1 - s0) / 16;
    const uint4* q = reinterpret_cast<const uint4*>(p + s0);
    unsigned long long acc = 0;
    uint32_t zero = 0, big = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 v = q[i];
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            acc += byte_sum4(w[k]);
            zero |= nz_bytes(w[k]) ^ 0x80808080u;
            big |= (((w[k] & 0x7f7f7f7fu) + 0x77777777u) | w[k]) & 0x80808080u;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x < ((s1 - s0) & 15)) {
        const uint8_t b = p[s0 + n16 * 16 + threadIdx.x];
        acc += b;
        zero |= b == 0;
        big |= b > 8;
    }
    uint32_t f = (zero ? 1u : 0u) | (big ? 2u : 0u);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        acc += __shfl_xor_sync(0xffffffffu, acc, d);
        f |= __shfl_xor_sync(0xffffffffu, f, d);
    }
    if ((threadIdx.x & 31) == 0) {
        if (acc) atomicAdd(sums + blockIdx.y, acc);
        if (f) atomicOr(sums + nseg, (unsigned long long)f);
    }
}

cudaError_t segment_sums_launch(const uint8_t* d_p, size_t n, size_t seg, uint64_t* d_sums, cudaStream_t stream)
{
    const size_t nseg = n ? (n + seg - 1) / seg : 0;
    cudaError_t e = cudaMemsetAsync(d_sums, 0, (nseg + 1) * sizeof(uint64_t), stream);
    if (e != cudaSuccess || n == 0) return e;
    int dev = 0, sms = 0;
    if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
    if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
    const size_t per = std::min<size_t>(seg, n);
    const unsigned parts = (unsigned)std::max<size_t>(1, std::min<size_t>((per / 16 + 2047) / 2048, (size_t)(8 * sms + nseg - 1) / nseg));
    segment_sums_kernel<<<dim3(parts, (unsigned)nseg), 256, 0, stream>>>(d_p, n, seg, reinterpret_cast<unsigned long long*>(d_sums), (unsigned)nseg);
    return cudaGetLastError();
}

cudaError_t sum_bytes_launch(const uint8_t* d_p, size_t n, uint64_t* d_sum, cudaStream_t stream)
{
    cudaError_t e = cudaMemsetAsync(d_sum, 0, sizeof(uint64_t), stream);
    if (e != cudaSuccess || n == 0) return e;
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const unsigned blocks = (unsigned)std::min<size_t>((n / 16 + 255) / 256 + 1, (size_t)sms * 8);
    sum_bytes_kernel<<<blocks, 256, 0, stream>>>(d_p, n, reinterpret_cast<unsigned long long*>(d_sum));
    return cudaGetLastError();
}

}  // namespace flrl
