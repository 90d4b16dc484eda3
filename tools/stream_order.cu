// stream_order.cu — does the ORDER in which persistent CTAs walk memory matter for DRAM throughput?
// 296 / 592 / 1184 CTAs x 256 threads copy `n` bytes in 16 KiB tiles (LDG.128 / STG.128, a warp moves 512 contiguous bytes per
// instruction).  mode 0: CTA c owns a contiguous range of tiles (as fl_decode does); mode 1: tile = k * G + c (all
// CTAs inside one moving window, as the kernels with an atomic tile counter do).  ratio: bytes read per byte written
// The mixes (percent of every tile read / written) are those of the four codec kernels.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o stream_order tools/stream_order.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256, 4) walk(const uint4* __restrict__ in, uint4* __restrict__ out, long long tiles, int mode, int ratio, int wratio)
{
    const long long G = gridDim.x, c = blockIdx.x;
    const long long per = (tiles + G - 1) / G;
    for (long long k = 0; k < per; ++k) {
        const long long tile = mode == 0 ? c * per + k : k * G + c;
        if (tile >= tiles) break;
        const uint4* src = in + tile * 1024;  // 16 KiB = 1024 uint4
        uint4* dst = out + tile * 1024;
        uint4 v[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int idx = i * 256 + threadIdx.x;
            v[i] = make_uint4(idx, tile, k, c);
            if ((idx * 100 >> 10) < ratio) v[i] = src[idx];  // the first ratio % of the tile are read
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int idx = i * 256 + threadIdx.x;
            if ((idx * 100 >> 10) < wratio) dst[idx] = v[i];  // the first wratio % of the tile are written
            else if (v[i].x == 0x12345679u) dst[idx] = v[i];  // (keeps the loads alive)
        }
    }
}

int main()
{
    const size_t n = 1ull << 30;
    uint4 *in, *out;
    cudaMalloc(&in, n);
    cudaMalloc(&out, n);
    cudaMemset(in, 1, n);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    // read % / written % of every 16 KiB tile: copy, FL decode, write only (RL decode, long runs), FL encode, read only (RL encode, long runs)
    const int mixes[][2] = {{100, 100}, {60, 100}, {0, 100}, {100, 52}, {100, 3}};
    for (int grid : {296, 592, 1184})
    for (auto& mix : mixes)
        for (int mode = 0; mode < 2; ++mode) {
            const int ratio = mix[0], wratio = mix[1];
            float best = 1e9f;
            for (int it = 0; it < 8; ++it) {
                cudaEventRecord(a);
                walk<<<grid, 256>>>(in, out, (long long)(n / 16384), mode, ratio, wratio);
                cudaEventRecord(b);
                cudaEventSynchronize(b);
                float ms;
                cudaEventElapsedTime(&ms, a, b);
                if (it >= 2 && ms < best) best = ms;
            }
            const double bytes = (double)n * (ratio / 100.0 + wratio / 100.0);
            printf("%4d CTAs  read %3d %% write %3d %%  %s  %.3f ms  %.0f GB/s of DRAM traffic\n", grid, ratio, wratio, mode == 0 ? "contiguous ranges" : "moving window    ", best,
                   bytes / best / 1e6);
        }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
