// pcie_ceiling.cu — what the box gives a pinned-host <-> device copy: H2D alone, D2H alone, both at once, on 1 / 2 / 4 / 8
// GPUs concurrently (one thread and two streams per GPU, 256 MiB per copy).  The yardstick for bench.py's end-to-end
// numbers (host buffers in, host buffers out).  build: nvcc -O2 -o gpurun_out/pcie_ceiling tools/pcie_ceiling.cu
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <thread>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s\n", cudaGetErrorString(e), #x); return 1; } } while (0)

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main()
{
    int have = 0;
    CK(cudaGetDeviceCount(&have));
    const size_t N = 256u << 20;
    const int reps = 8;
    std::vector<void*> hin(have), hout(have), din(have), dout(have);
    std::vector<cudaStream_t> s0(have), s1(have);
    for (int g = 0; g < have; ++g) {
        CK(cudaSetDevice(g));
        CK(cudaHostAlloc(&hin[g], N, cudaHostAllocPortable));
        CK(cudaHostAlloc(&hout[g], N, cudaHostAllocPortable));
        CK(cudaMalloc(&din[g], N));
        CK(cudaMalloc(&dout[g], N));
        CK(cudaStreamCreate(&s0[g]));
        CK(cudaStreamCreate(&s1[g]));
    }
    printf("{\"pcie_ceiling\": [");
    bool first = true;
    for (int G = 1; G <= have; G *= 2) {
        for (int mode = 0; mode < 3; ++mode) {  // 0 H2D, 1 D2H, 2 both
            auto run = [&](int g) {
                cudaSetDevice(g);
                for (int r = 0; r < reps; ++r) {
                    if (mode != 1) cudaMemcpyAsync(din[g], hin[g], N, cudaMemcpyHostToDevice, s0[g]);
                    if (mode != 0) cudaMemcpyAsync(hout[g], dout[g], N, cudaMemcpyDeviceToHost, s1[g]);
                }
                cudaStreamSynchronize(s0[g]);
                cudaStreamSynchronize(s1[g]);
            };
            for (int g = 0; g < G; ++g) run(g);  // warm-up
            const double t0 = now();
            std::vector<std::thread> th;
            for (int g = 0; g < G; ++g) th.emplace_back(run, g);
            for (auto& t : th) t.join();
            const double dt = now() - t0;
            const double bytes = (double)N * reps * G * (mode == 2 ? 2 : 1);
            printf("%s{\"gpus\": %d, \"mode\": \"%s\", \"GBps_total\": %.1f, \"GBps_per_gpu_per_direction\": %.1f}", first ? "" : ", ", G,
                   mode == 0 ? "h2d" : mode == 1 ? "d2h" : "h2d+d2h", bytes / dt / 1e9, (double)N * reps / dt / 1e9);
            first = false;
        }
    }
    printf("]}\n");
    return 0;
}
