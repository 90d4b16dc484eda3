// timers.cuh — progress lines in the reference's format.
//   CpuTimer   "[TIMER] Step: \"%s\", Time: %ld ms"  (src/timers/cpu_timer.cu:7-27)
//   device     "[TIMER] Step: \"%s\", Time: %f ms"   (src/timers/gpu_timer.cu:25) — the milliseconds come from the
//              cudaEvent pairs the host layer records around its kernels / whole call (flrl_last_timing).
#pragma once
#include <chrono>
#include <cstdio>

namespace Timers {

class CpuTimer {
  public:
    void start() { t0_ = std::chrono::high_resolution_clock::now(); }
    void end()
    {
        ms_ = (long)std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::high_resolution_clock::now() - t0_).count();
    }
    long ms() const { return ms_; }
    void printResult(const char* step) const { printf("[TIMER] Step: \"%s\", Time: %ld ms\n", step, ms_); }

  private:
    std::chrono::high_resolution_clock::time_point t0_{};
    long ms_ = 0;
};

inline void printDeviceResult(const char* step, float ms) { printf("[TIMER] Step: \"%s\", Time: %f ms\n", step, ms); }

// bytes / ms -> B, KB, MB or GB per second (src/timers/cpu_timer_with_transfer.cu:16-40)
inline void printTransfer(const char* step, double bytes, double ms)
{
    static const char* units[] = {"B/s", "KB/s", "MB/s", "GB/s"};
    double v = ms > 0 ? bytes / (ms / 1000.0) : 0.0;
    int u = 0;
    while (v >= 1024.0 && u < 3) {
        v /= 1024.0;
        ++u;
    }
    printf("[TIMER] Step: \"%s\", Transfer Speed: %lf [%s] \n", step, v, units[u]);
}

}  // namespace Timers
