// flrl_host.cu — HOST layer of include/flrl.h: host buffers in, host buffers out.
//
// Mirrors the reference's host drivers: FixedLength::gpuCompress / gpuDecompress
// (/root/reference/src/fl/fl_gpu.cu:289-423, :537-645) and, with ngpus > 1, gpuNCCLCompress
// (:76-287) — re-designed for one process driving all GPUs of a B200 box:
//   * contiguous 128-byte aligned shards (the loadFileMpi rule, file_io.cu:46-51, in 64-bit arithmetic), one
//     worker thread per GPU, no data-path collective: only the per-shard compressed sizes are exchanged — ONE
//     ncclAllGather of a u64 per GPU — and every GPU then delivers its payload at its exclusive-scan offset
//     (directly to the host buffer, or peer-copied into GPU 0's buffer first: flrl_set_concat_mode);
//   * every shard runs a CHUNK PIPELINE when the caller's buffers are pinned (flrl_host_alloc): copy-in, kernel
//     and copy-out of 64 MiB chunks overlap on three streams, so a GPU's PCIe link is busy in both directions;
//     chunk outputs concatenate to the single-stream output (FL: 128-byte aligned cuts; RL: prev_run + 255
//     bytes of look-ahead).  Pageable buffers take the whole-shard path through two pinned bounce buffers;
//   * what the decoders need before the first copy (value bytes per chunk = 16 x sum of its widths, output
//     bytes per shard = sum of its counts) is computed on the GPU (segment_sums_kernel), not by a host loop;
//   * device buffers, streams, events and pinned staging are cached per device (the reference cudaMallocs /
//     cudaFrees every buffer on every call, fl_gpu.cu:320-322, :409-412); flrl_release() frees all of it;
//   * timing is cudaEvent based (the reference's GpuTimer, src/timers/gpu_timer.cu:6-26), per GPU;
//     NVTX ranges mark copy-in / kernel / copy-out of every chunk for nsys.
// No CPU fallback: every failure is reported, never routed to the CPU codec.
// One host-layer call at a time per process (the per-device caches are shared): calls are serialised by a mutex.
#include <dlfcn.h>
#include <nvtx3/nvToolsExt.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <mutex>
#include <thread>
#include <vector>

#include "abi_common.cuh"
#include "fl_kernels.cuh"
#include "rl_kernels.cuh"

using namespace flrl;

namespace {

constexpr size_t BOUNCE = 32u << 20;  // pinned staging chunk for pageable host buffers
constexpr int MAX_DEV = 16;
constexpr size_t FRAME_LENGTH_ = FLRL_FRAME_LENGTH;

#define CK(call)                                      \
    do {                                              \
        cudaError_t e_ = (call);                      \
        if (e_ != cudaSuccess) return fail_cuda(e_, #call); \
    } while (0)

struct Nvtx {  // scoped NVTX range
    explicit Nvtx(const char* name) { nvtxRangePushA(name); }
    ~Nvtx() { nvtxRangePop(); }
};

struct DBuf {
    uint8_t* p = nullptr;
    size_t cap = 0;
    int ensure(size_t n)
    {
        n = (n + 255) & ~(size_t)255;
        if (n <= cap) return 0;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        CK(cudaMalloc(&p, n));
        cap = n;
        return 0;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

struct Dev {
    int id = -1;
    bool ready = false;
    cudaStream_t s = nullptr;
    cudaEvent_t ev[2] = {nullptr, nullptr}, t0 = nullptr, t1 = nullptr, k0 = nullptr, k1 = nullptr;
    uint8_t* bounce[2] = {nullptr, nullptr};
    uint64_t* h_scal = nullptr;  // pinned: small results read back from the device (64 words)
    uint64_t* h_sums = nullptr;  // pinned: per-chunk sums read back from the device
    size_t h_sums_cap = 0, n_sums = 0;
    DBuf in, a, b, ws, scal, gather, sums;  // device buffers (grow-only)
    // chunk pipeline: copy-in and copy-out streams beside the kernel stream `s`, two slots of device buffers,
    // per-slot events
    cudaStream_t s_in = nullptr, s_out = nullptr;
    cudaEvent_t in_done[2] = {nullptr, nullptr}, k_start[2] = {nullptr, nullptr}, k_done[2] = {nullptr, nullptr},
                out_done[2] = {nullptr, nullptr};
    DBuf pin[2], pa[2], pb[2];
    bool pipe_ready = false;
    // the last call's figures for this GPU
    float kernel_ms = 0, total_ms = 0;
    uint64_t bytes_in = 0, bytes_out = 0;  // host -> device, device -> host

    int init(int dev)
    {
        if (ready) return 0;
        id = dev;
        CK(cudaSetDevice(dev));
        CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
        for (auto* e : {&ev[0], &ev[1]}) CK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
        for (auto* e : {&t0, &t1, &k0, &k1}) CK(cudaEventCreate(e));
        CK(cudaHostAlloc(&h_scal, 64 * sizeof(uint64_t), cudaHostAllocPortable));
        if (scal.ensure(64 * sizeof(uint64_t))) return -1;
        ready = true;
        return 0;
    }
    int need_pipe()
    {
        if (pipe_ready) return 0;
        CK(cudaStreamCreateWithFlags(&s_in, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) {
            CK(cudaEventCreateWithFlags(&in_done[i], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&out_done[i], cudaEventDisableTiming));
            CK(cudaEventCreate(&k_start[i]));
            CK(cudaEventCreate(&k_done[i]));
        }
        pipe_ready = true;
        return 0;
    }
    int need_bounce()
    {
        for (int i = 0; i < 2; ++i)
            if (!bounce[i]) CK(cudaHostAlloc(&bounce[i], BOUNCE, cudaHostAllocPortable));
        return 0;
    }
    int need_sums(size_t n)  // n u64 on the device and pinned on the host
    {
        if (sums.ensure(n * sizeof(uint64_t))) return -1;
        if (n > h_sums_cap) {
            if (h_sums) cudaFreeHost(h_sums);
            h_sums = nullptr;
            h_sums_cap = 0;
            CK(cudaHostAlloc(&h_sums, n * sizeof(uint64_t), cudaHostAllocPortable));
            h_sums_cap = n;
        }
        return 0;
    }
    void release()
    {
        if (!ready) return;
        cudaSetDevice(id);
        cudaStreamSynchronize(s);
        for (DBuf* x : {&in, &a, &b, &ws, &scal, &gather, &sums, &pin[0], &pin[1], &pa[0], &pa[1], &pb[0], &pb[1]}) x->release();
        for (auto& p : bounce) {
            if (p) cudaFreeHost(p);
            p = nullptr;
        }
        if (h_scal) cudaFreeHost(h_scal);
        if (h_sums) cudaFreeHost(h_sums);
        h_scal = h_sums = nullptr;
        h_sums_cap = 0;
        for (cudaEvent_t e : {ev[0], ev[1], t0, t1, k0, k1}) cudaEventDestroy(e);
        cudaStreamDestroy(s);
        if (pipe_ready) {
            cudaStreamSynchronize(s_in);
            cudaStreamSynchronize(s_out);
            for (int i = 0; i < 2; ++i)
                for (cudaEvent_t e : {in_done[i], out_done[i], k_start[i], k_done[i]}) cudaEventDestroy(e);
            cudaStreamDestroy(s_in);
            cudaStreamDestroy(s_out);
            pipe_ready = false;
        }
        ready = false;
    }
};

Dev g_dev[MAX_DEV];
int g_map[MAX_DEV];  // shard index -> device ordinal for the running call (guarded by g_host_mu)
int g_last_gpus = 0;  // GPUs of the last call (flrl_last_timing_gpu)
int g_concat_peer = 0;
std::atomic<size_t> g_pipe_chunk{(size_t)64u << 20};
std::mutex g_host_mu;  // host-layer calls share the per-device caches: one call at a time
thread_local float tl_kernel_ms = 0, tl_total_ms = 0;

bool is_pinned(const void* p)
{
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeHost;
}

// host -> device on d.s; pageable sources are staged through the two pinned bounce buffers
int h2d(Dev& d, void* dst, const void* src, size_t n)
{
    if (n == 0) return 0;
    d.bytes_in += n;
    if (is_pinned(src)) {
        CK(cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, d.s));
        return 0;
    }
    if (d.need_bounce()) return -1;
    int slot = 0;
    for (size_t o = 0; o < n; o += BOUNCE, slot ^= 1) {
        const size_t m = std::min(BOUNCE, n - o);
        CK(cudaEventSynchronize(d.ev[slot]));  // the copy that last used this buffer has drained
        memcpy(d.bounce[slot], static_cast<const uint8_t*>(src) + o, m);
        CK(cudaMemcpyAsync(static_cast<uint8_t*>(dst) + o, d.bounce[slot], m, cudaMemcpyHostToDevice, d.s));
        CK(cudaEventRecord(d.ev[slot], d.s));
    }
    return 0;
}

// device -> host on d.s; returns after the data is in dst when dst is pageable, else asynchronous
int d2h(Dev& d, void* dst, const void* src, size_t n)
{
    if (n == 0) return 0;
    d.bytes_out += n;
    if (is_pinned(dst)) {
        CK(cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, d.s));
        return 0;
    }
    if (d.need_bounce()) return -1;
    size_t prev_o = 0, prev_m = 0;
    int slot = 0;
    for (size_t o = 0; o < n; o += BOUNCE, slot ^= 1) {
        const size_t m = std::min(BOUNCE, n - o);
        CK(cudaMemcpyAsync(d.bounce[slot], static_cast<const uint8_t*>(src) + o, m, cudaMemcpyDeviceToHost, d.s));
        CK(cudaEventRecord(d.ev[slot], d.s));
        if (prev_m) {  // drain the previous chunk while this one is in flight
            CK(cudaEventSynchronize(d.ev[slot ^ 1]));
            memcpy(static_cast<uint8_t*>(dst) + prev_o, d.bounce[slot ^ 1], prev_m);
        }
        prev_o = o;
        prev_m = m;
    }
    CK(cudaEventSynchronize(d.ev[slot ^ 1]));
    memcpy(static_cast<uint8_t*>(dst) + prev_o, d.bounce[slot ^ 1], prev_m);
    return 0;
}

// Restores the caller's current device when a host-layer call returns.
struct DeviceGuard {
    int dev = -1;
    DeviceGuard() { if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); dev = -1; } }
    ~DeviceGuard() { if (dev >= 0) cudaSetDevice(dev); }
};

// ---------------------------------------------------------------------------------------------
// Chunk pipeline of one shard (file_io's "pinned, double-buffered cudaMemcpyAsync"): the shard is cut into
// chunks that encode / decode independently, and chunk i+1 is copied in while chunk i runs and chunk i-1 is
// copied out — PCIe is full duplex, so a shard costs max(bytes in, bytes out) instead of their sum.
//   copy_in(i, slot)  enqueues the H2D copies of chunk i on d.s_in
//   launch(i, slot)   enqueues the kernel (+ the read-back of its 8-byte result / status into d.h_scal) on d.s
//   copy_out(i, slot) is called once the kernel of chunk i has finished (its result is in d.h_scal): enqueues the
//                     copies of its output on d.s_out
// Returns with all three streams drained.  chunk size: 64 MiB (about 1.1 ms of PCIe 5 x16 each way);
// flrl_set_pipe_chunk overrides it (tests use small chunks).
size_t pipe_chunk_bytes()
{
    size_t c = g_pipe_chunk.load() & ~(size_t)4095;
    return c < 65536 ? 65536 : c;
}
inline uint64_t* pipe_result(Dev& d, int slot) { return d.h_scal + 48 + 4 * slot; }              // pinned host
inline uint64_t* pipe_result_dev(Dev& d, int slot) { return reinterpret_cast<uint64_t*>(d.scal.p) + 48 + 4 * slot; }

template <class In, class Launch, class Out>
int run_pipeline(Dev& d, size_t nchunks, In copy_in, Launch launch, Out copy_out)
{
    if (d.need_pipe()) return -1;
    if (nchunks == 0) return 0;
    float kernel_ms = 0;
    CK(cudaEventRecord(d.ev[0], d.s));
    CK(cudaStreamWaitEvent(d.s_in, d.ev[0], 0));  // what the caller put on d.s comes first
    auto finish = [&](size_t i) -> int {
        const int slot = (int)(i & 1);
        CK(cudaEventSynchronize(d.k_done[slot]));
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, d.k_start[slot], d.k_done[slot]));
        kernel_ms += ms;
        {
            Nvtx r("flrl copy-out");
            if (int rc = copy_out(i, slot)) return rc;
        }
        CK(cudaEventRecord(d.out_done[slot], d.s_out));
        return 0;
    };
    for (size_t i = 0; i < nchunks; ++i) {
        const int slot = (int)(i & 1);
        if (i >= 2) {
            CK(cudaStreamWaitEvent(d.s_in, d.k_done[slot], 0));   // the kernel of chunk i-2 has read this slot's input
            CK(cudaStreamWaitEvent(d.s, d.out_done[slot], 0));    // chunk i-2's output has left this slot's buffers
        }
        {
            Nvtx r("flrl copy-in");
            if (int rc = copy_in(i, slot)) return rc;
        }
        CK(cudaEventRecord(d.in_done[slot], d.s_in));
        CK(cudaStreamWaitEvent(d.s, d.in_done[slot], 0));
        CK(cudaEventRecord(d.k_start[slot], d.s));
        {
            Nvtx r("flrl kernel");
            if (int rc = launch(i, slot)) return rc;
        }
        CK(cudaEventRecord(d.k_done[slot], d.s));
        CK(cudaStreamWaitEvent(d.s_out, d.k_done[slot], 0));
        if (i >= 1)
            if (int rc = finish(i - 1)) return rc;
    }
    if (int rc = finish(nchunks - 1)) return rc;
    CK(cudaStreamWaitEvent(d.s, d.out_done[(nchunks - 1) & 1], 0));
    if (nchunks >= 2) CK(cudaStreamWaitEvent(d.s, d.out_done[(nchunks - 2) & 1], 0));
    CK(cudaStreamSynchronize(d.s));
    d.kernel_ms += kernel_ms;
    return 0;
}

// Number of GPUs the call uses and their ordinals: one GPU = the caller's CURRENT device (so a rank of a
// multi-process job stays on its own GPU); several = devices 0 .. G-1.
int device_count(int ngpus, int* out)
{
    int have = 0;
    CK(cudaGetDeviceCount(&have));
    if (have <= 0) return fail((int)cudaErrorNoDevice, "CUDA Error: no CUDA device (the fl / rl methods have no CPU fallback)");
    int g = ngpus <= 0 ? have : std::min(ngpus, have);
    g = std::min(g, MAX_DEV);
    int cur = 0;
    CK(cudaGetDevice(&cur));
    for (int i = 0; i < MAX_DEV; ++i) g_map[i] = i;
    if (g == 1) g_map[0] = cur < MAX_DEV ? cur : 0;
    *out = g;
    return 0;
}
void single_gpu_map()
{
    int cur = 0;
    if (cudaGetDevice(&cur) != cudaSuccess) cur = 0;
    g_map[0] = cur < MAX_DEV ? cur : 0;
}

// contiguous 128-byte aligned shards: per = (n / (128 * G)) * 128, the last shard takes the rest
// (/root/reference/src/file_io.cu:46-51, in 64-bit arithmetic)
void shard_range(size_t n, int G, int g, size_t unit, size_t* begin, size_t* len)
{
    const size_t per = (n / (unit * (size_t)G)) * unit;
    *begin = per * (size_t)g;
    *len = g == G - 1 ? n - *begin : per;
}

// Number of bytes equal to data[pos] immediately before pos, looking back at most `limit` bytes: whole 4 KiB
// blocks with memcmp against a block of that byte, then byte by byte.
uint64_t run_before(const uint8_t* data, size_t pos, size_t limit)
{
    if (pos == 0 || limit == 0) return 0;
    limit = std::min(limit, pos);
    const uint8_t v = data[pos];
    uint64_t r = 0;
    if (data[pos - 1] != v) return 0;
    uint8_t block[4096];
    memset(block, v, sizeof block);
    while (limit - r >= sizeof block && memcmp(data + pos - r - sizeof block, block, sizeof block) == 0) r += sizeof block;
    while (r < limit && data[pos - 1 - r] == v) ++r;
    return r;
}

// ---- NCCL, resolved lazily: libflrl.so itself does not depend on libnccl -------------------------
struct Nccl {
    void* h = nullptr;
    void* comms[MAX_DEV + 1][MAX_DEV] = {};  // one communicator set per GPU count
    bool made[MAX_DEV + 1] = {};
    int (*CommInitAll)(void**, int, const int*) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    std::mutex mu;
} g_nccl;

int nccl_ensure(int G)
{
    std::lock_guard<std::mutex> lk(g_nccl.mu);
    if (g_nccl.made[G]) return 0;
    if (!g_nccl.h) {
        g_nccl.h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!g_nccl.h) return fail(FLRL_E_ARG, "NCCL Error: cannot load libnccl.so.2: %s", dlerror());
        g_nccl.CommInitAll = reinterpret_cast<decltype(g_nccl.CommInitAll)>(dlsym(g_nccl.h, "ncclCommInitAll"));
        g_nccl.AllGather = reinterpret_cast<decltype(g_nccl.AllGather)>(dlsym(g_nccl.h, "ncclAllGather"));
        g_nccl.GroupStart = reinterpret_cast<decltype(g_nccl.GroupStart)>(dlsym(g_nccl.h, "ncclGroupStart"));
        g_nccl.GroupEnd = reinterpret_cast<decltype(g_nccl.GroupEnd)>(dlsym(g_nccl.h, "ncclGroupEnd"));
        g_nccl.GetErrorString = reinterpret_cast<decltype(g_nccl.GetErrorString)>(dlsym(g_nccl.h, "ncclGetErrorString"));
        if (!g_nccl.CommInitAll || !g_nccl.AllGather || !g_nccl.GroupStart || !g_nccl.GroupEnd)
            return fail(FLRL_E_ARG, "NCCL Error: libnccl.so.2 lacks a required symbol");
    }
    int devs[MAX_DEV];
    for (int i = 0; i < G; ++i) devs[i] = i;
    const int rc = g_nccl.CommInitAll(g_nccl.comms[G], G, devs);
    if (rc != 0) return fail(FLRL_E_ARG, "NCCL Error: ncclCommInitAll: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?");
    g_nccl.made[G] = true;
    return 0;
}

// one u64 per GPU (device memory at d.scal + slot) -> all G values in every GPU's d.scal + 8 .. + 8 + G
int nccl_allgather_sizes(int G)
{
    constexpr int ncclUint64_ = 5;  // ncclDataType_t::ncclUint64
    int rc = g_nccl.GroupStart();
    for (int g = 0; g < G && rc == 0; ++g) {
        uint64_t* base = reinterpret_cast<uint64_t*>(g_dev[g_map[g]].scal.p);
        rc = g_nccl.AllGather(base, base + 8, 1, ncclUint64_, g_nccl.comms[G][g], g_dev[g_map[g]].s);
    }
    const int rc2 = g_nccl.GroupEnd();
    if (rc == 0) rc = rc2;
    if (rc != 0) return fail(FLRL_E_ARG, "NCCL Error: ncclAllGather: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?");
    return 0;
}

// run fn(g) for every shard: inline for one GPU, one worker thread per GPU otherwise
template <typename F>
int for_each_gpu(int G, F&& fn)
{
    if (G == 1) return fn(0);
    std::vector<int> rc(G, 0);
    std::vector<std::string> msg(G);
    std::vector<std::thread> th;
    for (int g = 0; g < G; ++g)
        th.emplace_back([&, g] {
            rc[g] = fn(g);
            if (rc[g]) msg[g] = last_error_slot();  // thread-local: carry it back to the caller's thread
        });
    for (auto& t : th) t.join();
    for (int g = 0; g < G; ++g)
        if (rc[g]) {
            last_error_slot() = msg[g];
            return rc[g];
        }
    return 0;
}

void collect_timing(int G)
{
    tl_kernel_ms = tl_total_ms = 0;
    for (int g = 0; g < G; ++g) {
        tl_kernel_ms = std::max(tl_kernel_ms, g_dev[g_map[g]].kernel_ms);
        tl_total_ms = std::max(tl_total_ms, g_dev[g_map[g]].total_ms);
    }
    g_last_gpus = G;
}

// a call starts on a GPU: device current, caches ready, figures reset, t0 recorded
int begin_gpu(int g, Dev** out)
{
    Dev& d = g_dev[g_map[g]];
    CK(cudaSetDevice(g_map[g]));
    if (d.init(g_map[g])) return -1;
    d.kernel_ms = d.total_ms = 0;
    d.bytes_in = d.bytes_out = 0;
    CK(cudaEventRecord(d.t0, d.s));
    *out = &d;
    return 0;
}
// ... and ends: everything on d.s has finished, total time taken
int end_gpu(Dev& d)
{
    CK(cudaSetDevice(d.id));
    CK(cudaEventRecord(d.t1, d.s));
    CK(cudaStreamSynchronize(d.s));
    CK(cudaEventElapsedTime(&d.total_ms, d.t0, d.t1));
    return 0;
}
// one kernel bracketed by k0 / k1 on d.s (whole-shard paths); call after the stream has drained
int add_kernel_time(Dev& d)
{
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, d.k0, d.k1));
    d.kernel_ms += ms;
    return 0;
}

// After every GPU has put its figure (compressed size, sum of widths, sum of counts) into scal[0] on ITS stream:
// exchange (NCCL all-gather when G > 1), read back.  sizes[g] = per-shard figure.
int exchange_sizes(int G, std::vector<uint64_t>& sizes)
{
    Nvtx r("flrl size all-gather");
    sizes.assign(G, 0);
    if (G == 1) {
        Dev& d = g_dev[g_map[0]];
        CK(cudaSetDevice(g_map[0]));
        CK(cudaMemcpyAsync(d.h_scal, d.scal.p, sizeof(uint64_t), cudaMemcpyDeviceToHost, d.s));
        CK(cudaStreamSynchronize(d.s));
        sizes[0] = d.h_scal[0];
        return 0;
    }
    if (int rc = nccl_ensure(G)) return rc;
    if (int rc = nccl_allgather_sizes(G)) return rc;
    for (int g = 0; g < G; ++g) {
        Dev& d = g_dev[g_map[g]];
        CK(cudaSetDevice(g_map[g]));
        CK(cudaMemcpyAsync(d.h_scal, d.scal.p + 8 * sizeof(uint64_t), G * sizeof(uint64_t), cudaMemcpyDeviceToHost, d.s));
    }
    for (int g = 0; g < G; ++g) {
        CK(cudaSetDevice(g_map[g]));
        CK(cudaStreamSynchronize(g_dev[g_map[g]].s));
    }
    for (int g = 0; g < G; ++g) sizes[g] = g_dev[g_map[0]].h_scal[g];
    for (int g = 1; g < G; ++g)  // every rank must have gathered the same vector
        for (int k = 0; k < G; ++k)
            if (g_dev[g_map[g]].h_scal[k] != sizes[k]) return fail(FLRL_E_ARG, "NCCL Error: size all-gather mismatch on GPU %d", g);
    return 0;
}
// the figure is known on the host (the pipeline read every chunk's size back): put it where the all-gather reads it
int post_size(Dev& d, uint64_t v)
{
    d.h_scal[40] = v;
    CK(cudaMemcpyAsync(d.scal.p, d.h_scal + 40, sizeof(uint64_t), cudaMemcpyHostToDevice, d.s));
    return 0;
}

// deliver every shard's payload (device pointer src[g], bytes len[g]; nullptr = already delivered) to host
// dst + off[g]: directly from each GPU, or peer-copied into GPU 0's gather buffer first (one D2H from GPU 0)
int deliver(int G, uint8_t* dst, const std::vector<const uint8_t*>& src, const std::vector<uint64_t>& off,
            const std::vector<uint64_t>& len, uint64_t total)
{
    Nvtx r("flrl concatenate");
    if (G > 1 && g_concat_peer) {
        Dev& d0 = g_dev[g_map[0]];
        CK(cudaSetDevice(g_map[0]));
        if (d0.gather.ensure(total)) return -1;
        for (int g = 0; g < G; ++g) {
            if (!len[g] || !src[g]) continue;
            CK(cudaSetDevice(g_map[g]));
            CK(cudaMemcpyPeerAsync(d0.gather.p + off[g], g_map[0], src[g], g_map[g], len[g], g_dev[g_map[g]].s));
        }
        for (int g = 0; g < G; ++g) {
            CK(cudaSetDevice(g_map[g]));
            CK(cudaStreamSynchronize(g_dev[g_map[g]].s));
        }
        CK(cudaSetDevice(g_map[0]));
        if (d2h(d0, dst, d0.gather.p, total)) return -1;
        CK(cudaStreamSynchronize(d0.s));  // the gather buffer is reused by the next delivery
        return 0;
    }
    return for_each_gpu(G, [&](int g) -> int {
        if (!src[g]) return 0;
        CK(cudaSetDevice(g_map[g]));
        return d2h(g_dev[g_map[g]], dst + off[g], src[g], len[g]);
    });
}

// the pipeline applies to inputs of more than one chunk per call and pinned host buffers
bool pipe_ok(size_t bytes, std::initializer_list<const void*> bufs)
{
    if (bytes <= pipe_chunk_bytes()) return false;
    for (const void* b : bufs)
        if (!is_pinned(b)) return false;
    return true;
}

}  // namespace

namespace {

// ---------------------------------------------------------------------------------- shards on the chunk pipeline
// FL compress of one shard.  The widths go straight to the host (their place is known a priori).  The values go
// straight to the host when `values_host` is given (the shard's place in the container is known: shard 0), else they
// are appended to d.b on the device and delivered after the size exchange.
int fl_compress_shard(Dev& d, const uint8_t* data, size_t n, uint8_t* bits_host, uint8_t* values_host, size_t values_cap, uint64_t* total)
{
    const size_t CH = pipe_chunk_bytes();
    const size_t nch = (n + CH - 1) / CH;
    for (int k = 0; k < 2; ++k)
        if (d.pin[k].ensure(CH) || d.pa[k].ensure(CH / 128) || d.pb[k].ensure(CH)) return -1;
    if (d.ws.ensure(fl_workspace_bytes(CH))) return -1;
    if (!values_host && d.b.ensure((n + 15) & ~(size_t)15)) return -1;
    uint64_t voff = 0;
    int rc = run_pipeline(
        d, nch,
        [&](size_t i, int k) -> int {
            const size_t c0 = i * CH, m = std::min(CH, n - c0);
            CK(cudaMemcpyAsync(d.pin[k].p, data + c0, m, cudaMemcpyHostToDevice, d.s_in));
            d.bytes_in += m;
            return 0;
        },
        [&](size_t i, int k) -> int {
            const size_t c0 = i * CH, m = std::min(CH, n - c0);
            CK(fl_encode_launch(d.pin[k].p, m, d.pa[k].p, d.pb[k].p, pipe_result_dev(d, k), d.ws.p, d.s));
            CK(cudaMemcpyAsync(pipe_result(d, k), pipe_result_dev(d, k), sizeof(uint64_t), cudaMemcpyDeviceToHost, d.s));
            return 0;
        },
        [&](size_t i, int k) -> int {
            const size_t c0 = i * CH, m = std::min(CH, n - c0);
            const uint64_t vs = *pipe_result(d, k);
            CK(cudaMemcpyAsync(bits_host + c0 / 128, d.pa[k].p, (m + 127) / 128, cudaMemcpyDeviceToHost, d.s_out));
            d.bytes_out += (m + 127) / 128;
            if (values_host) {
                if (voff + vs > values_cap) return fail(FLRL_E_CAPACITY, "flrl_fl_gpu_compress: values capacity %zu too small", values_cap);
                if (vs) CK(cudaMemcpyAsync(values_host + voff, d.pb[k].p, vs, cudaMemcpyDeviceToHost, d.s_out));
                d.bytes_out += vs;
            } else if (vs) {
                CK(cudaMemcpyAsync(d.b.p + voff, d.pb[k].p, vs, cudaMemcpyDeviceToDevice, d.s_out));
            }
            voff += vs;
            return 0;
        });
    *total = voff;
    return rc;
}

// FL decompress of one shard whose widths are in d.a and whose per-chunk width sums are in d.h_sums (phase A).
// `tail_bytes`: value bytes of the container's partial last frame when it is in this shard.
int fl_decompress_shard(Dev& d, size_t n, const uint8_t* values, uint64_t tail_bytes, uint8_t* out, bool* bad)
{
    const size_t CH = pipe_chunk_bytes();
    const size_t nch = (n + CH - 1) / CH;
    for (int k = 0; k < 2; ++k)
        if (d.pin[k].ensure(CH + 32) || d.pb[k].ensure(CH + 16)) return -1;
    if (d.ws.ensure(fl_workspace_bytes(CH))) return -1;
    std::vector<uint64_t> voff(nch + 1, 0);  // every width was checked to be <= 8: a chunk's values fit CH (+ the partial frame)
    for (size_t i = 0; i < nch; ++i) voff[i + 1] = voff[i] + (i < d.n_sums ? 16 * d.h_sums[i] : 0) + (i == nch - 1 ? tail_bytes : 0);
    return run_pipeline(
        d, nch,
        [&](size_t i, int k) -> int {
            const uint64_t vl = voff[i + 1] - voff[i];
            if (vl) CK(cudaMemcpyAsync(d.pin[k].p, values + voff[i], vl, cudaMemcpyHostToDevice, d.s_in));
            d.bytes_in += vl;
            return 0;
        },
        [&](size_t i, int k) -> int {
            const size_t c0 = i * CH, m = std::min(CH, n - c0);
            CK(fl_decode_launch(d.a.p + c0 / 128, (m + 127) / 128, d.pin[k].p, voff[i + 1] - voff[i], d.pb[k].p, m, d.ws.p, d.s));
            CK(cudaMemcpyAsync(pipe_result(d, k), fl_decode_status(d.ws.p, m), sizeof(int), cudaMemcpyDeviceToHost, d.s));
            return 0;
        },
        [&](size_t i, int k) -> int {
            const size_t c0 = i * CH, m = std::min(CH, n - c0);
            if (*reinterpret_cast<int*>(pipe_result(d, k))) *bad = true;
            CK(cudaMemcpyAsync(out + c0, d.pb[k].p, m, cudaMemcpyDeviceToHost, d.s_out));
            d.bytes_out += m;
            return 0;
        });
}

// RL compress of one shard: data[b0 .. b0 + n) of `size` available bytes (the bytes to encode + the caller's look-ahead),
// stream_prev = bytes equal to data[0] immediately before data[0].  Pairs go straight to the host when
// counts_host is given (shard 0), else they are appended to d.a / d.b and delivered after the size exchange.
int rl_compress_shard(Dev& d, const uint8_t* data, size_t size, uint64_t stream_prev, size_t b0, size_t n, uint8_t* counts_host,
                      uint8_t* values_host, size_t pair_cap, uint64_t* total)
{
    const size_t CH = pipe_chunk_bytes();
    const size_t nch = (n + CH - 1) / CH;
    for (int k = 0; k < 2; ++k)
        if (d.pin[k].ensure(CH + 256 + 16) || d.pa[k].ensure(CH) || d.pb[k].ensure(CH)) return -1;
    if (d.ws.ensure(rl_encode_workspace_bytes(CH))) return -1;
    if (!counts_host && (d.a.ensure((n + 15) & ~(size_t)15) || d.b.ensure((n + 15) & ~(size_t)15))) return -1;
    // the run every chunk continues: bytes equal to its first byte immediately before it (a run that covers the whole
    // previous chunk is carried over, so no chunk looks back further than one chunk — except the shard's first)
    std::vector<uint64_t> prev(nch, 0);
    prev[0] = run_before(data, b0, b0);
    if (prev[0] == b0) prev[0] += stream_prev;  // the run reaches back to the start of the data: it continues the caller's
    for (size_t i = 1; i < nch; ++i) {
        const size_t c0 = b0 + i * CH;
        uint64_t r = run_before(data, c0, CH);
        if (r == CH) r += prev[i - 1];  // the whole previous chunk is part of the run (its first byte equals ours)
        prev[i] = r;
    }
    uint64_t off = 0;
    int rc = run_pipeline(
        d, nch,
        [&](size_t i, int k) -> int {
            const size_t c0 = b0 + i * CH, m = std::min(CH, b0 + n - c0), ahead = std::min<size_t>(255, size - (c0 + m));
            CK(cudaMemcpyAsync(d.pin[k].p, data + c0, m + ahead, cudaMemcpyHostToDevice, d.s_in));
            d.bytes_in += m + ahead;
            return 0;
        },
        [&](size_t i, int k) -> int {
            const size_t c0 = b0 + i * CH, m = std::min(CH, b0 + n - c0), ahead = std::min<size_t>(255, size - (c0 + m));
            CK(rl_encode_launch(d.pin[k].p, m, m + ahead, prev[i], d.pa[k].p, d.pb[k].p, CH, pipe_result_dev(d, k), d.ws.p, d.s));
            CK(cudaMemcpyAsync(pipe_result(d, k), pipe_result_dev(d, k), sizeof(uint64_t), cudaMemcpyDeviceToHost, d.s));
            return 0;
        },
        [&](size_t, int k) -> int {
            const uint64_t np = *pipe_result(d, k);
            if (counts_host) {
                if (off + np > pair_cap) return fail(FLRL_E_CAPACITY, "flrl_rl_gpu_compress: pair capacity %zu too small", pair_cap);
                if (np) {
                    CK(cudaMemcpyAsync(counts_host + off, d.pa[k].p, np, cudaMemcpyDeviceToHost, d.s_out));
                    CK(cudaMemcpyAsync(values_host + off, d.pb[k].p, np, cudaMemcpyDeviceToHost, d.s_out));
                }
                d.bytes_out += 2 * np;
            } else if (np) {
                CK(cudaMemcpyAsync(d.a.p + off, d.pa[k].p, np, cudaMemcpyDeviceToDevice, d.s_out));
                CK(cudaMemcpyAsync(d.b.p + off, d.pb[k].p, np, cudaMemcpyDeviceToDevice, d.s_out));
            }
            off += np;
            return 0;
        });
    *total = off;
    return rc;
}

// RL decompress of one shard whose counts are in d.a and whose per-chunk output sizes are in d.h_sums (phase A):
// chunks of `cp` pairs; the values come in chunk by chunk, the output leaves chunk by chunk.
int rl_decompress_shard(Dev& d, size_t np, size_t cp, const uint8_t* values, uint8_t* out, bool* bad)
{
    const size_t nch = (np + cp - 1) / cp;
    uint64_t omax = 0;
    std::vector<uint64_t> ooff(nch + 1, 0);
    for (size_t i = 0; i < nch; ++i) {
        ooff[i + 1] = ooff[i] + d.h_sums[i];
        omax = std::max<uint64_t>(omax, d.h_sums[i]);
    }
    for (int k = 0; k < 2; ++k)
        if (d.pb[k].ensure(cp + 16) || d.pin[k].ensure(omax + 16)) return -1;
    if (d.ws.ensure(rl_decode_workspace_bytes(cp))) return -1;
    return run_pipeline(
        d, nch,
        [&](size_t i, int k) -> int {
            const size_t q0 = i * cp, m = std::min(cp, np - q0);
            CK(cudaMemcpyAsync(d.pb[k].p, values + q0, m, cudaMemcpyHostToDevice, d.s_in));
            d.bytes_in += m;
            return 0;
        },
        [&](size_t i, int k) -> int {
            const size_t q0 = i * cp, m = std::min(cp, np - q0);
            CK(rl_decode_launch(d.a.p + q0, d.pb[k].p, m, d.pin[k].p, ooff[i + 1] - ooff[i], pipe_result_dev(d, k), d.ws.p, d.s));
            CK(cudaMemcpyAsync(pipe_result(d, k) + 1, rl_decode_status(d.ws.p, m), sizeof(int), cudaMemcpyDeviceToHost, d.s));
            return 0;
        },
        [&](size_t i, int k) -> int {
            if (*reinterpret_cast<int*>(pipe_result(d, k) + 1)) *bad = true;
            const uint64_t nb = ooff[i + 1] - ooff[i];
            if (nb) CK(cudaMemcpyAsync(out + ooff[i], d.pin[k].p, nb, cudaMemcpyDeviceToHost, d.s_out));
            d.bytes_out += nb;
            return 0;
        });
}

// phase A of the decoders on one GPU: the shard's widths / counts (m bytes at `meta`) to d.a, their sums per segment of
// `seg` bytes to d.h_sums (pinned host; one extra word = flags: 1 a byte is 0, 2 a byte is > 8), their total to scal[0]
// for the size exchange.
int meta_sums(Dev& d, const uint8_t* meta, size_t m, size_t nsum, size_t seg, uint64_t* total, uint64_t* flags)
{
    Nvtx r("flrl widths/counts + sums");
    const size_t nseg = nsum ? (nsum + seg - 1) / seg : 0;
    if (d.a.ensure(m + 16) || d.need_sums(nseg + 2)) return -1;
    if (h2d(d, d.a.p, meta, m)) return -1;
    CK(segment_sums_launch(d.a.p, nsum, seg, reinterpret_cast<uint64_t*>(d.sums.p), d.s));
    CK(cudaMemcpyAsync(d.h_sums, d.sums.p, (nseg + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, d.s));
    CK(cudaStreamSynchronize(d.s));
    uint64_t t = 0;
    for (size_t i = 0; i < nseg; ++i) t += d.h_sums[i];
    d.n_sums = nseg;
    *total = t;
    *flags = d.h_sums[nseg];
    return post_size(d, t);
}

}  // namespace

extern "C" {

void flrl_set_concat_mode(int peer) { g_concat_peer = peer != 0; }
void flrl_set_pipe_chunk(size_t bytes) { g_pipe_chunk.store(bytes ? bytes : (size_t)64u << 20); }

void* flrl_host_alloc(size_t bytes)
{
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) {
        fail_cuda(cudaGetLastError(), "flrl_host_alloc");
        return nullptr;
    }
    return p;
}
void flrl_host_free(void* p)
{
    if (p) cudaFreeHost(p);
}
void flrl_last_timing(float* kernel_ms, float* total_ms)
{
    if (kernel_ms) *kernel_ms = tl_kernel_ms;
    if (total_ms) *total_ms = tl_total_ms;
}
int flrl_last_gpu_count(void) { return g_last_gpus; }
int flrl_last_timing_gpu(int g, float* kernel_ms, float* total_ms, uint64_t* bytes_in, uint64_t* bytes_out)
{
    if (g < 0 || g >= g_last_gpus) return fail(FLRL_E_ARG, "flrl_last_timing_gpu: GPU %d did not take part in the last call", g);
    const Dev& d = g_dev[g_map[g]];
    if (kernel_ms) *kernel_ms = d.kernel_ms;
    if (total_ms) *total_ms = d.total_ms;
    if (bytes_in) *bytes_in = d.bytes_in;
    if (bytes_out) *bytes_out = d.bytes_out;
    return FLRL_OK;
}

// ------------------------------------------------------------------------------------ FL compress
int flrl_fl_gpu_compress_into(const uint8_t* data, size_t size, int ngpus, uint8_t* bits, size_t bits_cap, uint8_t* values,
                              size_t values_cap, size_t* bits_size, size_t* values_size)
{
    if (!bits_size || !values_size) return fail(FLRL_E_ARG, "flrl_fl_gpu_compress: NULL size output");
    *bits_size = *values_size = 0;
    if (size == 0) return FLRL_OK;  // fl_gpu.cu:291-294
    if (!data) return fail(FLRL_E_ARG, "flrl_fl_gpu_compress: NULL input");
    std::lock_guard<std::mutex> lk(g_host_mu);
    DeviceGuard restore;
    int G = 1;
    if (int rc = device_count(ngpus, &G)) return rc;
    if (size < (size_t)G * 16384 && G > 1) { G = 1; single_gpu_map(); }
    const size_t frames = (size + FRAME_LENGTH_ - 1) / FRAME_LENGTH_;
    if (bits_cap < frames) return fail(FLRL_E_CAPACITY, "flrl_fl_gpu_compress: bits capacity %zu < %zu", bits_cap, frames);
    const bool piped = pipe_ok(size / G, {data, bits, values});
    std::vector<uint64_t> mine(G, 0);
    int rc = for_each_gpu(G, [&](int g) -> int {
        Dev* dp = nullptr;
        if (begin_gpu(g, &dp)) return -1;
        Dev& d = *dp;
        size_t b0, n;
        shard_range(size, G, g, 128, &b0, &n);
        if (piped) {
            const bool direct = g == 0 && !(G > 1 && g_concat_peer);  // shard 0 starts at offset 0 of the values
            if (int r = fl_compress_shard(d, data + b0, n, bits + b0 / 128, direct ? values : nullptr, values_cap, &mine[g])) return r;
            return post_size(d, mine[g]);
        }
        const size_t n16 = (n + 15) & ~(size_t)15;
        if (d.in.ensure(n16) || d.a.ensure((n + 127) / 128) || d.b.ensure(n16) || d.ws.ensure(fl_workspace_bytes(n))) return -1;
        if (h2d(d, d.in.p, data + b0, n)) return -1;
        CK(cudaEventRecord(d.k0, d.s));
        CK(fl_encode_launch(d.in.p, n, d.a.p, d.b.p, reinterpret_cast<uint64_t*>(d.scal.p), d.ws.p, d.s));
        CK(cudaEventRecord(d.k1, d.s));
        return 0;
    });
    if (rc) return rc;
    std::vector<uint64_t> sizes;
    if ((rc = exchange_sizes(G, sizes))) return rc;
    std::vector<uint64_t> voff(G), boff(G), blen(G);
    std::vector<const uint8_t*> vsrc(G), bsrc(G);
    uint64_t total = 0;
    for (int g = 0; g < G; ++g) {
        size_t b0, n;
        shard_range(size, G, g, 128, &b0, &n);
        if (piped && sizes[g] != mine[g]) return fail(FLRL_E_ARG, "NCCL Error: size all-gather returned %llu for GPU %d, expected %llu",
                                                      (unsigned long long)sizes[g], g, (unsigned long long)mine[g]);
        voff[g] = total;  // exclusive scan of the gathered sizes
        total += sizes[g];
        boff[g] = b0 / 128;
        blen[g] = (n + 127) / 128;
        const bool direct = piped && g == 0 && !(G > 1 && g_concat_peer);
        vsrc[g] = direct ? nullptr : g_dev[g_map[g]].b.p;
        bsrc[g] = piped ? nullptr : g_dev[g_map[g]].a.p;
    }
    if (values_cap < total) return fail(FLRL_E_CAPACITY, "flrl_fl_gpu_compress: values capacity %zu < %llu", values_cap, (unsigned long long)total);
    if (!piped && (rc = deliver(G, bits, bsrc, boff, blen, frames))) return rc;
    if ((rc = deliver(G, values, vsrc, voff, sizes, total))) return rc;
    for (int g = 0; g < G; ++g) {
        if (end_gpu(g_dev[g_map[g]])) return -1;
        if (!piped && add_kernel_time(g_dev[g_map[g]])) return -1;
    }
    collect_timing(G);
    *bits_size = frames;
    *values_size = total;
    return FLRL_OK;
}

int flrl_fl_gpu_compress(const uint8_t* data, size_t size, int ngpus, uint8_t** bits, size_t* bits_size, uint8_t** values,
                         size_t* values_size)
{
    if (!bits || !values) return fail(FLRL_E_ARG, "flrl_fl_gpu_compress: NULL output");
    *bits = *values = nullptr;
    if (size == 0) {
        if (bits_size) *bits_size = 0;
        if (values_size) *values_size = 0;
        return FLRL_OK;
    }
    const size_t frames = (size + FRAME_LENGTH_ - 1) / FRAME_LENGTH_;
    uint8_t* b = static_cast<uint8_t*>(malloc(frames));
    uint8_t* v = static_cast<uint8_t*>(malloc(size + 16));  // worst case: every frame 8 bits wide
    if (!b || !v) {
        free(b);
        free(v);
        return fail(FLRL_E_NOMEM, "flrl_fl_gpu_compress: out of memory");
    }
    const int rc = flrl_fl_gpu_compress_into(data, size, ngpus, b, frames, v, size + 16, bits_size, values_size);
    if (rc) {
        free(b);
        free(v);
        return rc;
    }
    uint8_t* shrunk = static_cast<uint8_t*>(realloc(v, *values_size ? *values_size : 1));
    *bits = b;
    *values = shrunk ? shrunk : v;
    return FLRL_OK;
}

// ---------------------------------------------------------------------------------- FL decompress
int flrl_fl_gpu_decompress_into(size_t output_size, const uint8_t* bits, size_t bits_size, const uint8_t* values,
                                size_t values_size, int ngpus, uint8_t* out, size_t out_cap, size_t* out_size)
{
    if (!out_size) return fail(FLRL_E_ARG, "flrl_fl_gpu_decompress: NULL size output");
    *out_size = 0;
    if (output_size == 0 || bits_size == 0 || values_size == 0) return FLRL_OK;  // fl_gpu.cu:539-543
    if (!bits || !values || !out) return fail(FLRL_E_ARG, "flrl_fl_gpu_decompress: NULL buffer");
    const size_t frames = (output_size + FRAME_LENGTH_ - 1) / FRAME_LENGTH_;
    if (bits_size != frames) return fail(FLRL_E_FORMAT, "flrl_fl_gpu_decompress: bitsSize != ceil(outputSize/128)");
    if (out_cap < output_size) return fail(FLRL_E_CAPACITY, "flrl_fl_gpu_decompress: output capacity too small");
    uint64_t tail_bytes = 0;  // value bytes of the partial last frame
    if (output_size % 128) {
        const unsigned w = bits[frames - 1];
        if (w < 1 || w > 8) return fail(FLRL_E_FORMAT, "flrl_fl_gpu_decompress: malformed container (width %u of the last frame)", w);
        tail_bytes = ((output_size % 128) * w + 7) / 8;
    }
    std::lock_guard<std::mutex> lk(g_host_mu);
    DeviceGuard restore;
    int G = 1;
    if (int rc = device_count(ngpus, &G)) return rc;
    if (output_size < (size_t)G * 16384 && G > 1) { G = 1; single_gpu_map(); }
    const bool piped = pipe_ok(output_size / G, {bits, values, out});
    const size_t CH = pipe_chunk_bytes();

    // phase A: widths of each shard to its GPU; sums of the widths per chunk -> value bytes of every chunk and shard
    std::vector<uint64_t> wsum(G, 0), flags(G, 0);
    int rc = for_each_gpu(G, [&](int g) -> int {
        Dev* dp = nullptr;
        if (begin_gpu(g, &dp)) return -1;
        size_t b0, n;
        shard_range(output_size, G, g, 128, &b0, &n);
        return meta_sums(*dp, bits + b0 / 128, (n + 127) / 128, n / 128 /* full frames */, piped ? CH / 128 : ((n / 128 + 15) & ~(size_t)15) + 16,
                         &wsum[g], &flags[g]);
    });
    if (rc) return rc;
    for (int g = 0; g < G; ++g)
        if (flags[g]) return fail(FLRL_E_FORMAT, "flrl_fl_gpu_decompress: malformed container (a width outside 1..8, GPU %d)", g);
    std::vector<uint64_t> sums;
    if ((rc = exchange_sizes(G, sums))) return rc;
    std::vector<uint64_t> voff(G), vlen(G);
    uint64_t total = 0;
    for (int g = 0; g < G; ++g) {
        if (sums[g] != wsum[g]) return fail(FLRL_E_ARG, "NCCL Error: size all-gather mismatch on GPU %d", g);
        voff[g] = total;
        vlen[g] = 16 * sums[g] + (g == G - 1 ? tail_bytes : 0);
        total += vlen[g];
    }
    if (total > values_size) return fail(FLRL_E_FORMAT, "flrl_fl_gpu_decompress: values shorter than the widths imply");
    // phase B: values of each shard to its GPU, decode, output back
    rc = for_each_gpu(G, [&](int g) -> int {
        Dev& d = g_dev[g_map[g]];
        CK(cudaSetDevice(g_map[g]));
        size_t b0, n;
        shard_range(output_size, G, g, 128, &b0, &n);
        bool bad = false;
        if (piped) {
            if (int r = fl_decompress_shard(d, n, values + voff[g], g == G - 1 ? tail_bytes : 0, out + b0, &bad)) return r;
        } else {
            const size_t nf = (n + 127) / 128;
            if (d.in.ensure(vlen[g] + 16) || d.b.ensure(n + 16) || d.ws.ensure(fl_workspace_bytes(n))) return -1;
            if (h2d(d, d.in.p, values + voff[g], vlen[g])) return -1;
            CK(cudaEventRecord(d.k0, d.s));
            CK(fl_decode_launch(d.a.p, nf, d.in.p, vlen[g], d.b.p, n, d.ws.p, d.s));
            CK(cudaEventRecord(d.k1, d.s));
            CK(cudaMemcpyAsync(d.h_scal + 32, fl_decode_status(d.ws.p, n), sizeof(int), cudaMemcpyDeviceToHost, d.s));
            if (d2h(d, out + b0, d.b.p, n)) return -1;
        }
        if (end_gpu(d)) return -1;
        if (!piped) {
            if (add_kernel_time(d)) return -1;
            bad = *reinterpret_cast<int*>(d.h_scal + 32) != 0;
        }
        if (bad) return fail(FLRL_E_FORMAT, "flrl_fl_gpu_decompress: malformed container (GPU %d)", g);
        return 0;
    });
    if (rc) return rc;
    collect_timing(G);
    *out_size = output_size;
    return FLRL_OK;
}

int flrl_fl_gpu_decompress(size_t output_size, const uint8_t* bits, size_t bits_size, const uint8_t* values,
                           size_t values_size, int ngpus, uint8_t** out, size_t* out_size)
{
    if (!out || !out_size) return fail(FLRL_E_ARG, "flrl_fl_gpu_decompress: NULL output");
    *out = nullptr;
    *out_size = 0;
    if (output_size == 0 || bits_size == 0 || values_size == 0) return FLRL_OK;
    uint8_t* o = static_cast<uint8_t*>(malloc(output_size));
    if (!o) return fail(FLRL_E_NOMEM, "flrl_fl_gpu_decompress: out of memory");
    const int rc = flrl_fl_gpu_decompress_into(output_size, bits, bits_size, values, values_size, ngpus, o, output_size, out_size);
    if (rc) {
        free(o);
        return rc;
    }
    *out = o;
    return FLRL_OK;
}

}  // extern "C"

extern "C" {

// ------------------------------------------------------------------------------------ RL compress
int flrl_rl_gpu_compress_part_into(const uint8_t* data, size_t size, size_t lookahead, uint64_t prev_run, int ngpus, uint8_t* counts,
                                   uint8_t* values, size_t pair_cap, size_t* pairs)
{
    if (!pairs) return fail(FLRL_E_ARG, "flrl_rl_gpu_compress: NULL size output");
    *pairs = 0;
    if (size == 0) return FLRL_OK;
    if (!data || !counts || !values) return fail(FLRL_E_ARG, "flrl_rl_gpu_compress: NULL buffer");
    const size_t avail = size + lookahead;  // bytes behind `data`: pairs start in [0, size), counts may look into the rest
    std::lock_guard<std::mutex> lk(g_host_mu);
    DeviceGuard restore;
    int G = 1;
    if (int rc = device_count(ngpus, &G)) return rc;
    if (size < (size_t)G * 16384 && G > 1) { G = 1; single_gpu_map(); }
    const bool piped = pipe_ok(size / G, {data, counts, values});
    std::vector<uint64_t> mine(G, 0);
    int rc = for_each_gpu(G, [&](int g) -> int {
        Dev* dp = nullptr;
        if (begin_gpu(g, &dp)) return -1;
        Dev& d = *dp;
        size_t b0, n;
        shard_range(size, G, g, 128, &b0, &n);
        if (piped) {
            const bool direct = g == 0 && !(G > 1 && g_concat_peer);  // shard 0's pairs start at pair 0
            if (int r = rl_compress_shard(d, data, avail, prev_run, b0, n, direct ? counts : nullptr, direct ? values : nullptr, pair_cap, &mine[g])) return r;
            return post_size(d, mine[g]);
        }
        // bytes equal to the shard's first byte immediately before it: the run the shard continues
        uint64_t prev = run_before(data, b0, b0);
        if (prev == b0) prev += prev_run;
        const size_t ahead = std::min<size_t>(255, avail - (b0 + n));  // look-ahead: a pair that starts here gets its full count
        const size_t n16 = (n + 15) & ~(size_t)15;
        if (d.in.ensure(n + ahead + 16) || d.a.ensure(n16) || d.b.ensure(n16) || d.ws.ensure(rl_encode_workspace_bytes(n))) return -1;
        if (h2d(d, d.in.p, data + b0, n + ahead)) return -1;
        CK(cudaEventRecord(d.k0, d.s));
        CK(rl_encode_launch(d.in.p, n, n + ahead, prev, d.a.p, d.b.p, n16, reinterpret_cast<uint64_t*>(d.scal.p), d.ws.p, d.s));
        CK(cudaEventRecord(d.k1, d.s));
        return 0;
    });
    if (rc) return rc;
    std::vector<uint64_t> sizes;
    if ((rc = exchange_sizes(G, sizes))) return rc;
    std::vector<uint64_t> off(G);
    std::vector<const uint8_t*> csrc(G), vsrc(G);
    uint64_t total = 0;
    for (int g = 0; g < G; ++g) {
        if (piped && sizes[g] != mine[g]) return fail(FLRL_E_ARG, "NCCL Error: size all-gather returned %llu for GPU %d, expected %llu",
                                                      (unsigned long long)sizes[g], g, (unsigned long long)mine[g]);
        off[g] = total;
        total += sizes[g];
        const bool direct = piped && g == 0 && !(G > 1 && g_concat_peer);
        csrc[g] = direct ? nullptr : g_dev[g_map[g]].a.p;
        vsrc[g] = direct ? nullptr : g_dev[g_map[g]].b.p;
    }
    if (pair_cap < total) return fail(FLRL_E_CAPACITY, "flrl_rl_gpu_compress: pair capacity %zu < %llu", pair_cap, (unsigned long long)total);
    if ((rc = deliver(G, counts, csrc, off, sizes, total))) return rc;
    if ((rc = deliver(G, values, vsrc, off, sizes, total))) return rc;
    for (int g = 0; g < G; ++g) {
        if (end_gpu(g_dev[g_map[g]])) return -1;
        if (!piped && add_kernel_time(g_dev[g_map[g]])) return -1;
    }
    collect_timing(G);
    *pairs = total;
    return FLRL_OK;
}

int flrl_rl_gpu_compress_into(const uint8_t* data, size_t size, int ngpus, uint8_t* counts, uint8_t* values, size_t pair_cap,
                              size_t* pairs)
{
    return flrl_rl_gpu_compress_part_into(data, size, 0, 0, ngpus, counts, values, pair_cap, pairs);
}

int flrl_rl_gpu_compress(const uint8_t* data, size_t size, int ngpus, uint8_t** counts, uint8_t** values, size_t* pairs)
{
    if (!counts || !values || !pairs) return fail(FLRL_E_ARG, "flrl_rl_gpu_compress: NULL output");
    *counts = *values = nullptr;
    *pairs = 0;
    if (size == 0) return FLRL_OK;
    uint8_t* c = static_cast<uint8_t*>(malloc(size));
    uint8_t* v = static_cast<uint8_t*>(malloc(size));
    if (!c || !v) {
        free(c);
        free(v);
        return fail(FLRL_E_NOMEM, "flrl_rl_gpu_compress: out of memory");
    }
    const int rc = flrl_rl_gpu_compress_into(data, size, ngpus, c, v, size, pairs);
    if (rc) {
        free(c);
        free(v);
        return rc;
    }
    uint8_t* c2 = static_cast<uint8_t*>(realloc(c, *pairs ? *pairs : 1));
    uint8_t* v2 = static_cast<uint8_t*>(realloc(v, *pairs ? *pairs : 1));
    *counts = c2 ? c2 : c;
    *values = v2 ? v2 : v;
    return FLRL_OK;
}

// ---------------------------------------------------------------------------------- RL decompress
int flrl_rl_gpu_decompress_into(const uint8_t* counts, const uint8_t* values, size_t pairs, int ngpus, uint8_t* out,
                                size_t out_cap, size_t* out_size)
{
    if (!out_size) return fail(FLRL_E_ARG, "flrl_rl_gpu_decompress: NULL size output");
    *out_size = 0;
    if (pairs == 0) return FLRL_OK;
    if (!counts || !values || !out) return fail(FLRL_E_ARG, "flrl_rl_gpu_decompress: NULL buffer");
    std::lock_guard<std::mutex> lk(g_host_mu);
    DeviceGuard restore;
    int G = 1;
    if (int rc = device_count(ngpus, &G)) return rc;
    if (pairs < (size_t)G * 4096 && G > 1) { G = 1; single_gpu_map(); }
    const bool piped = pipe_ok((2 * pairs + out_cap) / G, {counts, values, out});
    // pairs per chunk: about one pipeline chunk of pairs + output at the stream's mean count (out_cap is the decoded
    // size the caller expects)
    const size_t CH = pipe_chunk_bytes();
    const size_t mean = std::max<size_t>(1, out_cap / pairs);
    const size_t cp = std::min<size_t>(16u << 20, std::max<size_t>(64u << 10, (CH / (2 + mean)) & ~(size_t)4095));

    // phase A: counts of each shard to its GPU; their sums per chunk -> output bytes of every chunk and shard
    std::vector<uint64_t> csum(G, 0), flags(G, 0);
    int rc = for_each_gpu(G, [&](int g) -> int {
        Dev* dp = nullptr;
        if (begin_gpu(g, &dp)) return -1;
        size_t p0, np;
        shard_range(pairs, G, g, 16, &p0, &np);
        return meta_sums(*dp, counts + p0, np, np, piped ? cp : ((np + 15) & ~(size_t)15) + 16, &csum[g], &flags[g]);
    });
    if (rc) return rc;
    for (int g = 0; g < G; ++g)
        if (flags[g] & 1) return fail(FLRL_E_FORMAT, "flrl_rl_gpu_decompress: malformed pair stream (a count of 0, GPU %d)", g);
    std::vector<uint64_t> olen;
    if ((rc = exchange_sizes(G, olen))) return rc;
    std::vector<uint64_t> ooff(G);
    uint64_t total = 0;
    for (int g = 0; g < G; ++g) {
        if (olen[g] != csum[g]) return fail(FLRL_E_ARG, "NCCL Error: size all-gather mismatch on GPU %d", g);
        ooff[g] = total;
        total += olen[g];
    }
    if (total > out_cap) return fail(FLRL_E_CAPACITY, "flrl_rl_gpu_decompress: decoded size %llu exceeds the output capacity %zu",
                                     (unsigned long long)total, out_cap);
    rc = for_each_gpu(G, [&](int g) -> int {
        Dev& d = g_dev[g_map[g]];
        CK(cudaSetDevice(g_map[g]));
        size_t p0, np;
        shard_range(pairs, G, g, 16, &p0, &np);
        bool bad = false;
        if (piped) {
            if (int r = rl_decompress_shard(d, np, cp, values + p0, out + ooff[g], &bad)) return r;
        } else {
            if (d.b.ensure(np + 16) || d.in.ensure(olen[g] + 16) || d.ws.ensure(rl_decode_workspace_bytes(np))) return -1;
            if (h2d(d, d.b.p, values + p0, np)) return -1;
            uint64_t* d_size = reinterpret_cast<uint64_t*>(d.scal.p) + 40;
            CK(cudaEventRecord(d.k0, d.s));
            CK(rl_decode_launch(d.a.p, d.b.p, np, d.in.p, olen[g], d_size, d.ws.p, d.s));
            CK(cudaEventRecord(d.k1, d.s));
            CK(cudaMemcpyAsync(d.h_scal + 32, rl_decode_status(d.ws.p, np), sizeof(int), cudaMemcpyDeviceToHost, d.s));
            if (d2h(d, out + ooff[g], d.in.p, olen[g])) return -1;
        }
        if (end_gpu(d)) return -1;
        if (!piped) {
            if (add_kernel_time(d)) return -1;
            bad = *reinterpret_cast<int*>(d.h_scal + 32) != 0;
        }
        if (bad) return fail(FLRL_E_FORMAT, "flrl_rl_gpu_decompress: malformed pair stream (GPU %d)", g);
        return 0;
    });
    if (rc) return rc;
    collect_timing(G);
    *out_size = total;
    return FLRL_OK;
}

int flrl_rl_gpu_decompress(const uint8_t* counts, const uint8_t* values, size_t pairs, size_t expected_size, int ngpus,
                           uint8_t** out, size_t* out_size)
{
    if (!out || !out_size) return fail(FLRL_E_ARG, "flrl_rl_gpu_decompress: NULL output");
    *out = nullptr;
    *out_size = 0;
    if (pairs == 0) return FLRL_OK;
    const size_t cap = expected_size ? expected_size : pairs * 255;
    uint8_t* o = static_cast<uint8_t*>(malloc(cap));
    if (!o) return fail(FLRL_E_NOMEM, "flrl_rl_gpu_decompress: out of memory");
    const int rc = flrl_rl_gpu_decompress_into(counts, values, pairs, ngpus, o, cap, out_size);
    if (rc) {
        free(o);
        return rc;
    }
    *out = o;
    return FLRL_OK;
}

// One-time costs paid up front instead of inside the first codec call: CUDA context, streams and events of every GPU of
// the call, the kernels' modules (lazy loading: a tiny encode + decode of each codec per GPU) and, for several GPUs, the
// NCCL communicators (ncclCommInitAll: the largest item, seconds on an 8-GPU box).  Optional; the CLI calls it while its
// reader thread fetches the first part and prints the time as its own [TIMER] line.
int flrl_prepare(int ngpus)
{
    std::lock_guard<std::mutex> lk(g_host_mu);
    DeviceGuard restore;
    int G = 1;
    if (int rc = device_count(ngpus, &G)) return rc;
    int rc = for_each_gpu(G, [&](int g) -> int {
        Dev* dp = nullptr;
        if (begin_gpu(g, &dp)) return -1;
        Dev& d = *dp;
        if (d.need_pipe()) return -1;
        const size_t n = 4096;
        if (d.in.ensure(n + 256) || d.a.ensure(n) || d.b.ensure(n) || d.gather.ensure(n) ||
            d.ws.ensure(std::max(fl_workspace_bytes(n), std::max(rl_encode_workspace_bytes(n), rl_decode_workspace_bytes(n)))))
            return -1;
        uint64_t* d_size = reinterpret_cast<uint64_t*>(d.scal.p) + 40;
        CK(cudaMemsetAsync(d.in.p, 0, n + 256, d.s));
        CK(fl_encode_launch(d.in.p, n, d.a.p, d.b.p, d_size, d.ws.p, d.s));           // 32 frames of width 1 -> 512 value bytes
        CK(fl_decode_launch(d.a.p, n / 128, d.b.p, 16 * (n / 128), d.gather.p, n, d.ws.p, d.s));
        CK(rl_encode_launch(d.in.p, n, n, 0, d.a.p, d.b.p, n, d_size, d.ws.p, d.s));  // 17 pairs
        CK(rl_decode_launch(d.a.p, d.b.p, (n + 254) / 255, d.gather.p, n, d_size, d.ws.p, d.s));
        CK(segment_sums_launch(d.a.p, 16, 16, reinterpret_cast<uint64_t*>(d.scal.p) + 32, d.s));
        CK(cudaStreamSynchronize(d.s));
        return 0;
    });
    if (rc) return rc;
    if (G > 1) {
        if ((rc = nccl_ensure(G))) return rc;
        for (int g = 0; g < G; ++g) {
            CK(cudaSetDevice(g_map[g]));
            CK(cudaMemsetAsync(g_dev[g_map[g]].scal.p, 0, sizeof(uint64_t), g_dev[g_map[g]].s));
        }
        std::vector<uint64_t> sizes;
        if ((rc = exchange_sizes(G, sizes))) return rc;  // the first collective sets up NCCL's channels
    }
    return FLRL_OK;
}

// release every cached device buffer, stream, event and pinned staging buffer, of the pipeline too
void flrl_release(void)
{
    std::lock_guard<std::mutex> lk(g_host_mu);
    DeviceGuard restore;
    for (int g = 0; g < MAX_DEV; ++g) g_dev[g].release();
}

}  // extern "C"
