// stream_io.cu — the GPU methods of the CLI, file to file, through bounded pinned buffers.
//
// Replaces the reference's whole-file malloc + fread, codec call, whole-file fwrite
// (/root/reference/src/file_io.cu:73-115, :194-280; src/main.cu:72-129): the file is cut into PARTS (128 MiB per GPU,
// 256 to 512 MiB in all); a reader thread fills part j+1 (parallel preads into pinned memory) while the host
// layer of include/flrl.h runs part j on the GPUs (its own chunk pipeline and shards inside) and a writer thread
// writes part j-1 (parallel pwrites).  Peak host memory is three parts, not the file.  Part outputs concatenate to
// the single-stream containers: FL parts are cut at multiples of 128 bytes; RL parts carry the run that is open at
// their start (prev_run) and 255 bytes of look-ahead.
#include "stream_io.cuh"

#include <cuda_runtime.h>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <memory>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "../../include/flrl.h"
#include "timers.cuh"

namespace StreamIO {

namespace {

constexpr int SLOTS = 3;
constexpr size_t IO_SLICE = 8u << 20;  // bytes per pread / pwrite call of one I/O thread

double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

struct Fd {
    int fd = -1;
    Fd(const char* path, int flags, const char* err)
    {
        fd = open(path, flags, 0644);
        if (fd < 0) throw std::runtime_error(err);
    }
    ~Fd() { if (fd >= 0) close(fd); }
    size_t size() const
    {
        struct stat st;
        if (fstat(fd, &st) != 0) throw std::runtime_error("[FileIO] Cannot read file content");
        return (size_t)st.st_size;
    }
};

// n bytes at file offset `off`, in slices on up to 4 threads (a single memcpy from / to the page cache is the bottleneck)
void io_all(int fd, uint8_t* buf, size_t n, size_t off, bool write)
{
    if (n == 0) return;
    const int T = (int)std::min<size_t>(4, (n + IO_SLICE - 1) / IO_SLICE);
    std::vector<std::thread> th;
    std::vector<int> bad(T, 0);
    auto work = [&](int t) {
        for (size_t o = (size_t)t * IO_SLICE; o < n; o += (size_t)T * IO_SLICE) {
            size_t m = std::min(IO_SLICE, n - o), done = 0;
            while (done < m) {
                const ssize_t r = write ? pwrite(fd, buf + o + done, m - done, (off_t)(off + o + done)) : pread(fd, buf + o + done, m - done, (off_t)(off + o + done));
                if (r <= 0) {
                    bad[t] = 1;
                    return;
                }
                done += (size_t)r;
            }
        }
    };
    for (int t = 1; t < T; ++t) th.emplace_back(work, t);
    work(0);
    for (auto& t : th) t.join();
    for (int b : bad)
        if (b) throw std::runtime_error(write ? "[FileIO] Cannot write to file" : "[FileIO] Cannot read file content");
}

struct Pinned {
    uint8_t* p = nullptr;
    explicit Pinned(size_t n) : p(static_cast<uint8_t*>(flrl_host_alloc(n ? n : 1)))
    {
        if (!p) throw std::runtime_error(flrl_last_error());
    }
    ~Pinned() { flrl_host_free(p); }
    Pinned(const Pinned&) = delete;
};

// Three stages (read, GPU, write) over a ring of SLOTS slots.  stage s may work on part j once stage s-1 has finished
// part j and stage s has finished part j-1; the reader may refill a slot once the writer has finished part j-SLOTS.
struct Ring {
    std::mutex mu;
    std::condition_variable cv;
    long done[3] = {-1, -1, -1};  // last part finished by the reader / the GPU stage / the writer
    std::string error;
    double busy_ms[3] = {0, 0, 0};
    void wait_turn(int stage, long j)
    {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return !error.empty() || (stage == 0 ? done[2] >= j - SLOTS : done[stage - 1] >= j); });
        if (!error.empty()) throw std::runtime_error(error);
    }
    void finish(int stage, long j)
    {
        std::lock_guard<std::mutex> lk(mu);
        done[stage] = j;
        cv.notify_all();
    }
    void fail(const std::string& what)
    {
        std::lock_guard<std::mutex> lk(mu);
        if (error.empty()) error = what;
        cv.notify_all();
    }
};

// runs read(j), gpu(j), write(j) for j = 0 .. parts-1 on three threads; gpu() runs on the calling thread
template <class R, class G, class W>
void run_stages(Ring& ring, long parts, R read, G gpu, W write)
{
    auto stage = [&](int s, auto& fn) {
        try {
            for (long j = 0; j < parts; ++j) {
                ring.wait_turn(s, j);
                const double t0 = now_ms();
                fn(j, (int)(j % SLOTS));
                ring.busy_ms[s] += now_ms() - t0;
                ring.finish(s, j);
            }
        } catch (const std::exception& e) {
            ring.fail(e.what());
        }
    };
    std::thread tr([&] { stage(0, read); }), tw([&] { stage(2, write); });
    stage(1, gpu);
    tr.join();
    tw.join();
    if (!ring.error.empty()) throw std::runtime_error(ring.error);
}

void check(int rc)
{
    if (rc != FLRL_OK) throw std::runtime_error(flrl_last_error());
}

size_t part_bytes(int gpus)
{
    // test hook of the CLI (not of the library): small parts make a small file take the many-parts path
    if (const char* e = getenv("FLRL_CLI_PART_BYTES")) {
        const size_t v = strtoull(e, nullptr, 10) & ~(size_t)4095;
        if (v >= 65536) return v;
    }
    // 128 MiB per GPU (two pipeline chunks per shard), at least 256 MiB, at most 512 MiB: three slots of in + out buffers
    // are pinned up front, and pinning costs about 0.6 s per GiB — more than the codec needs for the whole file
    const size_t g = (size_t)std::max(1, gpus);
    return std::min<size_t>((size_t)512u << 20, std::max<size_t>((size_t)256u << 20, g * ((size_t)128u << 20)));
}

int gpus_of(int ngpus)
{
    int have = 0;
    if (cudaGetDeviceCount(&have) != cudaSuccess || have < 1) have = 1;  // the host layer reports the error itself
    return ngpus > 0 ? std::min(ngpus, have) : have;  // 0: every GPU of the box
}

struct Acc {  // device timers of the parts, summed (max over the GPUs of a part, as the host layer reports them)
    float kernel_ms = 0, total_ms = 0;
    void add()
    {
        float k = 0, t = 0;
        flrl_last_timing(&k, &t);
        kernel_ms += k;
        total_ms += t;
    }
};

void print_timers(const char* what, const Ring& ring, const Acc& acc, size_t bytes, double wall_ms)
{
    printf("[TIMER] Step: \"Load data from file\", Time: %ld ms\n", (long)ring.busy_ms[0]);
    char step[96];
    snprintf(step, sizeof step, "%s (kernels)", what);
    Timers::printDeviceResult(step, acc.kernel_ms);
    snprintf(step, sizeof step, "%s (kernels + transfers)", what);
    Timers::printDeviceResult(step, acc.total_ms);
    Timers::printTransfer(what, (double)bytes, acc.total_ms);
    printf("[TIMER] Step: \"Save data to file\", Time: %ld ms\n", (long)ring.busy_ms[2]);
    snprintf(step, sizeof step, "%s (file to file, read / GPU / write overlapped)", what);
    printf("[TIMER] Step: \"%s\", Time: %ld ms\n", step, (long)wall_ms);
    snprintf(step, sizeof step, "%s file to file", what);
    Timers::printTransfer(step, (double)bytes, wall_ms);
}

}  // namespace

// ------------------------------------------------------------------------------------------------ FL
void compressFL(const char* input, const char* output, int ngpus)
{
    const double t0 = now_ms();
    Fd in(input, O_RDONLY, "[FileIO] Cannot open file");
    Fd out(output, O_WRONLY | O_CREAT | O_TRUNC, "[FileIO] Cannot open file");
    const size_t n = in.size(), frames = (n + 127) / 128;
    const size_t S = part_bytes(gpus_of(ngpus));
    const long parts = (long)((n + S - 1) / S);
    const size_t cap = std::min(S, n);
    std::vector<std::unique_ptr<Pinned>> din, dbits, dvals;
    for (int k = 0; k < SLOTS && k < std::max(1L, parts); ++k) {
        din.emplace_back(new Pinned(cap));
        dbits.emplace_back(new Pinned(cap / 128 + 1));
        dvals.emplace_back(new Pinned(cap + 16));
    }
    std::vector<size_t> nv(std::max(1L, parts), 0), voff(std::max(1L, parts) + 1, 0);
    Ring ring;
    Acc acc;
    auto len = [&](long j) { return std::min(S, n - (size_t)j * S); };
    run_stages(
        ring, parts, [&](long j, int k) { io_all(in.fd, din[k]->p, len(j), (size_t)j * S, false); },
        [&](long j, int k) {
            size_t nb = 0;
            check(flrl_fl_gpu_compress_into(din[k]->p, len(j), ngpus, dbits[k]->p, cap / 128 + 1, dvals[k]->p, cap + 16, &nb, &nv[j]));
            voff[j + 1] = voff[j] + nv[j];
            acc.add();
        },
        [&](long j, int k) {
            io_all(out.fd, dbits[k]->p, (len(j) + 127) / 128, 24 + (size_t)j * (S / 128), true);
            io_all(out.fd, dvals[k]->p, nv[j], 24 + frames + voff[j], true);
        });
    const uint64_t hdr[3] = {n, n ? frames : 0, voff[std::max(0L, parts)]};  // src/file_io.cu:236-252: sizes first
    io_all(out.fd, reinterpret_cast<uint8_t*>(const_cast<uint64_t*>(hdr)), sizeof hdr, 0, true);
    print_timers("Compression", ring, acc, n, now_ms() - t0);
}

void decompressFL(const char* input, const char* output, int ngpus)
{
    const double t0 = now_ms();
    Fd in(input, O_RDONLY, "[FileIO] Cannot open file");
    const size_t total = in.size();
    uint64_t hdr[3] = {0, 0, 0};
    if (total < sizeof hdr) throw std::runtime_error("[FileIO] Cannot read file content");
    io_all(in.fd, reinterpret_cast<uint8_t*>(hdr), sizeof hdr, 0, false);
    const size_t n = hdr[0], nbits = hdr[1], nvals = hdr[2];
    if (nbits > total || nvals > total || sizeof hdr + nbits + nvals > total) throw std::runtime_error("[FileIO] Cannot read file content");
    Fd out(output, O_WRONLY | O_CREAT | O_TRUNC, "[FileIO] Cannot open file");
    if (n == 0 || nbits == 0 || nvals == 0) return;  // fl_gpu.cu:539-543: nothing to decode
    if (nbits != (n + 127) / 128) throw std::runtime_error("flrl_fl_gpu_decompress: bitsSize != ceil(outputSize/128)");
    // the widths (1/128 of the output) are read whole: the place of every part's values is 16 x the sum of the widths before it
    std::vector<uint8_t> bits(nbits);
    io_all(in.fd, bits.data(), nbits, sizeof hdr, false);
    const size_t S = part_bytes(gpus_of(ngpus));
    const long parts = (long)((n + S - 1) / S);
    std::vector<size_t> voff(parts + 1, 0);
    for (long j = 0; j < parts; ++j) {
        const size_t f0 = (size_t)j * (S / 128), f1 = std::min(nbits, f0 + S / 128);
        size_t sum = 0;
        for (size_t f = f0; f < f1; ++f) sum += bits[f];
        size_t v = 16 * sum;
        if (j == parts - 1 && n % 128) v = v - 16 * bits[nbits - 1] + ((n % 128) * bits[nbits - 1] + 7) / 8;  // the partial last frame
        voff[j + 1] = voff[j] + v;
        if (v > S + 16) throw std::runtime_error("flrl_fl_gpu_decompress: malformed container (a width outside 1..8)");
    }
    if (voff[parts] > nvals) throw std::runtime_error("flrl_fl_gpu_decompress: values shorter than the widths imply");
    const size_t cap = std::min(S, n);
    std::vector<std::unique_ptr<Pinned>> dbits, dvals, dout;
    for (int k = 0; k < SLOTS && k < parts; ++k) {
        dbits.emplace_back(new Pinned(cap / 128 + 1));
        dvals.emplace_back(new Pinned(cap + 16));
        dout.emplace_back(new Pinned(cap));
    }
    Ring ring;
    Acc acc;
    auto len = [&](long j) { return std::min(S, n - (size_t)j * S); };
    run_stages(
        ring, parts,
        [&](long j, int k) {
            memcpy(dbits[k]->p, bits.data() + (size_t)j * (S / 128), (len(j) + 127) / 128);
            io_all(in.fd, dvals[k]->p, voff[j + 1] - voff[j], sizeof hdr + nbits + voff[j], false);
        },
        [&](long j, int k) {
            size_t got = 0;
            check(flrl_fl_gpu_decompress_into(len(j), dbits[k]->p, (len(j) + 127) / 128, dvals[k]->p, voff[j + 1] - voff[j], ngpus, dout[k]->p, cap, &got));
            if (got != len(j)) throw std::runtime_error("[FL] decoded size differs from the size in the header");
            acc.add();
        },
        [&](long j, int k) { io_all(out.fd, dout[k]->p, len(j), (size_t)j * S, true); });
    print_timers("Decompression", ring, acc, n, now_ms() - t0);
}

// ------------------------------------------------------------------------------------------------ RL
void compressRL(const char* input, const char* output, int ngpus)
{
    const double t0 = now_ms();
    Fd in(input, O_RDONLY, "[FileIO] Cannot open file");
    Fd out(output, O_RDWR | O_CREAT | O_TRUNC, "[FileIO] Cannot open file");  // read back once: the parked values
    const size_t n = in.size();
    const size_t S = part_bytes(gpus_of(ngpus));
    const long parts = (long)((n + S - 1) / S);
    const size_t cap = std::min(S, n);
    std::vector<std::unique_ptr<Pinned>> din, dc, dv;
    for (int k = 0; k < SLOTS && k < std::max(1L, parts); ++k) {
        din.emplace_back(new Pinned(cap + 256));
        dc.emplace_back(new Pinned(cap + 16));
        dv.emplace_back(new Pinned(cap + 16));
    }
    std::vector<size_t> np(std::max(1L, parts), 0), poff(std::max(1L, parts) + 1, 0);
    std::vector<uint64_t> prev(std::max(1L, parts), 0);
    // the run that is open at the end of the part read last: its byte and its length (carried over whole parts)
    int tail_byte = -1;
    uint64_t tail_run = 0;
    // the container puts all counts before all values, and the pair count P is known only at the end: the values are parked in
    // the output file itself, behind the place of the last possible count (P <= n), and moved down to 16 + P when the last
    // part is done (P bytes through one pinned buffer; little for the inputs RL is meant for).  Host memory stays O(parts).
    const size_t park = 16 + n;
    Ring ring;
    Acc acc;
    auto len = [&](long j) { return std::min(S, n - (size_t)j * S); };
    auto ahead = [&](long j) { return std::min<size_t>(255, n - ((size_t)j * S + len(j))); };
    run_stages(
        ring, parts,
        [&](long j, int k) {
            const uint8_t* d = din[k]->p;
            const size_t m = len(j);
            io_all(in.fd, din[k]->p, m + ahead(j), (size_t)j * S, false);
            prev[j] = (tail_byte == d[0]) ? tail_run : 0;
            size_t t = 1;
            while (t < m && d[m - 1 - t] == d[m - 1]) ++t;  // bytes equal to the last one at the end of the part
            tail_run = (t == m && tail_byte == d[m - 1]) ? tail_run + m : t;
            tail_byte = d[m - 1];
        },
        [&](long j, int k) {
            check(flrl_rl_gpu_compress_part_into(din[k]->p, len(j), ahead(j), prev[j], ngpus, dc[k]->p, dv[k]->p, cap + 16, &np[j]));
            poff[j + 1] = poff[j] + np[j];
            acc.add();
        },
        [&](long j, int k) {
            io_all(out.fd, dc[k]->p, np[j], 16 + poff[j], true);
            io_all(out.fd, dv[k]->p, np[j], park + poff[j], true);
        });
    const size_t P = poff[std::max(0L, parts)];
    const double tw = now_ms();
    if (P && parts > 0) {  // values: from park .. park + P down to 16 + P .. 16 + 2 P, ascending (source is always ahead of the target)
        uint8_t* buf = dv[0]->p;
        for (size_t o = 0; o < P; o += cap) {
            const size_t m = std::min(cap, P - o);
            io_all(out.fd, buf, m, park + o, false);
            io_all(out.fd, buf, m, 16 + P + o, true);
        }
    }
    if (ftruncate(out.fd, (off_t)(16 + 2 * P)) != 0) throw std::runtime_error("[FileIO] Cannot write to file");
    const uint64_t hdr[2] = {n, P};
    io_all(out.fd, reinterpret_cast<uint8_t*>(const_cast<uint64_t*>(hdr)), sizeof hdr, 0, true);
    ring.busy_ms[2] += now_ms() - tw;
    print_timers("Compression", ring, acc, n, now_ms() - t0);
}

void decompressRL(const char* input, const char* output, int ngpus)
{
    const double t0 = now_ms();
    Fd in(input, O_RDONLY, "[FileIO] Cannot open file");
    const size_t total = in.size();
    uint64_t hdr[2] = {0, 0};
    if (total < sizeof hdr) throw std::runtime_error("[FileIO] Cannot read file content");
    io_all(in.fd, reinterpret_cast<uint8_t*>(hdr), sizeof hdr, 0, false);
    const size_t n = hdr[0], P = hdr[1];
    if (P > total || sizeof hdr + 2 * P > total) throw std::runtime_error("[FileIO] Cannot read file content");
    Fd out(output, O_WRONLY | O_CREAT | O_TRUNC, "[FileIO] Cannot open file");
    if (P == 0) {
        if (n != 0) throw std::runtime_error("[RL] decoded size differs from the size in the header");
        return;
    }
    // parts of pairs: as many pairs as decode to at most one part of output (the reader sums the counts it has read)
    const size_t S = part_bytes(gpus_of(ngpus));
    const size_t pcap = std::min(P, S);  // pairs per part at most (a pair decodes to >= 1 byte)
    std::vector<std::unique_ptr<Pinned>> dc, dv, dout;
    for (int k = 0; k < SLOTS; ++k) {
        dc.emplace_back(new Pinned(pcap + 16));
        dv.emplace_back(new Pinned(pcap + 16));
        dout.emplace_back(new Pinned(S + 16));
    }
    // the number of parts is not known in advance: the reader decides part by part and marks the last one
    std::vector<size_t> p0s, nps, olen, ooff;
    size_t next_pair = 0, next_out = 0;
    Ring ring;
    Acc acc;
    std::mutex plan_mu;
    // stage loops need a part count: run them until the reader says "done" by using an upper bound and empty parts
    const long max_parts = (long)((n + S - 1) / S + (P + pcap - 1) / pcap + 2);
    auto part = [&](long j, size_t* p0, size_t* m, size_t* ol, size_t* oo) {
        std::lock_guard<std::mutex> lk(plan_mu);
        *p0 = p0s[j];
        *m = nps[j];
        *ol = olen[j];
        *oo = ooff[j];
    };
    run_stages(
        ring, max_parts,
        [&](long, int k) {
            size_t m = std::min(pcap, P - next_pair), sum = 0, take = 0;
            if (m) io_all(in.fd, dc[k]->p, m, sizeof hdr + next_pair, false);
            const uint8_t* c = dc[k]->p;
            while (take < m && sum + c[take] <= S) sum += c[take++];
            if (take) io_all(in.fd, dv[k]->p, take, sizeof hdr + P + next_pair, false);
            std::lock_guard<std::mutex> lk(plan_mu);
            p0s.push_back(next_pair);
            nps.push_back(take);
            olen.push_back(sum);
            ooff.push_back(next_out);
            next_pair += take;
            next_out += sum;
        },
        [&](long j, int k) {
            size_t p0, m, ol, oo, got = 0;
            part(j, &p0, &m, &ol, &oo);
            if (!m) return;
            check(flrl_rl_gpu_decompress_into(dc[k]->p, dv[k]->p, m, ngpus, dout[k]->p, S + 16, &got));
            if (got != ol) throw std::runtime_error("[RL] decoded size differs from the sum of the counts");
            acc.add();
        },
        [&](long j, int k) {
            size_t p0, m, ol, oo;
            part(j, &p0, &m, &ol, &oo);
            if (m) io_all(out.fd, dout[k]->p, ol, oo, true);
        });
    if (next_pair != P || next_out != n) throw std::runtime_error("[RL] decoded size differs from the size in the header");
    print_timers("Decompression", ring, acc, n, now_ms() - t0);
}

}  // namespace StreamIO
