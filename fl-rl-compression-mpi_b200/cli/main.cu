// main.cu — `./compress operation method input_file output_file`
//
// Same CLI, progress lines, error convention and on-disk FL container as the reference
// (/root/reference/src/main.cu:18-169): failures are printed as "[ERROR]: <what>" on stderr and the process
// still exits 0 (:95-99, :122-125); only usage errors exit 1.  The MPI bootstrap (:35-70) is gone: one process
// drives every GPU of the box through the host layer of include/flrl.h.
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <stdexcept>

#include "../../include/flrl.h"
#include "args_parser.cuh"
#include "file_io.cuh"
#include "timers.cuh"

namespace {

using ArgsParser::Method;

void check(int rc)
{
    if (rc != FLRL_OK) throw std::runtime_error(flrl_last_error());
}

bool usesGpu(Method m) { return m == Method::FixedLength || m == Method::FixedLengthMulti || m == Method::RunLength; }

int gpuCount(Method m)
{
    if (m == Method::FixedLengthMulti) return 0;  // all visible devices
    const char* e = getenv("FLRL_NGPUS");
    return e ? atoi(e) : 1;
}

void printDeviceTimers(const char* what, size_t bytes)
{
    float kernel_ms = 0, total_ms = 0;
    flrl_last_timing(&kernel_ms, &total_ms);
    char step[96];
    snprintf(step, sizeof step, "%s (kernels)", what);
    Timers::printDeviceResult(step, kernel_ms);
    snprintf(step, sizeof step, "%s (kernels + transfers)", what);
    Timers::printDeviceResult(step, total_ms);
    Timers::printTransfer(what, (double)bytes, total_ms);
    // one line pair per GPU, as the reference prints per rank (src/timers/cpu_timer_with_transfer.cu:16-40): the time of
    // that GPU's shard and the rate at which it moved its bytes over PCIe (both directions)
    const int gpus = flrl_last_gpu_count();
    for (int g = 0; gpus > 1 && g < gpus; ++g) {
        float k = 0, t = 0;
        uint64_t in = 0, out = 0;
        if (flrl_last_timing_gpu(g, &k, &t, &in, &out) != FLRL_OK) break;
        snprintf(step, sizeof step, "%s GPU %d (kernels)", what, g);
        Timers::printDeviceResult(step, k);
        snprintf(step, sizeof step, "%s GPU %d (kernels + transfers)", what, g);
        Timers::printDeviceResult(step, t);
        snprintf(step, sizeof step, "%s GPU %d host<->device", what, g);
        Timers::printTransfer(step, (double)(in + out), t);
    }
}

void compress(Method method, const char* input, const char* output)
{
    const bool gpu = usesGpu(method);
    FileIO::Buffer content, meta, payload;
    uint8_t *m = nullptr, *p = nullptr;  // malloc-owned results of the CPU methods
    try {
        content = FileIO::loadFile(input, gpu);
        const size_t n = content.size;
        Timers::CpuTimer t;
        t.start();
        switch (method) {
            case Method::FixedLength:
            case Method::FixedLengthMulti: {
                size_t nb = 0, nv = 0;
                meta = FileIO::allocate((n + 127) / 128, true);
                payload = FileIO::allocate(n + 16, true);
                check(flrl_fl_gpu_compress_into(content.data, n, gpuCount(method), meta.data, meta.size, payload.data, payload.size, &nb, &nv));
                printDeviceTimers("Compression", n);
                FileIO::saveCompressedFL(output, n, meta.data, nb, payload.data, nv);
                break;
            }
            case Method::FixedLengthCPU: {
                size_t nb = 0, nv = 0;
                check(flrl_fl_cpu_compress(content.data, n, &m, &nb, &p, &nv));
                t.end();
                t.printResult("Compression");
                FileIO::saveCompressedFL(output, n, m, nb, p, nv);
                break;
            }
            case Method::RunLength: {
                size_t pairs = 0;
                meta = FileIO::allocate(n, true);
                payload = FileIO::allocate(n, true);
                check(flrl_rl_gpu_compress_into(content.data, n, gpuCount(method), meta.data, payload.data, n, &pairs));
                printDeviceTimers("Compression", n);
                FileIO::saveCompressedRL(output, n, meta.data, payload.data, pairs);
                break;
            }
            case Method::RunLengthCPU: {
                size_t pairs = 0;
                check(flrl_rl_cpu_compress(content.data, n, &m, &p, &pairs));
                t.end();
                t.printResult("Compression");
                FileIO::saveCompressedRL(output, n, m, p, pairs);
                break;
            }
        }
    } catch (const std::exception& e) {
        std::cerr << "[ERROR]: " << e.what() << '\n';
    }
    flrl_free(m);
    flrl_free(p);
    content.release();
    meta.release();
    payload.release();
}

void decompress(Method method, const char* input, const char* output)
{
    const bool gpu = usesGpu(method);
    const bool rl = method == Method::RunLength || method == Method::RunLengthCPU;
    FileIO::FLFile fl;
    FileIO::RLFile rlf;
    FileIO::Buffer out;
    uint8_t* o = nullptr;  // malloc-owned result of the CPU methods
    try {
        size_t n = 0;
        Timers::CpuTimer t;
        if (rl) {
            rlf = FileIO::loadCompressedRL(input, gpu);
            t.start();
            if (gpu) {
                out = FileIO::allocate(rlf.inputSize, true);
                check(flrl_rl_gpu_decompress_into(rlf.counts(), rlf.values(), rlf.pairCount, gpuCount(method), out.data, out.size, &n));
                printDeviceTimers("Decompression", n);
            } else {
                check(flrl_rl_cpu_decompress(rlf.counts(), rlf.values(), rlf.pairCount, &o, &n));
                t.end();
                t.printResult("Decompression");
            }
            if (n != rlf.inputSize) throw std::runtime_error("[RL] decoded size differs from the size in the header");
        } else {
            fl = FileIO::loadCompressedFL(input, gpu);
            t.start();
            if (gpu) {
                out = FileIO::allocate(fl.inputSize, true);
                check(flrl_fl_gpu_decompress_into(fl.inputSize, fl.bits(), fl.bitsSize, fl.values(), fl.valuesSize, gpuCount(method),
                                                  out.data, out.size, &n));
                printDeviceTimers("Decompression", n);
            } else {
                check(flrl_fl_cpu_decompress(fl.inputSize, fl.bits(), fl.bitsSize, fl.values(), fl.valuesSize, &o, &n));
                t.end();
                t.printResult("Decompression");
            }
        }
        FileIO::saveFile(output, gpu ? out.data : o, n);
    } catch (const std::exception& e) {
        std::cerr << "[ERROR]: " << e.what() << '\n';
    }
    flrl_free(o);
    out.release();
    fl.blob.release();
    rlf.blob.release();
}

}  // namespace

int main(int argc, char** argv)
{
    const ArgsParser::Args args = ArgsParser::parseArguments(argc, argv);
    if (args.operation == ArgsParser::Operation::Compression)
        compress(args.method, args.inputFile, args.outputFile);
    else
        decompress(args.method, args.inputFile, args.outputFile);
    return 0;
}
