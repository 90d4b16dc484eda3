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
#include "stream_io.cuh"
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

// contexts, kernel modules and (several GPUs) NCCL communicators, timed apart from the codec: in a short-lived process
// this set-up, not the codec, is most of a multi-GPU run on a small file
void prepareGpus(Method m)
{
    Timers::CpuTimer t;
    t.start();
    check(flrl_prepare(gpuCount(m)));
    t.end();
    t.printResult("GPU set-up (contexts, kernel modules, NCCL communicators)");
}

// one line pair per GPU of the last part, as the reference prints per rank (src/timers/cpu_timer_with_transfer.cu:16-40):
// the time of that GPU's shard and the rate at which it moved its bytes over PCIe (both directions)
void printGpuTimers(const char* what)
{
    char step[96];
    const int gpus = flrl_last_gpu_count();
    for (int g = 0; gpus > 1 && g < gpus; ++g) {
        float k = 0, t = 0;
        uint64_t in = 0, out = 0;
        if (flrl_last_timing_gpu(g, &k, &t, &in, &out) != FLRL_OK) break;
        snprintf(step, sizeof step, "%s GPU %d, last part (kernels)", what, g);
        Timers::printDeviceResult(step, k);
        snprintf(step, sizeof step, "%s GPU %d, last part (kernels + transfers)", what, g);
        Timers::printDeviceResult(step, t);
        snprintf(step, sizeof step, "%s GPU %d, last part host<->device", what, g);
        Timers::printTransfer(step, (double)(in + out), t);
    }
}

void compress(Method method, const char* input, const char* output)
{
    if (usesGpu(method)) {  // file to file through bounded pinned buffers: read, GPU and write overlap (stream_io.cu)
        try {
            prepareGpus(method);
            if (method == Method::RunLength)
                StreamIO::compressRL(input, output, gpuCount(method));
            else
                StreamIO::compressFL(input, output, gpuCount(method));
            printGpuTimers("Compression");
        } catch (const std::exception& e) {
            std::cerr << "[ERROR]: " << e.what() << '\n';
            remove(output);  // no partial container is left behind
        }
        return;
    }
    FileIO::Buffer content;
    uint8_t *m = nullptr, *p = nullptr;  // malloc-owned results of the CPU methods
    try {
        content = FileIO::loadFile(input, false);
        const size_t n = content.size;
        Timers::CpuTimer t;
        t.start();
        if (method == Method::RunLengthCPU) {
            size_t pairs = 0;
            check(flrl_rl_cpu_compress(content.data, n, &m, &p, &pairs));
            t.end();
            t.printResult("Compression");
            FileIO::saveCompressedRL(output, n, m, p, pairs);
        } else {
            size_t nb = 0, nv = 0;
            check(flrl_fl_cpu_compress(content.data, n, &m, &nb, &p, &nv));
            t.end();
            t.printResult("Compression");
            FileIO::saveCompressedFL(output, n, m, nb, p, nv);
        }
    } catch (const std::exception& e) {
        std::cerr << "[ERROR]: " << e.what() << '\n';
    }
    flrl_free(m);
    flrl_free(p);
    content.release();
}

void decompress(Method method, const char* input, const char* output)
{
    const bool rl = method == Method::RunLength || method == Method::RunLengthCPU;
    if (usesGpu(method)) {
        try {
            prepareGpus(method);
            if (rl)
                StreamIO::decompressRL(input, output, gpuCount(method));
            else
                StreamIO::decompressFL(input, output, gpuCount(method));
            printGpuTimers("Decompression");
        } catch (const std::exception& e) {
            std::cerr << "[ERROR]: " << e.what() << '\n';
            remove(output);
        }
        return;
    }
    FileIO::FLFile fl;
    FileIO::RLFile rlf;
    uint8_t* o = nullptr;  // malloc-owned result of the CPU methods
    try {
        size_t n = 0;
        Timers::CpuTimer t;
        if (rl) {
            rlf = FileIO::loadCompressedRL(input, false);
            t.start();
            check(flrl_rl_cpu_decompress(rlf.counts(), rlf.values(), rlf.pairCount, &o, &n));
            t.end();
            t.printResult("Decompression");
            if (n != rlf.inputSize) throw std::runtime_error("[RL] decoded size differs from the size in the header");
        } else {
            fl = FileIO::loadCompressedFL(input, false);
            t.start();
            check(flrl_fl_cpu_decompress(fl.inputSize, fl.bits(), fl.bitsSize, fl.values(), fl.valuesSize, &o, &n));
            t.end();
            t.printResult("Decompression");
        }
        FileIO::saveFile(output, o, n);
    } catch (const std::exception& e) {
        std::cerr << "[ERROR]: " << e.what() << '\n';
    }
    flrl_free(o);
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
