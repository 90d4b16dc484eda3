// args_parser.cuh — command line of the `compress` tool.
// Same contract as the reference (/root/reference/src/args_parser.cuh, args_parser.cu:8-68):
// exactly four positional arguments, usage on stderr + exit(1) on anything else.
#pragma once

namespace ArgsParser {

enum class Operation { Compression, Decompression };

enum class Method {
    FixedLength,      // fl       GPU, 1 device (FLRL_NGPUS overrides)
    FixedLengthCPU,   // fl-cpu
    FixedLengthMulti, // fl-nccl / fl-mpi: every GPU of the box, NCCL all-gather of the shard sizes
    RunLength,        // rl       GPU
    RunLengthCPU,     // rl-cpu
};

struct Args {
    Operation operation;
    Method method;
    const char* inputFile;
    const char* outputFile;
};

Args parseArguments(int argc, char** argv);
[[noreturn]] void usage(const char* program);

}  // namespace ArgsParser
