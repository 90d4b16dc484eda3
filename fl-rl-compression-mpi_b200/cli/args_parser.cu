#include "args_parser.cuh"

#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace ArgsParser {

namespace {
struct Name {
    const char* text;
    Method method;
};
// README.md:22-26 names fl, fl-cpu, rl, rl-cpu; the fork also parses fl-mpi, fl-nccl and fl-shmem
// (src/args_parser.cu:29-49).  fl-mpi / fl-nccl keep working (on_cluster.sh:18-33) and mean "all GPUs of
// this box"; fl-shmem is a stub that runs the CPU codec in the reference (src/main.cu:90-92, :116-118).
const Name kMethods[] = {
    {"fl", Method::FixedLength},          {"fl-cpu", Method::FixedLengthCPU}, {"rl", Method::RunLength},
    {"rl-cpu", Method::RunLengthCPU},     {"fl-nccl", Method::FixedLengthMulti}, {"fl-mpi", Method::FixedLengthMulti},
    {"fl-shmem", Method::FixedLengthCPU},
};
}  // namespace

Args parseArguments(int argc, char** argv)
{
    if (argc != 5) usage(argv[0]);
    Args a{};
    if (strcmp(argv[1], "c") == 0)
        a.operation = Operation::Compression;
    else if (strcmp(argv[1], "d") == 0)
        a.operation = Operation::Decompression;
    else
        usage(argv[0]);
    bool found = false;
    for (const Name& n : kMethods)
        if (strcmp(argv[2], n.text) == 0) {
            a.method = n.method;
            found = true;
        }
    if (!found) usage(argv[0]);
    a.inputFile = argv[3];
    a.outputFile = argv[4];
    return a;
}

void usage(const char* s)
{
    fprintf(stderr, "USAGE: %s operation method input_file output_file\n", s);
    fprintf(stderr, "operation - c (compress) or d (decompress)\n");
    fprintf(stderr,
            "method - fl (fixed-length using gpu), fl-cpu (fixed-length using cpu), rl (run-length using gpu), "
            "rl-cpu (run-length using cpu), fl-nccl / fl-mpi (fixed-length using every gpu of this box)\n");
    std::exit(1);
}

}  // namespace ArgsParser
