// stream_io.cuh — the GPU methods of the CLI, file to file through bounded pinned buffers (stream_io.cu).
// Errors are thrown as std::runtime_error with the reference's messages; the caller prints "[ERROR]: what".
#pragma once

namespace StreamIO {

// ngpus: 1 = one GPU, 0 = every GPU of the box (the fl-nccl / fl-mpi methods)
void compressFL(const char* input, const char* output, int ngpus);
void decompressFL(const char* input, const char* output, int ngpus);
void compressRL(const char* input, const char* output, int ngpus);
void decompressRL(const char* input, const char* output, int ngpus);

}  // namespace StreamIO
