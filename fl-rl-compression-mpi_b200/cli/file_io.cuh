// file_io.cuh — raw files and the two compressed containers.
//   FL  u64 inputSize, u64 bitsSize, u64 valuesSize, bits[bitsSize], values[valuesSize]
//       byte-identical to the reference's writer / reader (/root/reference/src/file_io.cu:222-280, :117-192)
//   RL  u64 inputSize, u64 pairCount, counts[pairCount], values[pairCount]
//       (build decision — the reference has no RL container; same "sizes, metadata, payload" order)
// Buffers that feed the GPU methods are pinned (flrl_host_alloc) so the host layer's cudaMemcpyAsync goes
// straight from / to them; the CPU methods use malloc.
#pragma once
#include <cstddef>
#include <cstdint>

namespace FileIO {

struct Buffer {
    uint8_t* data = nullptr;
    size_t size = 0;
    bool pinned = false;
    void release();
};
Buffer allocate(size_t size, bool pinned);  // throws std::runtime_error("[FileIO] Cannot allocate memory")

Buffer loadFile(const char* path, bool pinned);
void saveFile(const char* path, const uint8_t* data, size_t size);

struct FLFile {
    size_t inputSize = 0, bitsSize = 0, valuesSize = 0;
    Buffer blob;  // bits followed by values
    const uint8_t* bits() const { return blob.data; }
    const uint8_t* values() const { return blob.data + bitsSize; }
};
FLFile loadCompressedFL(const char* path, bool pinned);
void saveCompressedFL(const char* path, size_t inputSize, const uint8_t* bits, size_t bitsSize, const uint8_t* values,
                      size_t valuesSize);

struct RLFile {
    size_t inputSize = 0, pairCount = 0;
    Buffer blob;  // counts followed by values
    const uint8_t* counts() const { return blob.data; }
    const uint8_t* values() const { return blob.data + pairCount; }
};
RLFile loadCompressedRL(const char* path, bool pinned);
void saveCompressedRL(const char* path, size_t inputSize, const uint8_t* counts, const uint8_t* values, size_t pairCount);

}  // namespace FileIO
