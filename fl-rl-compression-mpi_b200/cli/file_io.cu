#include "file_io.cuh"

#include <cstdio>
#include <cstdlib>
#include <stdexcept>

#include "../../include/flrl.h"
#include "timers.cuh"

namespace FileIO {

namespace {

struct File {  // closes on every path (the reference leaks the FILE* on success in loadCompressedFL, file_io.cu:188-191)
    FILE* f;
    File(const char* path, const char* mode) : f(fopen(path, mode))
    {
        if (!f) throw std::runtime_error("[FileIO] Cannot open file");
    }
    ~File()
    {
        if (f) fclose(f);
    }
    size_t size()
    {
        if (fseeko(f, 0, SEEK_END) != 0) throw std::runtime_error("[FileIO] Cannot read file content");
        const off_t n = ftello(f);
        if (n < 0 || fseeko(f, 0, SEEK_SET) != 0) throw std::runtime_error("[FileIO] Cannot read file content");
        return (size_t)n;
    }
    void read(void* dst, size_t n)
    {
        if (n && fread(dst, 1, n, f) != n) throw std::runtime_error("[FileIO] Cannot read file content");
    }
    void write(const void* src, size_t n)
    {
        if (n && fwrite(src, 1, n, f) != n) throw std::runtime_error("[FileIO] Cannot write to file");
    }
};

}  // namespace

void Buffer::release()
{
    if (pinned)
        flrl_host_free(data);
    else
        free(data);
    data = nullptr;
    size = 0;
}

Buffer allocate(size_t size, bool pinned)
{
    Buffer b;
    b.pinned = pinned;
    b.size = size;
    b.data = static_cast<uint8_t*>(pinned ? flrl_host_alloc(size ? size : 1) : malloc(size ? size : 1));
    if (!b.data) throw std::runtime_error(pinned ? flrl_last_error() : "[FileIO] Cannot allocate memory");
    return b;
}

Buffer loadFile(const char* path, bool pinned)
{
    Timers::CpuTimer t;
    t.start();
    File f(path, "rb");
    Buffer b = allocate(f.size(), pinned);
    try {
        f.read(b.data, b.size);
    } catch (...) {
        b.release();
        throw;
    }
    t.end();
    t.printResult("Load data from file");
    return b;
}

void saveFile(const char* path, const uint8_t* data, size_t size)
{
    Timers::CpuTimer t;
    t.start();
    File f(path, "wb");
    f.write(data, size);
    t.end();
    t.printResult("Save data to file");
}

FLFile loadCompressedFL(const char* path, bool pinned)
{
    Timers::CpuTimer t;
    t.start();
    File f(path, "rb");
    const size_t total = f.size();
    uint64_t hdr[3];
    f.read(hdr, sizeof hdr);  // inputSize, bitsSize, valuesSize — native little-endian size_t (file_io.cu:236-252)
    FLFile c;
    c.inputSize = hdr[0];
    c.bitsSize = hdr[1];
    c.valuesSize = hdr[2];
    if (c.bitsSize > total || c.valuesSize > total || sizeof hdr + c.bitsSize + c.valuesSize > total)
        throw std::runtime_error("[FileIO] Cannot read file content");
    c.blob = allocate(c.bitsSize + c.valuesSize, pinned);
    try {
        f.read(c.blob.data, c.bitsSize + c.valuesSize);
    } catch (...) {
        c.blob.release();
        throw;
    }
    t.end();
    t.printResult("Load data from file");
    return c;
}

void saveCompressedFL(const char* path, size_t inputSize, const uint8_t* bits, size_t bitsSize, const uint8_t* values,
                      size_t valuesSize)
{
    Timers::CpuTimer t;
    t.start();
    File f(path, "wb");
    const uint64_t hdr[3] = {inputSize, bitsSize, valuesSize};
    f.write(hdr, sizeof hdr);
    f.write(bits, bitsSize);
    f.write(values, valuesSize);
    t.end();
    t.printResult("Save data to file");
}

RLFile loadCompressedRL(const char* path, bool pinned)
{
    Timers::CpuTimer t;
    t.start();
    File f(path, "rb");
    const size_t total = f.size();
    uint64_t hdr[2];
    f.read(hdr, sizeof hdr);
    RLFile c;
    c.inputSize = hdr[0];
    c.pairCount = hdr[1];
    if (c.pairCount > total || sizeof hdr + 2 * c.pairCount > total) throw std::runtime_error("[FileIO] Cannot read file content");
    c.blob = allocate(2 * c.pairCount, pinned);
    try {
        f.read(c.blob.data, 2 * c.pairCount);
    } catch (...) {
        c.blob.release();
        throw;
    }
    t.end();
    t.printResult("Load data from file");
    return c;
}

void saveCompressedRL(const char* path, size_t inputSize, const uint8_t* counts, const uint8_t* values, size_t pairCount)
{
    Timers::CpuTimer t;
    t.start();
    File f(path, "wb");
    const uint64_t hdr[2] = {inputSize, pairCount};
    f.write(hdr, sizeof hdr);
    f.write(counts, pairCount);
    f.write(values, pairCount);
    t.end();
    t.printResult("Save data to file");
}

}  // namespace FileIO
