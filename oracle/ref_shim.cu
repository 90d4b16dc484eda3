/*
 * ref_shim.cu — extern "C" doorway into the UNMODIFIED reference CPU codec.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/flrl_oracle.c header).  This file holds
 * no codec logic: it includes the reference's own header from where it lies
 * (-I/root/reference/src, see oracle/Makefile) and forwards to
 *   FixedLength::cpuCompress    /root/reference/src/fl/fl_cpu.cu:9-90
 *   FixedLength::cpuDecompress  /root/reference/src/fl/fl_cpu.cu:92-147
 * so that ctypes can call them.  Built into oracle/_ref/libflref.so.
 */
#include <cstdint>
#include <cstdlib>
#include <cstring>

#include "fl/fl_cpu.cuh"

extern "C" {

/* Returns 0 on success.  *bits / *values are malloc-owned by the reference and
 * must be released with flref_free (the reference's callers use free(),
 * src/main.cu:126-127). */
int flref_fl_cpu_compress(const uint8_t *data, size_t n, uint8_t **bits, size_t *bitsSize,
                          uint8_t **values, size_t *valuesSize)
{
    try {
        FixedLength::FLCompressed c = FixedLength::cpuCompress(const_cast<uint8_t *>(data), n);
        *bits = c.outputBits;
        *bitsSize = c.bitsSize;
        *values = c.outputValues;
        *valuesSize = c.valuesSize;
        return 0;
    } catch (...) {
        return 1;
    }
}

int flref_fl_cpu_decompress(size_t outputSize, const uint8_t *bits, size_t bitsSize,
                            const uint8_t *values, size_t valuesSize, uint8_t **out, size_t *outSize)
{
    try {
        FixedLength::FLDecompressed d = FixedLength::cpuDecompress(
            outputSize, const_cast<uint8_t *>(bits), bitsSize, const_cast<uint8_t *>(values), valuesSize);
        *out = d.data;
        *outSize = d.size;
        return 0;
    } catch (...) {
        return 1;
    }
}

void flref_free(void *p) { free(p); }

} /* extern "C" */
