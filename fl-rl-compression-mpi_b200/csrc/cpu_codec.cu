// cpu_codec.cu — the CLI's CPU methods (`fl-cpu`, `rl-cpu`).
//
// These are the product's counterparts of FixedLength::cpuCompress / cpuDecompress
// (/root/reference/src/fl/fl_cpu.cu:9-147) and of the run-length spec
// (IMPLEMENTATION-PLAN.md:81-179).  They are CPU methods by definition; the GPU methods
// (`fl`, `rl`) never route through them.  The FL packer reuses the word-parallel lane
// arithmetic of flrl_bits.cuh (16 elements per step) instead of the reference's per-bit loop.
#include <stdlib.h>
#include <string.h>

#include "abi_common.cuh"
#include "flrl_bits.cuh"

using namespace flrl;

namespace {

inline uint32_t load32(const uint8_t* p)
{
    uint32_t v;
    memcpy(&v, p, 4);
    return v;
}

// width of the frame in[0..len): max(1, 8 - clz8(OR of the bytes))   (fl_cpu.cu:37-50)
uint32_t frame_width(const uint8_t* in, size_t len)
{
    uint64_t acc = 0;
    size_t i = 0;
    for (; i + 8 <= len; i += 8) {
        uint64_t v;
        memcpy(&v, in + i, 8);
        acc |= v;
    }
    uint32_t o = (uint32_t)(acc | (acc >> 32));
    o |= o >> 16;
    o |= o >> 8;
    for (; i < len; ++i) o |= in[i];
    return width_of_or(o);
}

// full frame: 8 groups of 16 elements, each 2*b bytes (fl_cpu.cu:64-84, LSB first)
void pack_frame(const uint8_t* in, uint32_t b, uint8_t* out)
{
    const PackK k = make_pack_k(b);
    for (int g = 0; g < 8; ++g) {
        uint64_t lo, hi;
        pack16(load32(in + 16 * g), load32(in + 16 * g + 4), load32(in + 16 * g + 8), load32(in + 16 * g + 12), k, lo, hi);
        uint8_t tmp[16];
        memcpy(tmp, &lo, 8);
        memcpy(tmp + 8, &hi, 8);
        memcpy(out + 2 * b * g, tmp, 2 * b);
    }
}

void unpack_frame(const uint8_t* in, uint32_t b, uint8_t* out)
{
    const UnpackK k = make_unpack_k(b);
    for (int g = 0; g < 8; ++g) {
        uint8_t tmp[16] = {0};
        memcpy(tmp, in + 2 * b * g, 2 * b);
        uint64_t lo, hi;
        memcpy(&lo, tmp, 8);
        memcpy(&hi, tmp + 8, 8);
        uint32_t x[4];
        unpack16(lo, hi, k, x);
        memcpy(out + 16 * g, x, 16);
    }
}

// partial last frame: plain bit loop
size_t pack_tail(const uint8_t* in, size_t len, uint32_t b, uint8_t* out)
{
    const size_t nbytes = (len * b + 7) / 8;
    memset(out, 0, nbytes);
    for (size_t i = 0; i < len; ++i) {
        const size_t bit = i * b;
        const uint32_t v = (uint32_t)in[i] << (bit & 7);
        out[bit >> 3] |= (uint8_t)v;
        if ((bit & 7) + b > 8) out[(bit >> 3) + 1] |= (uint8_t)(v >> 8);
    }
    return nbytes;
}

}  // namespace

extern "C" {

void flrl_free(void* p) { free(p); }

int flrl_fl_cpu_compress(const uint8_t* data, size_t size, uint8_t** bits, size_t* bits_size, uint8_t** values,
                         size_t* values_size)
{
    if (!bits || !bits_size || !values || !values_size) return fail(FLRL_E_ARG, "flrl_fl_cpu_compress: NULL output");
    *bits = *values = nullptr;
    *bits_size = *values_size = 0;
    if (size == 0) return FLRL_OK;  // fl_cpu.cu:11-14
    if (!data) return fail(FLRL_E_ARG, "flrl_fl_cpu_compress: NULL input");
    const size_t frames = (size + FRAME - 1) / FRAME;
    uint8_t* b = static_cast<uint8_t*>(malloc(frames));
    uint8_t* v = static_cast<uint8_t*>(malloc(size + 16));
    if (!b || !v) {
        free(b);
        free(v);
        return fail(FLRL_E_NOMEM, "flrl_fl_cpu_compress: out of memory");
    }
    size_t o = 0;
    for (size_t f = 0; f < frames; ++f) {
        const size_t len = size - f * FRAME < FRAME ? size - f * FRAME : FRAME;
        const uint32_t w = frame_width(data + f * FRAME, len);
        b[f] = (uint8_t)w;
        if (len == FRAME) {
            pack_frame(data + f * FRAME, w, v + o);
            o += 16 * w;
        } else {
            o += pack_tail(data + f * FRAME, len, w, v + o);
        }
    }
    uint8_t* shrunk = static_cast<uint8_t*>(realloc(v, o ? o : 1));
    *bits = b;
    *bits_size = frames;
    *values = shrunk ? shrunk : v;
    *values_size = o;
    return FLRL_OK;
}

int flrl_fl_cpu_decompress(size_t output_size, const uint8_t* bits, size_t bits_size, const uint8_t* values,
                           size_t values_size, uint8_t** out, size_t* out_size)
{
    if (!out || !out_size) return fail(FLRL_E_ARG, "flrl_fl_cpu_decompress: NULL output");
    *out = nullptr;
    *out_size = 0;
    if (output_size == 0 || bits_size == 0 || values_size == 0) return FLRL_OK;  // fl_cpu.cu:99-103
    if (!bits || !values) return fail(FLRL_E_ARG, "flrl_fl_cpu_decompress: NULL input");
    const size_t frames = (output_size + FRAME - 1) / FRAME;
    if (bits_size != frames) return fail(FLRL_E_FORMAT, "flrl_fl_cpu_decompress: bitsSize != ceil(outputSize/128)");
    uint8_t* o = static_cast<uint8_t*>(malloc(output_size));
    if (!o) return fail(FLRL_E_NOMEM, "flrl_fl_cpu_decompress: out of memory");
    size_t at = 0;
    for (size_t f = 0; f < frames; ++f) {
        const size_t len = output_size - f * FRAME < FRAME ? output_size - f * FRAME : FRAME;
        const uint32_t w = bits[f];
        const size_t need = (len * w + 7) / 8;
        if (w < 1 || w > 8 || at + need > values_size) {
            free(o);
            return fail(FLRL_E_FORMAT, "flrl_fl_cpu_decompress: malformed container at frame %zu", f);
        }
        if (len == FRAME) {
            unpack_frame(values + at, w, o + f * FRAME);
        } else {
            for (size_t i = 0; i < len; ++i) {
                const size_t bit = i * w;
                uint32_t x = values[at + (bit >> 3)];
                if ((bit >> 3) + 1 < need) x |= (uint32_t)values[at + (bit >> 3) + 1] << 8;
                o[f * FRAME + i] = (uint8_t)((x >> (bit & 7)) & ((1u << w) - 1u));
            }
        }
        at += need;
    }
    *out = o;
    *out_size = output_size;
    return FLRL_OK;
}

int flrl_rl_cpu_compress(const uint8_t* data, size_t size, uint8_t** counts, uint8_t** values, size_t* pairs)
{
    if (!counts || !values || !pairs) return fail(FLRL_E_ARG, "flrl_rl_cpu_compress: NULL output");
    *counts = *values = nullptr;
    *pairs = 0;
    if (size == 0) return FLRL_OK;
    if (!data) return fail(FLRL_E_ARG, "flrl_rl_cpu_compress: NULL input");
    // pass 1: pair count (a run of length L yields ceil(L/255) pairs, IMPLEMENTATION-PLAN.md:125)
    size_t p = 0;
    for (size_t i = 0; i < size;) {
        const uint8_t* e = data + i + 1;
        const uint8_t* end = data + size;
        while (e < end && *e == data[i]) ++e;
        const size_t len = (size_t)(e - (data + i));
        p += (len + 254) / 255;
        i += len;
    }
    uint8_t* c = static_cast<uint8_t*>(malloc(p));
    uint8_t* v = static_cast<uint8_t*>(malloc(p));
    if (!c || !v) {
        free(c);
        free(v);
        return fail(FLRL_E_NOMEM, "flrl_rl_cpu_compress: out of memory");
    }
    size_t k = 0;
    for (size_t i = 0; i < size;) {
        size_t j = i + 1;
        while (j < size && data[j] == data[i]) ++j;
        size_t len = j - i;
        const size_t full = len / 255, rem = len % 255;
        memset(c + k, 255, full);
        memset(v + k, data[i], full + (rem ? 1 : 0));
        k += full;
        if (rem) c[k++] = (uint8_t)rem;
        i = j;
    }
    *counts = c;
    *values = v;
    *pairs = p;
    return FLRL_OK;
}

int flrl_rl_cpu_decompress(const uint8_t* counts, const uint8_t* values, size_t pairs, uint8_t** out, size_t* out_size)
{
    if (!out || !out_size) return fail(FLRL_E_ARG, "flrl_rl_cpu_decompress: NULL output");
    *out = nullptr;
    *out_size = 0;
    if (pairs == 0) return FLRL_OK;
    if (!counts || !values) return fail(FLRL_E_ARG, "flrl_rl_cpu_decompress: NULL input");
    size_t n = 0;
    for (size_t k = 0; k < pairs; ++k) {
        if (counts[k] == 0) return fail(FLRL_E_FORMAT, "flrl_rl_cpu_decompress: count 0 at pair %zu", k);
        n += counts[k];
    }
    uint8_t* o = static_cast<uint8_t*>(malloc(n));
    if (!o) return fail(FLRL_E_NOMEM, "flrl_rl_cpu_decompress: out of memory");
    size_t at = 0;
    for (size_t k = 0; k < pairs; ++k) {
        memset(o + at, values[k], counts[k]);
        at += counts[k];
    }
    *out = o;
    *out_size = n;
    return FLRL_OK;
}

}  // extern "C"
