/*
 * flrl_oracle.c — CPU restatement of the reference's FL / RL codecs.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing on the product path (the CUDA library, the
 * `compress` CLI's `fl`/`rl` methods, bench.py's measured arm) may call into
 * this file.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs use it, and only as the checker / the timed baseline.
 *
 * Parity status
 *   FL:  PINNED.  tests/test_oracle_golden.py checks this file against every
 *        golden vector of SURVEY.md §8c (produced by the reference binary) and,
 *        when oracle/_ref/libflref.so exists, against the reference's own
 *        cpuCompress/cpuDecompress compiled from /root/reference/src.
 *   RL:  the reference has NO run-length code (src/args_parser.cu:29-52 has no
 *        `rl` method).  The pair stream follows IMPLEMENTATION-PLAN.md:81-179 and
 *        is pinned by that document's three worked examples (:87-89, :125,
 *        :158-160); the on-disk RL layout is a build decision — "parity
 *        unpinned" for the container, pinned for the (counts, values) stream.
 *
 * Each function cites the reference lines it restates.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#define FLRL_FRAME 128u /* src/fl/fl_common.cuh:9  FRAME_LENGTH */

/* src/fl/fl_common.cuh:198-212 countLeadingZeroes: width of a byte = 8 - clz8,
 * 0 for value 0 (the callers clamp the frame width to >= 1). */
static unsigned byte_width(uint8_t v)
{
    unsigned w = 0;
    while (v) {
        ++w;
        v >>= 1;
    }
    return w;
}

/* src/fl/fl_cpu.cu:21 — framesCount = ceil(size / FRAME_LENGTH). */
size_t flrl_oracle_fl_frames(size_t n) { return (n + FLRL_FRAME - 1) / FLRL_FRAME; }

/*
 * src/fl/fl_cpu.cu:9-90 cpuCompress.
 *   bits[f]  = max(1, max_i(8 - clz8(x_i)))                     (:37-48)
 *   values   = LSB-first concatenation of bits[f]-wide fields   (:64-84)
 *   size     = ceil(sum_f bits[f] * len_f / 8)                  (:49,:53)
 * `values` must have room for n bytes (all-8-bit worst case); returns the
 * number of value bytes written.  n == 0 -> 0 (:11-14).
 */
uint64_t flrl_oracle_fl_compress(const uint8_t *in, size_t n, uint8_t *bits, uint8_t *values)
{
    if (n == 0) return 0;
    const size_t frames = flrl_oracle_fl_frames(n);
    uint64_t acc = 0;  /* pending bits, LSB = next output bit */
    unsigned fill = 0; /* number of valid bits in acc (< 8 between elements) */
    size_t o = 0;
    for (size_t f = 0; f < frames; ++f) {
        const size_t lo = f * FLRL_FRAME;
        const size_t hi = (lo + FLRL_FRAME < n) ? lo + FLRL_FRAME : n;
        uint8_t any = 0;
        for (size_t i = lo; i < hi; ++i) any |= in[i];
        unsigned w = byte_width(any); /* max of widths == width of the OR */
        if (w == 0) w = 1;
        bits[f] = (uint8_t)w;
        for (size_t i = lo; i < hi; ++i) {
            acc |= (uint64_t)in[i] << fill;
            fill += w;
            while (fill >= 8) {
                values[o++] = (uint8_t)acc;
                acc >>= 8;
                fill -= 8;
            }
        }
    }
    if (fill) values[o++] = (uint8_t)acc; /* zero padding bits, :53-55 */
    return o;
}

/*
 * src/fl/fl_cpu.cu:92-147 cpuDecompress.  Returns the number of bytes produced:
 * 0 when bitsSize == 0 || valuesSize == 0 (:94-97), else n.
 */
uint64_t flrl_oracle_fl_decompress(const uint8_t *bits, size_t nbits, const uint8_t *values,
                                   size_t nvalues, uint8_t *out, size_t n)
{
    if (nbits == 0 || nvalues == 0) return 0;
    uint64_t bitpos = 0;
    for (size_t f = 0; f < nbits; ++f) {
        const unsigned w = bits[f];
        const unsigned mask = (1u << w) - 1u;
        for (size_t i = 0; i < FLRL_FRAME && f * FLRL_FRAME + i < n; ++i) {
            const size_t byte = (size_t)(bitpos >> 3);
            const unsigned sh = (unsigned)(bitpos & 7);
            unsigned v = values[byte] >> sh;
            if (sh + w > 8) v |= (unsigned)values[byte + 1] << (8 - sh); /* :130-136 */
            out[f * FLRL_FRAME + i] = (uint8_t)(v & mask);
            bitpos += w;
        }
    }
    return n;
}

/*
 * IMPLEMENTATION-PLAN.md:83-152 run-length encode.  A pair starts where the
 * byte differs from its predecessor (:95-102) and every 255 bytes inside a
 * longer run (:125 "256 -> 255 + 1", greedy).  counts[k] in 1..255.
 * `counts` / `values` need room for n entries.  Returns the pair count.
 */
uint64_t flrl_oracle_rl_compress(const uint8_t *in, size_t n, uint8_t *counts, uint8_t *values)
{
    uint64_t p = 0;
    size_t i = 0;
    while (i < n) {
        const uint8_t v = in[i];
        size_t j = i + 1;
        while (j < n && in[j] == v) ++j;
        size_t len = j - i;
        while (len > 255) {
            counts[p] = 255;
            values[p] = v;
            ++p;
            len -= 255;
        }
        counts[p] = (uint8_t)len;
        values[p] = v;
        ++p;
        i = j;
    }
    return p;
}

/* Number of pairs flrl_oracle_rl_compress would emit (sizing helper). */
uint64_t flrl_oracle_rl_count_pairs(const uint8_t *in, size_t n)
{
    uint64_t p = 0;
    size_t i = 0;
    while (i < n) {
        size_t j = i + 1;
        while (j < n && in[j] == in[i]) ++j;
        p += (j - i + 254) / 255;
        i = j;
    }
    return p;
}

/*
 * IMPLEMENTATION-PLAN.md:154-179 run-length decode: out[i] = values[j] for
 * start[j] <= i < start[j+1], start = exclusive scan of counts; output length
 * = sum of counts (:171).  Returns the bytes produced, or UINT64_MAX when the
 * sum exceeds `cap` (nothing past cap is written).
 */
uint64_t flrl_oracle_rl_decompress(const uint8_t *counts, const uint8_t *values, size_t pairs,
                                   uint8_t *out, size_t cap)
{
    uint64_t o = 0;
    for (size_t k = 0; k < pairs; ++k) {
        const unsigned c = counts[k];
        if (o + c > cap) return UINT64_MAX;
        memset(out + o, values[k], c);
        o += c;
    }
    return o;
}

/* Sum of counts (decoded length) without producing output. */
uint64_t flrl_oracle_rl_decoded_size(const uint8_t *counts, size_t pairs)
{
    uint64_t s = 0;
    for (size_t k = 0; k < pairs; ++k) s += counts[k];
    return s;
}
