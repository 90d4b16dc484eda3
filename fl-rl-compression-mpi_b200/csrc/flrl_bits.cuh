// flrl_bits.cuh — lane-level bit packing / unpacking for the fixed-length codec.
//
// Format being produced (reference: /root/reference/src/fl/fl_cpu.cu:64-84, decode :117-141):
// a frame is 128 input bytes; with frame width b (1..8) element j of the frame occupies
// stream bits [j*b, (j+1)*b), LSB first, so a full frame is exactly 16*b bytes = 4*b
// 32-bit words.  A lane that owns 32 consecutive elements therefore owns exactly b whole
// output words, at word offset b*t for quarter t of the frame: no cross-lane carries.
//
// Everything here is branch-free in b (the width differs per frame, i.e. per 4-lane group),
// and is written as __host__ __device__ so tests/host_sim can run the very same arithmetic
// on the CPU against the oracle.  The multiplies are deliberate: on sm_100a IMAD issues on
// the FMA pipe and LOP3/SHF/PRMT on the ALU pipe (each 0.5 warp-instr/clk/SMSP), so
// "x + (x & m) * (2^k - 1)" (one LOP3 + one IMAD) moves a field left by k bits for half the
// ALU-pipe cost of mask/shift/or.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define FLRL_HD __host__ __device__ __forceinline__
#else
#define FLRL_HD inline
#endif

namespace flrl {

static constexpr uint32_t FRAME = 128;  // src/fl/fl_common.cuh:9 FRAME_LENGTH

// ---- shifts with PTX clamp semantics (amount >= width gives 0) ------------------------
FLRL_HD uint32_t shl32c(uint32_t x, uint32_t s)
{
#if defined(__CUDA_ARCH__)
    return __funnelshift_lc(0u, x, s);
#else
    return s >= 32 ? 0u : x << s;
#endif
}
FLRL_HD uint32_t shr32c(uint32_t x, uint32_t s)
{
#if defined(__CUDA_ARCH__)
    return __funnelshift_rc(x, 0u, s);
#else
    return s >= 32 ? 0u : x >> s;
#endif
}
// low 32 bits of (hi:lo) >> min(s,32)
FLRL_HD uint32_t funnel_rc(uint32_t lo, uint32_t hi, uint32_t s)
{
#if defined(__CUDA_ARCH__)
    return __funnelshift_rc(lo, hi, s);
#else
    if (s >= 32) return hi;
    return s == 0 ? lo : (lo >> s) | (hi << (32 - s));
#endif
}
FLRL_HD uint64_t shl64c(uint64_t x, uint32_t s)
{
#if defined(__CUDA_ARCH__)
    uint64_t r;
    asm("shl.b64 %0, %1, %2;" : "=l"(r) : "l"(x), "r"(s));
    return r;
#else
    return s >= 64 ? 0ull : x << s;
#endif
}
FLRL_HD uint64_t shr64c(uint64_t x, uint32_t s)
{
#if defined(__CUDA_ARCH__)
    uint64_t r;
    asm("shr.u64 %0, %1, %2;" : "=l"(r) : "l"(x), "r"(s));
    return r;
#else
    return s >= 64 ? 0ull : x >> s;
#endif
}
// byte permute: result byte i = byte (sel nibble i) of the 8-byte pool {a = 0..3, b = 4..7}
FLRL_HD uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel)
{
#if defined(__CUDA_ARCH__)
    return __byte_perm(a, b, sel);
#else
    uint64_t pool = ((uint64_t)b << 32) | a;
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) r |= (uint32_t)((pool >> (8 * ((sel >> (4 * i)) & 7))) & 0xff) << (8 * i);
    return r;
#endif
}

// width of a frame from the OR of its bytes: max(1, 8 - clz8(or))   (fl_cpu.cu:41-48)
FLRL_HD uint32_t width_of_or(uint32_t or8)
{
#if defined(__CUDA_ARCH__)
    uint32_t w = 32u - (uint32_t)__clz((int)(or8 & 0xffu));
#else
    uint32_t w = 0, v = or8 & 0xffu;
    while (v) { ++w; v >>= 1; }
#endif
    return w ? w : 1u;
}

// Per-frame constants of the multiply-based packer; for b == 8 the multipliers are 0, 0, 1 and the shift 0 (identity).
struct PackK {
    uint32_t k1;  // 2^(8-b)   - 1 : bytes   -> 2b-bit pairs in 16-bit lanes (left aligned)
    uint32_t k2;  // 2^(16-2b) - 1 : pairs   -> 4b-bit quads in 32-bit words
    uint32_t k3;  // 2^(32-4b)     : quads   -> 8b-bit octets in 64-bit words
    uint32_t ra;  // 56 - 7b      : right-align shift of an octet
    uint32_t s8;  // 8b
};
FLRL_HD PackK make_pack_k(uint32_t b)
{
    PackK k;
    k.k1 = (1u << (8 - b)) - 1u;
    k.k2 = (1u << (16 - 2 * b)) - 1u;
    k.k3 = 1u << (32 - 4 * b);
    k.ra = 56 - 7 * b;
    k.s8 = 8 * b;
    return k;
}

// octet = c0 * k3 + (c1 << 32): one IMAD.WIDE.U32 and one add into its high word (no carry: the low word is the product's)
FLRL_HD uint64_t pack_octet(uint32_t c0, uint32_t c1, uint32_t k3)
{
    const uint64_t m = (uint64_t)c0 * k3;
    return (m & 0xffffffffull) | ((uint64_t)((uint32_t)(m >> 32) + c1) << 32);
}

// 16 input bytes (x[0..3], little endian) of width b -> 16b bits, right aligned in (lo,hi).
//   stage 1: t  = x + (x & 0x00ff00ff) * k1   -> pair (v0 + v1<<b) at bits [8-b, 8+b) of each half
//   stage 2: c  = t + (t & 0x0000ffff) * k2   -> quad at bits [24-3b, 24+b)
//   stage 3: W  = (c1 << 32) + c0 * k3        -> octet at bits [56-7b, 56+b)
//   stage 4: Q  = P0 | P1 << 8b               (P = W >> (56-7b))
FLRL_HD void pack16(uint32_t x0, uint32_t x1, uint32_t x2, uint32_t x3, const PackK& k, uint64_t& qlo, uint64_t& qhi)
{
    uint32_t t0 = x0 + (x0 & 0x00ff00ffu) * k.k1;
    uint32_t t1 = x1 + (x1 & 0x00ff00ffu) * k.k1;
    uint32_t t2 = x2 + (x2 & 0x00ff00ffu) * k.k1;
    uint32_t t3 = x3 + (x3 & 0x00ff00ffu) * k.k1;
    uint32_t c0 = t0 + (t0 & 0xffffu) * k.k2;
    uint32_t c1 = t1 + (t1 & 0xffffu) * k.k2;
    uint32_t c2 = t2 + (t2 & 0xffffu) * k.k2;
    uint32_t c3 = t3 + (t3 & 0xffffu) * k.k2;
    uint64_t p0 = pack_octet(c0, c1, k.k3) >> k.ra;
    uint64_t p1 = pack_octet(c2, c3, k.k3) >> k.ra;
    qlo = p0 | shl64c(p1, k.s8);
    qhi = p1 >> (64 - k.s8);
}

// 32 input bytes (two uint4: a = elements 0..15, c = 16..31 of this lane) -> b output words
// handed to st(pred, j, word) for j in [0, b): word j of the lane (frame_word_base + b*t + j).
// Q0 (first 16b bits) supplies words [0, b/2); when b is odd the word at b/2 is shared: its low
// half is the top 16 bits of Q0 and Q1 moves up 16 bits — both done by PRMTs whose selector
// depends on the parity of b.  All eight stores are issued predicated, no branches.
template <typename Store>
FLRL_HD void pack32_store(const uint32_t a[4], const uint32_t c[4], uint32_t b, const PackK& k, Store&& st)
{
    uint64_t q0lo, q0hi, q1lo, q1hi;
    pack16(a[0], a[1], a[2], a[3], k, q0lo, q0hi);
    pack16(c[0], c[1], c[2], c[3], k, q1lo, q1hi);
    const uint32_t A0 = (uint32_t)q0lo, A1 = (uint32_t)(q0lo >> 32), A2 = (uint32_t)q0hi, A3 = (uint32_t)(q0hi >> 32);
    const uint32_t C0 = (uint32_t)q1lo, C1 = (uint32_t)(q1lo >> 32), C2 = (uint32_t)q1hi, C3 = (uint32_t)(q1hi >> 32);
    const uint32_t q = b >> 1;
    const bool odd = (b & 1u) != 0;
    const uint32_t sel = odd ? 0x5432u : 0x7654u;
    const uint32_t sel0 = odd ? 0x5410u : 0x7654u;  // even b: F0 = C0, the pick below is ignored
    const uint32_t a01 = (q & 1u) ? A1 : A0;
    const uint32_t a23 = (q & 1u) ? A3 : A2;
    const uint32_t aq = (q & 2u) ? a23 : a01;       // word floor(b/2) of Q0 (b odd, q <= 3)
    const uint32_t F0 = prmt(aq, C0, sel0);
    const uint32_t F1 = prmt(C0, C1, sel);
    const uint32_t F2 = prmt(C1, C2, sel);
    const uint32_t F3 = prmt(C2, C3, sel);
    const uint32_t nf = b - q;  // ceil(b/2) words come from F
    st(0 < q, 0u, A0);
    st(1 < q, 1u, A1);
    st(2 < q, 2u, A2);
    st(3 < q, 3u, A3);
    st(b != 0, q, F0);  // b == 0 marks "no frame here" (tile tail): nothing is stored
    st(1 < nf, q + 1, F1);
    st(2 < nf, q + 2, F2);
    st(3 < nf, q + 3, F3);
}

// ---- decode -----------------------------------------------------------------------------
struct UnpackK {
    uint32_t m4;   // mask of 4b bits
    uint32_t mb;   // (2^b - 1) * 0x00010001
    uint32_t f2;   // 2^16 - 2^(2b)
    uint32_t f1;   // 2^8  - 2^b
    uint32_t s8, s4, s2, s1;
};
FLRL_HD UnpackK make_unpack_k(uint32_t b)
{
    UnpackK k;
    k.m4 = 0xffffffffu >> (32 - 4 * b);
    k.mb = ((1u << b) - 1u) * 0x00010001u;
    k.f2 = 65536u - (1u << (2 * b));
    k.f1 = 256u - (1u << b);
    k.s8 = 8 * b; k.s4 = 4 * b; k.s2 = 2 * b; k.s1 = b;
    return k;
}
// 4b-bit quad -> 4 bytes
FLRL_HD uint32_t unpack_quad(uint32_t c, const UnpackK& k)
{
    uint32_t t = c + (c >> k.s2) * k.f2;                 // pairs into 16-bit lanes
    return t + ((t >> k.s1) & k.mb) * k.f1;              // elements into bytes
}
// 16b bits (lo,hi; garbage above bit 16b allowed) -> 16 bytes
FLRL_HD void unpack16(uint64_t lo, uint64_t hi, const UnpackK& k, uint32_t x[4])
{
    uint64_t p1 = shr64c(lo, k.s8) | shl64c(hi, 64 - k.s8);
    uint32_t c0 = (uint32_t)lo & k.m4;
    uint32_t c1 = funnel_rc((uint32_t)lo, (uint32_t)(lo >> 32), k.s4) & k.m4;
    uint32_t c2 = (uint32_t)p1 & k.m4;
    uint32_t c3 = funnel_rc((uint32_t)p1, (uint32_t)(p1 >> 32), k.s4) & k.m4;
    x[0] = unpack_quad(c0, k);
    x[1] = unpack_quad(c1, k);
    x[2] = unpack_quad(c2, k);
    x[3] = unpack_quad(c3, k);
}
// The lane's words, already in registers: r0..r3 = words 0..3, g0..g4 = words q..q+4 (q = b/2) -> 32 bytes.
FLRL_HD void unpack32_regs(uint32_t b, const UnpackK& k, const uint32_t r[4], const uint32_t g[5], uint32_t a[4], uint32_t c[4])
{
    const uint32_t sel = (b & 1u) ? 0x5432u : 0x3210u;
    unpack16(((uint64_t)r[1] << 32) | r[0], ((uint64_t)r[3] << 32) | r[2], k, a);
    uint32_t h0 = prmt(g[0], g[1], sel), h1 = prmt(g[1], g[2], sel), h2 = prmt(g[2], g[3], sel), h3 = prmt(g[3], g[4], sel);
    unpack16(((uint64_t)h1 << 32) | h0, ((uint64_t)h3 << 32) | h2, k, c);
}
// The lane's b words are read through `ld(j)` (j relative to the lane's first word; the
// callee must return 0 or anything for j >= b, never fault).  Produces 32 bytes.
template <typename Load>
FLRL_HD void unpack32_load(uint32_t b, const UnpackK& k, Load&& ld, uint32_t a[4], uint32_t c[4])
{
    const uint32_t q = b >> 1;
    const uint32_t r[4] = {ld(0u), ld(1u), ld(2u), ld(3u)};
    const uint32_t g[5] = {ld(q), ld(q + 1), ld(q + 2), ld(q + 3), ld(q + 4)};
    unpack32_regs(b, k, r, g, a, c);
}

}  // namespace flrl
