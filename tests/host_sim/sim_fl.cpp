// tests/host_sim/sim_fl.cpp — TEST INFRASTRUCTURE ONLY.
// Runs the lane-level arithmetic of csrc/flrl_bits.cuh (the same header the CUDA kernels
// include) on the CPU, one emulated lane at a time, so the bit math can be checked against
// the oracle without a GPU.  It is not a CPU code path of the product: nothing outside
// tests/ builds or loads it.
#include <cstring>
#include <vector>
#include "../../fl-rl-compression-mpi_b200/csrc/flrl_bits.cuh"

using namespace flrl;

extern "C" {

// full frames only (n % 128 == 0).  Returns value bytes written.
uint64_t sim_fl_encode(const uint8_t* in, size_t n, uint8_t* bits, uint8_t* values)
{
    size_t frames = n / FRAME;
    uint64_t off16 = 0;  // running output offset in 16-byte units
    for (size_t f = 0; f < frames; ++f) {
        uint32_t o = 0;
        for (int i = 0; i < 128; ++i) o |= in[f * 128 + i];
        uint32_t b = width_of_or(o);
        bits[f] = (uint8_t)b;
        PackK k = make_pack_k(b);
        uint32_t* wbase = reinterpret_cast<uint32_t*>(values + off16 * 16);
        for (uint32_t t = 0; t < 4; ++t) {
            uint32_t a[4], c[4];
            memcpy(a, in + f * 128 + t * 32, 16);
            memcpy(c, in + f * 128 + t * 32 + 16, 16);
            uint32_t* w = wbase + b * t;
            pack32_store(a, c, b, k, [&](bool p, uint32_t j, uint32_t v) { if (p) w[j] = v; });
        }
        off16 += b;
    }
    return off16 * 16;
}

uint64_t sim_fl_decode(const uint8_t* bits, size_t frames, const uint8_t* values, size_t nvalues, uint8_t* out)
{
    std::vector<uint8_t> padded(nvalues + 64, 0xAB);  // garbage past the end must not matter
    memcpy(padded.data(), values, nvalues);
    uint64_t off16 = 0;
    for (size_t f = 0; f < frames; ++f) {
        uint32_t b = bits[f];
        UnpackK k = make_unpack_k(b);
        const uint32_t* wbase = reinterpret_cast<const uint32_t*>(padded.data() + off16 * 16);
        for (uint32_t t = 0; t < 4; ++t) {
            const uint32_t* w = wbase + b * t;
            uint32_t a[4], c[4];
            unpack32_load(b, k, [&](uint32_t j) { uint32_t v; memcpy(&v, w + j, 4); return v; }, a, c);
            memcpy(out + f * 128 + t * 32, a, 16);
            memcpy(out + f * 128 + t * 32 + 16, c, 16);
        }
        off16 += b;
    }
    return frames * 128;
}
}
