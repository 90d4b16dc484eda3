// fl_kernels.cu — what the two fixed-length kernels share: the tile count and the workspace size.
//
// The fixed-length codec replaces the reference's four kernels + thrust scan + byte-atomic CAS loops
// (/root/reference/src/fl/fl_gpu.cu:648-808) with one kernel per direction: fl_encode.cu and fl_decode.cu.
// Both work on tiles of 128 frames (16 KiB of input).  Every full frame packs to 16*b bytes, so every tile's value
// range starts and ends on a 16-byte boundary and all bulk copies are legal (size % 16 == 0, 16-byte aligned); only the
// final partial frame (n % 128 != 0) goes through a small scalar path in the CTA that owns the last tile.
#include "fl_kernels.cuh"

#include <algorithm>

#include "flrl_bits.cuh"

namespace flrl {

static constexpr int TILE_FRAMES = 128;

size_t fl_tiles(size_t n)
{
    const size_t frames = (n + FRAME - 1) / FRAME;
    return (frames + TILE_FRAMES - 1) / TILE_FRAMES;
}

size_t fl_encode_workspace_bytes(size_t n);  // fl_encode.cu
// one workspace serves both directions: the decoder's status word at [tiles] (fl_decode_status), the first value unit of
// every tile and the pre-pass's look-back descriptors behind it, or the encoder's scanner arrays
size_t fl_workspace_bytes(size_t n)
{
    const size_t t = fl_tiles(n);
    return std::max((2 * t + t / 256 + 8) * sizeof(uint64_t) + 16, fl_encode_workspace_bytes(n));
}

}  // namespace flrl
