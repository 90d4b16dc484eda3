// flrl_abi.cu — DEVICE layer of include/flrl.h: argument checks, then the kernel launchers.
#include <stdarg.h>

#include "abi_common.cuh"
#include "fl_kernels.cuh"
#include "rl_kernels.cuh"

namespace flrl {

std::string& last_error_slot()
{
    static thread_local std::string s;
    return s;
}
int fail(int code, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    last_error_slot() = buf;
    return code;
}
// Same wording as the reference's CHECK_CUDA (/root/reference/src/utils.cuh:6-16).
int fail_cuda(cudaError_t e, const char* what)
{
    return fail((int)e, "CUDA Error: %s in %s", cudaGetErrorString(e), what);
}

}  // namespace flrl

using namespace flrl;

extern "C" {

const char* flrl_last_error(void) { return last_error_slot().c_str(); }
int flrl_version(void) { return 100; }

size_t flrl_fl_frames(size_t n) { return (n + FLRL_FRAME_LENGTH - 1) / FLRL_FRAME_LENGTH; }
size_t flrl_fl_workspace_bytes(size_t n) { return fl_workspace_bytes(n); }

int flrl_fl_encode(const uint8_t* d_in, size_t n, uint8_t* d_bits, uint8_t* d_values, size_t values_cap,
                   uint64_t* d_values_size, void* d_ws, size_t ws_bytes, flrl_stream_t stream)
{
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (!d_values_size) return fail(FLRL_E_ARG, "flrl_fl_encode: d_values_size is NULL");
    if (n == 0) {
        cudaError_t e = cudaMemsetAsync(d_values_size, 0, sizeof(uint64_t), s);
        return e == cudaSuccess ? FLRL_OK : fail_cuda(e, "flrl_fl_encode/memset");
    }
    if (!d_in || !d_bits || !d_values || !d_ws) return fail(FLRL_E_ARG, "flrl_fl_encode: NULL buffer");
    if (!aligned16(d_in) || !aligned16(d_values) || !aligned16(d_ws))
        return fail(FLRL_E_ALIGN, "flrl_fl_encode: d_in, d_values and d_ws must be 16-byte aligned");
    if (values_cap < ((n + 15) & ~(size_t)15)) return fail(FLRL_E_CAPACITY, "flrl_fl_encode: values_cap < round_up(n, 16)");
    if (ws_bytes < fl_workspace_bytes(n)) return fail(FLRL_E_WORKSPACE, "flrl_fl_encode: workspace too small");
    cudaError_t e = fl_encode_launch(d_in, n, d_bits, d_values, d_values_size, d_ws, s);
    return e == cudaSuccess ? FLRL_OK : fail_cuda(e, "flrl_fl_encode");
}

int flrl_fl_decode(const uint8_t* d_bits, size_t nbits, const uint8_t* d_values, size_t nvalues, uint8_t* d_out, size_t n,
                   int* d_status, void* d_ws, size_t ws_bytes, flrl_stream_t stream)
{
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (n == 0) {
        if (d_status) {
            cudaError_t e = cudaMemsetAsync(d_status, 0, sizeof(int), s);
            if (e != cudaSuccess) return fail_cuda(e, "flrl_fl_decode/memset");
        }
        return FLRL_OK;
    }
    if (!d_bits || !d_values || !d_out || !d_ws) return fail(FLRL_E_ARG, "flrl_fl_decode: NULL buffer");
    if (nbits != flrl_fl_frames(n)) return fail(FLRL_E_FORMAT, "flrl_fl_decode: bitsSize != ceil(outputSize/128)");
    if (!aligned16(d_values) || !aligned16(d_out) || !aligned16(d_ws))
        return fail(FLRL_E_ALIGN, "flrl_fl_decode: d_values, d_out and d_ws must be 16-byte aligned");
    if (ws_bytes < fl_workspace_bytes(n)) return fail(FLRL_E_WORKSPACE, "flrl_fl_decode: workspace too small");
    cudaError_t e = fl_decode_launch(d_bits, nbits, d_values, nvalues, d_out, n, d_ws, s);
    if (e != cudaSuccess) return fail_cuda(e, "flrl_fl_decode");
    if (d_status) {
        e = cudaMemcpyAsync(d_status, fl_decode_status(d_ws, n), sizeof(int), cudaMemcpyDeviceToDevice, s);
        if (e != cudaSuccess) return fail_cuda(e, "flrl_fl_decode/status");
    }
    return FLRL_OK;
}

size_t flrl_rl_workspace_bytes(size_t n) { return rl_encode_workspace_bytes(n) + 16; }
size_t flrl_rl_decode_workspace_bytes(size_t pairs) { return rl_decode_workspace_bytes(pairs) + 16; }

int flrl_rl_encode(const uint8_t* d_in, size_t n, size_t n_avail, uint64_t prev_run, uint8_t* d_counts, uint8_t* d_values,
                   size_t pair_cap, uint64_t* d_pairs, void* d_ws, size_t ws_bytes, flrl_stream_t stream)
{
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (!d_pairs) return fail(FLRL_E_ARG, "flrl_rl_encode: d_pairs is NULL");
    if (n == 0) {
        cudaError_t e = cudaMemsetAsync(d_pairs, 0, sizeof(uint64_t), s);
        return e == cudaSuccess ? FLRL_OK : fail_cuda(e, "flrl_rl_encode/memset");
    }
    if (!d_in || !d_counts || !d_values || !d_ws) return fail(FLRL_E_ARG, "flrl_rl_encode: NULL buffer");
    if (n_avail < n) return fail(FLRL_E_ARG, "flrl_rl_encode: n_avail < n");
    if (!aligned16(d_in) || !aligned16(d_counts) || !aligned16(d_values) || !aligned16(d_ws))
        return fail(FLRL_E_ALIGN, "flrl_rl_encode: d_in, d_counts, d_values and d_ws must be 16-byte aligned");
    if (ws_bytes < rl_encode_workspace_bytes(n)) return fail(FLRL_E_WORKSPACE, "flrl_rl_encode: workspace too small");
    cudaError_t e = rl_encode_launch(d_in, n, n_avail, prev_run, d_counts, d_values, pair_cap, d_pairs, d_ws, s);
    return e == cudaSuccess ? FLRL_OK : fail_cuda(e, "flrl_rl_encode");
}

int flrl_rl_decode(const uint8_t* d_counts, const uint8_t* d_values, size_t pairs, uint8_t* d_out, size_t out_cap,
                   uint64_t* d_out_size, int* d_status, void* d_ws, size_t ws_bytes, flrl_stream_t stream)
{
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (!d_out_size) return fail(FLRL_E_ARG, "flrl_rl_decode: d_out_size is NULL");
    if (pairs == 0) {
        cudaError_t e = cudaMemsetAsync(d_out_size, 0, sizeof(uint64_t), s);
        if (e == cudaSuccess && d_status) e = cudaMemsetAsync(d_status, 0, sizeof(int), s);
        return e == cudaSuccess ? FLRL_OK : fail_cuda(e, "flrl_rl_decode/memset");
    }
    if (!d_counts || !d_values || !d_ws || (!d_out && out_cap)) return fail(FLRL_E_ARG, "flrl_rl_decode: NULL buffer");
    if (!aligned16(d_counts) || !aligned16(d_values) || !aligned16(d_ws))
        return fail(FLRL_E_ALIGN, "flrl_rl_decode: d_counts, d_values and d_ws must be 16-byte aligned");
    if (ws_bytes < rl_decode_workspace_bytes(pairs)) return fail(FLRL_E_WORKSPACE, "flrl_rl_decode: workspace too small");
    cudaError_t e = rl_decode_launch(d_counts, d_values, pairs, d_out, out_cap, d_out_size, d_ws, s);
    if (e != cudaSuccess) return fail_cuda(e, "flrl_rl_decode");
    if (d_status) {
        e = cudaMemcpyAsync(d_status, rl_decode_status(d_ws, pairs), sizeof(int), cudaMemcpyDeviceToDevice, s);
        if (e != cudaSuccess) return fail_cuda(e, "flrl_rl_decode/status");
    }
    return FLRL_OK;
}

}  // extern "C"
