// abi_common.cuh — error plumbing shared by the extern "C" entry points.
#pragma once
#include <cuda_runtime.h>
#include <stdio.h>
#include <string>

#include "../../include/flrl.h"

namespace flrl {

std::string& last_error_slot();
int fail(int code, const char* fmt, ...);
int fail_cuda(cudaError_t e, const char* what);

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace flrl
