// Host-side helpers shared by the C-ABI launchers.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include "../../include/geotx_b200.h"

namespace geotx {
void set_error(const char* fmt, ...);  // defined in abi.cu (thread-local buffer)

inline int check_cuda(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return GEOTX_OK;
    set_error("%s: %s", what, cudaGetErrorString(e));
    return GEOTX_E_CUDA;
}
inline int launch_status(const char* what) { return check_cuda(cudaGetLastError(), what); }
}  // namespace geotx

#define GEOTX_REQUIRE(cond, ...)              \
    do {                                      \
        if (!(cond)) {                        \
            geotx::set_error(__VA_ARGS__);    \
            return GEOTX_E_ARG;               \
        }                                     \
    } while (0)
