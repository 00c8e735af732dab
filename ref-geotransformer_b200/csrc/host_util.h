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

// SM count of the current device (148 on B200), asked once per device: every persistent grid is sized from it.
inline int sm_count() {
    static int cached[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int sms = 0;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) return 148;
        cached[dev] = sms;
    }
    return cached[dev];
}
}  // namespace geotx

#define GEOTX_REQUIRE(cond, ...)              \
    do {                                      \
        if (!(cond)) {                        \
            geotx::set_error(__VA_ARGS__);    \
            return GEOTX_E_ARG;               \
        }                                     \
    } while (0)
