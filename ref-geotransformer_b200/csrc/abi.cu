// Library-wide pieces of the C ABI: error text, version.
#include "host_util.h"
#include <string.h>

namespace geotx {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace geotx

extern "C" const char* geotx_last_error(void) { return geotx::g_err; }
extern "C" int geotx_abi_version(void) { return 1; }

// Device-to-device copy by the copy engines, on the caller's stream, between two pointers that are valid in THIS
// process's address space: local memory or a peer GPU's memory mapped through CUDA IPC (dist.PeerGather).  A plain
// cudaMemcpyAsync: torch's own cross-device copy_ would create a context on the peer GPU first.
extern "C" int geotx_copy_async(void* dst, const void* src, int64_t bytes, void* stream) {
    GEOTX_REQUIRE(dst && src && bytes > 0, "geotx_copy_async: bad arguments");
    int rc = geotx::check_cuda(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDefault, (cudaStream_t)stream),
                               "geotx_copy_async");
    return rc;
}
