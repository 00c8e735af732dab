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
