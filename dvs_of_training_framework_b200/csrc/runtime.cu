// runtime.cu -- error string + version for the C ABI (include/dvsof.h).
#include <stdarg.h>
#include "common.cuh"

namespace dvsof {
static thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
}  // namespace dvsof

extern "C" const char* dvsof_last_error(void) { return dvsof::g_err; }
extern "C" int dvsof_version(void) { return 100; }
