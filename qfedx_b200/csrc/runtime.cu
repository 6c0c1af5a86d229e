// Error plumbing and small runtime queries shared by every entry point.
#include "common.cuh"
#include "../../include/qfedx_b200.h"
#include <stdarg.h>
#include <string.h>

namespace qfx {
static thread_local char g_last_error[512] = "";
void set_last_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
    va_end(ap);
}
}  // namespace qfx

extern "C" const char* qfx_last_error(void) { return qfx::g_last_error; }
extern "C" int qfx_version(void) { return 100; }
