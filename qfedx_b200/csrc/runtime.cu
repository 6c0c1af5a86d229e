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

__global__ void feistel_kernel(int n, unsigned long long key, int* out) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) out[j] = (int)qfx::feistel_perm((uint32_t)j, (uint32_t)n, key);
}
extern "C" int qfx_debug_feistel(int n, unsigned long long key, int* out_dev, cudaStream_t stream) {
    QFX_ARG_CHECK(n >= 1 && out_dev, QFX_ERR_BAD_ARG, "bad arguments");
    feistel_kernel<<<(n + 255) / 256, 256, 0, stream>>>(n, key, out_dev);
    return qfx::check_launch("feistel_kernel");
}
