// bx_api.cu — error text, version and device queries of the C ABI (include/burgers_b200.h)
#include "bx_common.cuh"
#include <cstring>

namespace bx {
static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace bx

extern "C" int bx_version(void) { return 100; }

extern "C" const char *bx_last_error(void) { return bx::g_err; }

extern "C" int bx_device_sm_count(void) {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return BX_ERR_CUDA;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return BX_ERR_CUDA;
    return sms;
}
