// runtime.cu -- error string, device queries.  Part of libexplainn_b200.so.
#include "common.cuh"
#include <stdarg.h>
#include "../../include/explainn_b200.h"

static thread_local char g_err[512] = "";

void xb_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int xb_check_launch(const char *what)
{
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        xb_set_error("%s: %s", what, cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    return XB_OK;
}

int xb_num_sms()
{
    static thread_local int cached_dev = -1, cached = 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (dev != cached_dev) {
        cudaDeviceProp p;
        if (cudaGetDeviceProperties(&p, dev) != cudaSuccess) return 148;
        cached = p.multiProcessorCount;
        cached_dev = dev;
    }
    return cached;
}

extern "C" const char *explainn_last_error(void) { return g_err; }
extern "C" int explainn_abi_version(void) { return EXPLAINN_B200_ABI_VERSION; }
extern "C" int explainn_num_sms(void) { return xb_num_sms(); }
