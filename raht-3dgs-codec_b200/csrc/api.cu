// api.cu -- error reporting, version and device queries of libraht_b200.
#include "common.cuh"
#include <string.h>

static thread_local char g_err[512] = "";
static long long g_launches = 0;

void raht_count_launches(int n) { __atomic_fetch_add(&g_launches, (long long)n, __ATOMIC_RELAXED); }
RAHT_API long long raht_launch_count(void) { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

void raht_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

RAHT_API int raht_abi_version(void) { return RAHT_B200_ABI_VERSION; }
RAHT_API const char *raht_last_error(void) { return g_err; }

RAHT_API int raht_device_info(int *sm_count, int *cc)
{
    int dev = 0;
    RAHT_CUDA(cudaGetDevice(&dev));
    cudaDeviceProp p;
    RAHT_CUDA(cudaGetDeviceProperties(&p, dev));
    if (sm_count) *sm_count = p.multiProcessorCount;
    if (cc) *cc = p.major * 10 + p.minor;
    if (p.major != 10) {
        raht_set_error("device %s is sm_%d%d; libraht_b200 is built for sm_100a only", p.name, p.major, p.minor);
        return RAHT_E_NODEVICE;
    }
    return RAHT_OK;
}
