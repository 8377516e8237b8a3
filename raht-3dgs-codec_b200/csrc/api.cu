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

// ---- small results to the host without a copy engine ----------------------------------------------------------------
// A size that must reach the host (voxel count, status word, stream lengths) is a few bytes; as a cudaMemcpyAsync it
// queues behind whatever bulk device-to-host copy another stream has in flight on the same copy engine (measured: the
// 46 MB read-back of the previous scene's streams delayed every synchronisation point of the next encode, +0.65 ms per
// scene).  This kernel stores the bytes straight into pinned host memory over PCIe instead.
static __global__ void publish_kernel(const unsigned char *__restrict__ src, volatile unsigned char *dst, unsigned bytes)
{
    if (((((uintptr_t)src) | ((uintptr_t)dst) | bytes) & 3) == 0) {
        const unsigned *s4 = (const unsigned *)src;
        volatile unsigned *d4 = (volatile unsigned *)dst;
        for (unsigned i = threadIdx.x; i < bytes / 4; i += blockDim.x) d4[i] = s4[i];
    } else {
        for (unsigned i = threadIdx.x; i < bytes; i += blockDim.x) dst[i] = src[i];
    }
    __threadfence_system();
}

RAHT_API int raht_publish_host(const void *src, void *pinned_host, size_t bytes, void *stream)
{
    RAHT_REQUIRE(bytes <= (1u << 20), "raht_publish_host: meant for small results (<= 1 MiB)");
    if (bytes == 0) return RAHT_OK;
    RAHT_REQUIRE(src && pinned_host, "raht_publish_host: null pointer");
    void *mapped = nullptr;
    if (cudaHostGetDevicePointer(&mapped, pinned_host, 0) != cudaSuccess || !mapped) {
        (void)cudaGetLastError();                                    // not mapped memory: plain asynchronous copy
        RAHT_CUDA(cudaMemcpyAsync(pinned_host, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
        return RAHT_OK;
    }
    raht_count_launches(1);
    publish_kernel<<<1, 128, 0, (cudaStream_t)stream>>>((const unsigned char *)src, (volatile unsigned char *)mapped, (unsigned)bytes);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}
