// api.cu -- error reporting, version and device queries of libraht_b200.
#include "common.cuh"
#include <string.h>
#include <stdlib.h>
#include <mutex>

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

// ---- per-device side stream -------------------------------------------------------------------------------------------
// Two legs of a call that touch disjoint data (the table passes of the sparse RLGR channels next to the second
// speculation sweep; the decoder's tree next to its entropy decode) run side by side: raht_side_fork makes the device's
// side stream wait for everything enqueued on `st` so far and returns it, raht_side_join makes `st` wait for everything
// enqueued on the side stream so far.  Events only (legal under stream capture); stream and events are created on first
// use and never destroyed.  Each function enqueues its record + wait pair under a lock, so host threads that share a
// device share the side stream safely (at worst they wait for each other's side work).
namespace {
struct SideStream {
    cudaStream_t s = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr;
    bool ready = false;
};
std::mutex g_side_mutex;
SideStream *side_locked()
{
    static SideStream tab[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    SideStream &S = tab[dev];
    if (!S.ready) {
        // highest priority: the side legs are short, latency-bound kernels; their CTAs should take the slots the bulk kernel
        // on the caller's stream frees instead of queueing behind its whole grid (RAHT_SIDE_PRIO=0: default priority)
        int lo = 0, hi = 0;
        static const int want_prio = getenv("RAHT_SIDE_PRIO") ? atoi(getenv("RAHT_SIDE_PRIO")) : 1;
        if (cudaDeviceGetStreamPriorityRange(&lo, &hi) != cudaSuccess) { lo = hi = 0; (void)cudaGetLastError(); }
        if (cudaStreamCreateWithPriority(&S.s, cudaStreamNonBlocking, want_prio ? hi : 0) != cudaSuccess) return nullptr;
        if (cudaEventCreateWithFlags(&S.fork, cudaEventDisableTiming) != cudaSuccess) return nullptr;
        if (cudaEventCreateWithFlags(&S.join, cudaEventDisableTiming) != cudaSuccess) return nullptr;
        S.ready = true;
    }
    return &S;
}
}  // namespace

cudaStream_t raht_side_fork(cudaStream_t st)
{
    std::lock_guard<std::mutex> lock(g_side_mutex);
    SideStream *S = side_locked();
    if (!S || cudaEventRecord(S->fork, st) != cudaSuccess || cudaStreamWaitEvent(S->s, S->fork, 0) != cudaSuccess) {
        (void)cudaGetLastError();
        return nullptr;
    }
    return S->s;
}

int raht_side_join(cudaStream_t st)
{
    std::lock_guard<std::mutex> lock(g_side_mutex);
    SideStream *S = side_locked();
    RAHT_REQUIRE(S != nullptr, "raht_side_join: no side stream on this device");
    RAHT_CUDA(cudaEventRecord(S->join, S->s));
    RAHT_CUDA(cudaStreamWaitEvent(st, S->join, 0));
    return RAHT_OK;
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
