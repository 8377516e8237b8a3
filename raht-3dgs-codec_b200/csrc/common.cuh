// common.cuh -- shared helpers for libraht_b200 (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <stdarg.h>
#include "../../include/raht_b200.h"

#define RAHT_API extern "C" __attribute__((visibility("default")))

void raht_set_error(const char *fmt, ...);
void raht_count_launches(int n);   // bookkeeping for raht_launch_count() (bench.py reports it)
cudaStream_t raht_side_fork(cudaStream_t st);   // api.cu: the device's side stream, waiting for what `st` holds so far (nullptr: unavailable)
int raht_side_join(cudaStream_t st);             // `st` waits for what the side stream holds so far
void raht_merge_share_gpu();                     // merge.cu: the next raht_merge_segments of this thread runs as short-lived CTAs
void raht_rlgr_decode_share_gpu();               // rlgr.cu: the next raht_rlgr_decode of this thread leaves room for a side-stream leg

#define RAHT_REQUIRE(cond, ...)                                                                    \
    do {                                                                                           \
        if (!(cond)) {                                                                             \
            raht_set_error(__VA_ARGS__);                                                           \
            return RAHT_E_INVALID;                                                                 \
        }                                                                                          \
    } while (0)

#define RAHT_CUDA(call)                                                                            \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess) {                                                                  \
            raht_set_error("%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));    \
            return RAHT_E_CUDA;                                                                    \
        }                                                                                          \
    } while (0)

#define RAHT_TRY(expr)                                                                             \
    do {                                                                                           \
        int rc_ = (expr);                                                                          \
        if (rc_ != RAHT_OK) return rc_;                                                            \
    } while (0)

#define RAHT_LAUNCH_CHECK()                                                                        \
    do {                                                                                           \
        cudaError_t e__ = cudaGetLastError();                                                      \
        if (e__ != cudaSuccess) {                                                                  \
            raht_set_error("%s:%d launch: %s", __FILE__, __LINE__, cudaGetErrorString(e__));       \
            return RAHT_E_CUDA;                                                                    \
        }                                                                                          \
    } while (0)

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// bump allocator over the caller's workspace
struct Arena {
    char *base;
    size_t off, cap;
    bool ok;
    Arena(void *p, size_t bytes) : base((char *)p), off(0), cap(bytes), ok(true) {}
    template <typename T> T *take(size_t count)
    {
        off = align_up(off, 256);
        T *r = (T *)(base + off);
        off += count * sizeof(T);
        if (off > cap || (base == nullptr && count)) ok = false;
        return r;
    }
};

static inline int sm_count_cached()
{
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (sms <= 0) sms = 148;
    }
    return sms;
}

// grid sized as a multiple of the SM count for grid-stride kernels
static inline int grid_for(int64_t work_items, int per_block, int ctas_per_sm)
{
    int64_t need = (work_items + per_block - 1) / per_block;
    int64_t cap = (int64_t)sm_count_cached() * ctas_per_sm;
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
}

// ---- programmatic dependent launch (PDL) ---------------------------------------------------------------------------------
// A kernel launched with pdl_launch may start while the kernel before it in the stream is still running (once every CTA of
// that kernel has executed pdl_trigger or exited): it runs its prologue -- loads of data that the predecessor does not
// write -- and must execute pdl_wait before it touches anything the predecessor produces (the wait returns when the
// predecessor grid has completed and its writes are visible).  Both instructions are no-ops in a plain launch.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename... KArgs, typename... Args>
static inline cudaError_t pdl_launch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// Launch-gap hiding for chains of short dependent kernels: a kernel whose first statement is pdl_enter() (trigger, then
// wait) and that is launched with RAHT_PDL_LAUNCH becomes resident while the kernel before it in the stream drains and
// starts the moment that one completes.  Semantics are those of a plain launch (nothing runs ahead of the wait).  Only
// behind another KERNEL of the same stream: after a memset, a copy or an event the site uses a plain launch.
__device__ __forceinline__ void pdl_enter() { pdl_trigger(); pdl_wait(); }
// RAHT_PDL: bit mask of the launch groups that use it (A/B): 1 transform stages, 2 front leg (keys, sort, voxel boundaries),
// 4 tree build, 8 RLGR encoder, 16 quantisation.
enum { PDL_TRANSFORM = 1, PDL_FRONT = 2, PDL_TREE = 4, PDL_RLGR = 8, PDL_QUANT = 16, PDL_DEFAULT = 1 | 2 | 16 };
static inline bool raht_use_pdl(int group)
{
    static const int v = getenv("RAHT_PDL") ? atoi(getenv("RAHT_PDL")) : PDL_DEFAULT;
    return (v & group) != 0;
}
#define RAHT_PDL_LAUNCH(group, kern, grid, block, smem, st, ...)                                   \
    do {                                                                                           \
        raht_count_launches(1);                                                                    \
        if (raht_use_pdl(group)) RAHT_CUDA(pdl_launch(kern, dim3(grid), dim3(block), (size_t)(smem), st, __VA_ARGS__)); \
        else kern<<<grid, block, smem, st>>>(__VA_ARGS__);                                         \
    } while (0)

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ unsigned lanemask_lt()
{
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// ---- device-wide index compaction (three launches, no inter-CTA spinning) -----------------------
// pass 1: per-block counts; pass 2: one-block exclusive scan of the counts; pass 3: scatter.
constexpr int SEL_THREADS = 256;
constexpr int SEL_ITEMS = 8;
constexpr int SEL_TILE = SEL_THREADS * SEL_ITEMS;

static inline int64_t sel_num_blocks(int64_t n) { return (n + SEL_TILE - 1) / SEL_TILE; }
static inline size_t sel_workspace_bytes(int64_t n) { return align_up((sel_num_blocks(n) + 1) * 8, 256) + 256; }

template <typename Pred>
__global__ void __launch_bounds__(SEL_THREADS) sel_count_kernel(int64_t n, Pred pred, unsigned long long *block_counts)
{
    pdl_enter();
    int64_t base = (int64_t)blockIdx.x * SEL_TILE;
    int cnt = 0;
#pragma unroll
    for (int k = 0; k < SEL_ITEMS; ++k) {
        int64_t i = base + k * SEL_THREADS + threadIdx.x;
        if (i < n && pred(i)) ++cnt;
    }
    __shared__ int wsum[SEL_THREADS / 32];
    for (int o = 16; o; o >>= 1) cnt += __shfl_xor_sync(~0u, cnt, o);
    if (lane_id() == 0) wsum[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < SEL_THREADS / 32; ++w) t += wsum[w];
        block_counts[blockIdx.x] = (unsigned long long)t;
    }
}

// exclusive scan of `m` counts in place by ONE block; total -> *total_out (and counts[m]).
// A thread owns SCAN_IPT consecutive counts per sweep (8192 per sweep of the block): the 22-bucket x 489-block table of
// order_RAGFT at 1 M rows is two sweeps (one count per thread and sweep made it eleven, four barriers each: 14 us).
constexpr int SCAN_IPT = 8;
static __global__ void __launch_bounds__(1024) scan_counts_kernel(unsigned long long *counts, int64_t m, int64_t *total_out)
{
    __shared__ unsigned long long wtot[32];
    __shared__ unsigned long long carry;
    pdl_enter();
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int64_t base = 0; base < m; base += 1024 * SCAN_IPT) {
        const int64_t i0 = base + (int64_t)threadIdx.x * SCAN_IPT;
        unsigned long long v[SCAN_IPT], x = 0;
#pragma unroll
        for (int t = 0; t < SCAN_IPT; ++t) {
            v[t] = (i0 + t < m) ? counts[i0 + t] : 0ull;
            x += v[t];
        }
        const unsigned long long mine = x;
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long y = __shfl_up_sync(~0u, x, o);
            if ((int)lane_id() >= o) x += y;
        }
        if (lane_id() == 31) wtot[threadIdx.x >> 5] = x;
        __syncthreads();
        if (threadIdx.x < 32) {
            unsigned long long t = wtot[threadIdx.x];
            unsigned long long s = t;
            for (int o = 1; o < 32; o <<= 1) {
                unsigned long long y = __shfl_up_sync(~0u, s, o);
                if ((int)lane_id() >= o) s += y;
            }
            wtot[threadIdx.x] = s - t;   // exclusive warp offsets
        }
        __syncthreads();
        unsigned long long excl = carry + wtot[threadIdx.x >> 5] + (x - mine);
#pragma unroll
        for (int t = 0; t < SCAN_IPT; ++t) {
            if (i0 + t < m) counts[i0 + t] = excl;
            excl += v[t];
        }
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        counts[m] = carry;
        if (total_out) *total_out = (int64_t)carry;
    }
}

template <typename Pred, typename Emit>
__global__ void __launch_bounds__(SEL_THREADS) sel_scatter_kernel(int64_t n, Pred pred, Emit emit,
                                                                 const unsigned long long *block_offsets)
{
    pdl_enter();
    int64_t base = (int64_t)blockIdx.x * SEL_TILE;
    __shared__ int wsum[SEL_THREADS / 32];
    __shared__ int running;
    if (threadIdx.x == 0) running = 0;
    __syncthreads();
    unsigned long long boff = block_offsets[blockIdx.x];
    // item order inside the tile: k-major, then thread => ascending i
    for (int k = 0; k < SEL_ITEMS; ++k) {
        int64_t i = base + k * SEL_THREADS + threadIdx.x;
        bool p = (i < n) && pred(i);
        unsigned b = __ballot_sync(~0u, p);
        int wcount = __popc(b);
        if (lane_id() == 0) wsum[threadIdx.x >> 5] = wcount;
        __syncthreads();
        int woff = 0, tot = 0;
        for (int w = 0; w < SEL_THREADS / 32; ++w) {
            int c = wsum[w];
            if (w < (int)(threadIdx.x >> 5)) woff += c;
            tot += c;
        }
        int run = running;
        if (p) emit(boff + (unsigned long long)(run + woff + __popc(b & lanemask_lt())), i);
        __syncthreads();
        if (threadIdx.x == 0) running = run + tot;
        __syncthreads();
    }
}

// Runs the three launches.  ws must hold sel_workspace_bytes(n).  count_out: device int64 (may be null).
template <typename Pred, typename Emit>
static inline int device_select(int64_t n, Pred pred, Emit emit, int64_t *count_out, void *ws, cudaStream_t st)
{
    int64_t nb = sel_num_blocks(n);
    raht_count_launches(n ? 2 : 0);
    unsigned long long *counts = (unsigned long long *)ws;
    if (n == 0) {
        (raht_count_launches(1), scan_counts_kernel)<<<1, 1024, 0, st>>>(counts, 0, count_out);
        return 0;
    }
    RAHT_PDL_LAUNCH(PDL_FRONT, (sel_count_kernel<Pred>), (unsigned)nb, SEL_THREADS, 0, st, n, pred, counts);
    raht_count_launches(-1);
    RAHT_PDL_LAUNCH(PDL_FRONT, scan_counts_kernel, 1, 1024, 0, st, counts, nb, count_out);
    RAHT_PDL_LAUNCH(PDL_FRONT, (sel_scatter_kernel<Pred, Emit>), (unsigned)nb, SEL_THREADS, 0, st, n, pred, emit, (const unsigned long long *)counts);
    raht_count_launches(-1);                                 // (count and scatter were counted above)
    return 0;
}

// ---- mbarrier + 1-D bulk async copy (TMA, SASS UBLKCP) helpers ----------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// global -> shared bulk copy, completion counted on `bar`.  16-byte aligned, bytes % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gmem_src, unsigned bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// shared -> global bulk copy (bulk async-group completion)
__device__ __forceinline__ void bulk_s2g(void *gmem_dst, const void *smem_src, unsigned bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
