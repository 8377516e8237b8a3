// tree.cu -- RAHT tree construction.
// Replaces python/RAHT_param.py:6-83 (RAHT_param) and :191-279 (RAHT_param_reorder_fast): instead
// of up to 3*depth rounds of gather / xor / mask / boolean-compaction with a host sync each, one
// XOR+CLZ pass gives level[] (App. D), from which List/Flags/weights (any level, on demand),
// order_RAGFT (one stable 22-way partition), w, and the shared-memory transform plan are derived.
#include "plan.cuh"
#include <math.h>
#include <new>

namespace {

// ---- level[] ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) levels_kernel(const int64_t *__restrict__ MC, int64_t n, uint8_t *__restrict__ level,
                                                     int32_t *status)
{
    bool bad = false;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (i == 0) { level[0] = 255; continue; }
        int64_t a = MC[i - 1], b = MC[i];
        bad |= !(a < b);
        uint64_t x = (uint64_t)a ^ (uint64_t)b;
        level[i] = x ? (uint8_t)(63 - __clzll((long long)x)) : (uint8_t)0;
    }
    if (__any_sync(~0u, bad) && lane_id() == 0 && status) atomicOr(status, RAHT_ST_NOT_SORTED_UNIQUE);
}

// ---- one level of List / Flags / weights (RAHT_param.py:226-248) -----------------------------------
struct LevelPred {
    const uint8_t *lev;
    int l;
    __device__ bool operator()(int64_t i) const { return i == 0 || lev[i] >= l; }
};
struct LevelEmit {
    int64_t *list;
    __device__ void operator()(unsigned long long pos, int64_t i) const { list[pos] = i; }
};
__global__ void __launch_bounds__(256) level_fill_kernel(const uint8_t *__restrict__ lev, int64_t n, int l,
                                                         const int64_t *__restrict__ list, const int64_t *count,
                                                         uint8_t *__restrict__ flags, int64_t *__restrict__ weights)
{
    const int64_t L = *count;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < L; k += (int64_t)gridDim.x * blockDim.x) {
        int64_t cur = list[k];
        int64_t nxt = (k + 1 < L) ? list[k + 1] : n;                    // :228-230
        if (weights) weights[k] = nxt - cur;
        if (flags) flags[k] = (k + 1 < L) && (lev[nxt] == l);           // :238-243
    }
}

__global__ void max_level_init(int32_t *m) { *m = -1; }
__global__ void __launch_bounds__(256) max_level_kernel(const uint8_t *__restrict__ lev, int64_t n, int32_t *out)
{
    int m = -1;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x + 1; i < n; i += (int64_t)gridDim.x * blockDim.x)
        m = max(m, (int)lev[i]);
    for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(~0u, m, o));
    if (lane_id() == 0 && m >= 0) atomicMax(out, m);
}

// ---- order_RAGFT: stable partition into 22 buckets (RAHT_param.py:219-274) -------------------------
constexpr int OR_BUCKETS = 22;       // bucket 0 = root, 1 + (20 - level/3) otherwise (coarse -> fine)
constexpr int OR_THREADS = 256;
constexpr int OR_IPT = 8;
constexpr int OR_TILE = OR_THREADS * OR_IPT;
constexpr int OR_WARPS = OR_THREADS / 32;

__device__ __forceinline__ int or_bucket(int64_t i, uint8_t lev) { return i == 0 ? 0 : 1 + (20 - min((int)lev, 62) / 3); }

__global__ void __launch_bounds__(OR_THREADS) order_count_kernel(const uint8_t *__restrict__ lev, int64_t n, int64_t nblocks,
                                                                unsigned long long *counts /* [bucket][block] */)
{
    __shared__ int hist[OR_BUCKETS];
    if (threadIdx.x < OR_BUCKETS) hist[threadIdx.x] = 0;
    __syncthreads();
    int64_t base = (int64_t)blockIdx.x * OR_TILE;
    for (int k = 0; k < OR_IPT; ++k) {
        int64_t i = base + k * OR_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&hist[or_bucket(i, lev[i])], 1);
    }
    __syncthreads();
    if (threadIdx.x < OR_BUCKETS) counts[(size_t)threadIdx.x * nblocks + blockIdx.x] = (unsigned long long)hist[threadIdx.x];
}

__global__ void __launch_bounds__(OR_THREADS) order_scatter_kernel(const uint8_t *__restrict__ lev, int64_t n, int64_t nblocks,
                                                                  const unsigned long long *__restrict__ offsets,
                                                                  int64_t *__restrict__ order, int32_t *__restrict__ position)
{
    __shared__ int whist[OR_WARPS][OR_BUCKETS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < OR_WARPS * OR_BUCKETS; i += OR_THREADS) (&whist[0][0])[i] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * OR_TILE;
    int bucket[OR_IPT], rank[OR_IPT];
#pragma unroll
    for (int j = 0; j < OR_IPT; ++j) {
        int64_t i = base + warp * (32 * OR_IPT) + j * 32 + lane;     // warp-striped: (warp, j, lane) == memory order
        int b = (i < n) ? or_bucket(i, lev[i]) : OR_BUCKETS;          // OR_BUCKETS = "none"
        unsigned peers = __match_any_sync(~0u, b);
        int leader = __ffs(peers) - 1;
        int before = __popc(peers & lanemask_lt());
        int old = 0;
        if (lane == leader && b < OR_BUCKETS) {
            old = whist[warp][b];
            whist[warp][b] = old + __popc(peers);
        }
        old = __shfl_sync(~0u, old, leader);
        bucket[j] = b;
        rank[j] = old + before;
        __syncwarp();
    }
    __syncthreads();
    if (tid < OR_BUCKETS) {
        int run = 0;
        for (int w = 0; w < OR_WARPS; ++w) {
            int c = whist[w][tid];
            whist[w][tid] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < OR_IPT; ++j) {
        int64_t i = base + warp * (32 * OR_IPT) + j * 32 + lane;
        if (i < n) {
            int b = bucket[j];
            unsigned long long dst = offsets[(size_t)b * nblocks + blockIdx.x] + (unsigned)(whist[warp][b] + rank[j]);
            if (order) order[dst] = i;
            if (position) position[i] = (int32_t)dst;
        }
    }
}

// ---- transform plan: one stage ------------------------------------------------------------------
constexpr int PB_THREADS = 256;

__global__ void __launch_bounds__(PB_THREADS)
plan_stage_kernel(const uint32_t *__restrict__ rows, const uint8_t *__restrict__ lev, int64_t n_entries, int64_t n_total,
                  int R, uint8_t *__restrict__ to_next, uint32_t *__restrict__ sched, double2 *__restrict__ ab,
                  TileHdr *__restrict__ hdr, double *__restrict__ w)
{
    extern __shared__ __align__(16) unsigned char pb_smem[];
    uint32_t *s_orig = (uint32_t *)pb_smem;          // R + 1
    uint16_t *s_s = (uint16_t *)(s_orig + R + 1);    // R   local index of the left row, 0xffff = not in tile
    uint16_t *s_t = s_s + R;                         // R   local index of the right boundary
    uint8_t *s_lev = (uint8_t *)(s_t + R);           // R + 1
    __shared__ int hist[PLAN_MAX_ROUNDS], offs[PLAN_MAX_ROUNDS], cur[PLAN_MAX_ROUNDS];
    __shared__ int s_nbf, s_nr;

    const int tid = threadIdx.x;
    const int64_t tile = blockIdx.x;
    const int64_t base = tile * R;
    const int n_rows = (int)min((int64_t)R, n_entries - base);
    const bool last = (tile == (int64_t)gridDim.x - 1);
    const bool multi = gridDim.x > 1;

    if (tid < PLAN_MAX_ROUNDS) { hist[tid] = 0; cur[tid] = 0; }
    for (int j = tid; j < n_rows; j += PB_THREADS) {
        s_lev[j] = lev[base + j];
        s_orig[j] = rows ? rows[base + j] : (uint32_t)(base + j);
    }
    if (tid == 0 && last) { s_lev[n_rows] = 255; s_orig[n_rows] = (uint32_t)n_total; }   // virtual end of the cloud
    __syncthreads();

    const int lim = last ? n_rows + 1 : n_rows;
    for (int j = tid; j < n_rows; j += PB_THREADS) {
        const int l = s_lev[j];
        if (l == 255) {                                   // the root: never a right sibling
            s_s[j] = 0xffff;
            to_next[base + j] = multi ? 1 : 0;
            if (!multi && w) w[s_orig[j]] = (double)n_total;
            continue;
        }
        int s = -1, t = -1;
        for (int i = j - 1; i >= 0; --i)
            if (s_lev[i] > l) { s = i; break; }
        if (s >= 0)
            for (int i = j + 1; i < lim; ++i)
                if (s_lev[i] > l) { t = i; break; }
        const bool in_tile = (s >= 0) && (t >= 0);
        s_s[j] = in_tile ? (uint16_t)s : (uint16_t)0xffff;
        s_t[j] = (uint16_t)(t >= 0 ? t : 0);
        to_next[base + j] = in_tile ? 0 : 1;
        if (in_tile) atomicAdd(&hist[min(l, PLAN_MAX_ROUNDS - 1)], 1);
    }
    __syncthreads();
    if (tid < 32) {                                       // scan of the 64-bin level histogram by one warp
        int h0 = hist[tid], h1 = hist[tid + 32];
        int x0 = h0, x1 = h1;
        for (int o = 1; o < 32; o <<= 1) {
            int y0 = __shfl_up_sync(~0u, x0, o), y1 = __shfl_up_sync(~0u, x1, o);
            if (tid >= o) { x0 += y0; x1 += y1; }
        }
        int tot0 = __shfl_sync(~0u, x0, 31);
        offs[tid] = x0 - h0;
        offs[tid + 32] = tot0 + x1 - h1;
        unsigned m0 = __ballot_sync(~0u, h0 > 0), m1 = __ballot_sync(~0u, h1 > 0);
        TileHdr *H = hdr + tile;
        if (h0 > 0) H->round_end[__popc(m0 & lanemask_lt())] = (uint16_t)x0;
        if (h1 > 0) H->round_end[__popc(m0) + __popc(m1 & lanemask_lt())] = (uint16_t)(tot0 + x1);
        if (tid == 31) { s_nbf = tot0 + x1; s_nr = __popc(m0) + __popc(m1); }
    }
    __syncthreads();
    for (int j = tid; j < n_rows; j += PB_THREADS) {
        const int s = s_s[j];
        if (s == 0xffff) continue;
        const int l = s_lev[j], t = s_t[j];
        const int p = offs[l] + atomicAdd(&cur[l], 1);
        const double w0 = (double)(s_orig[j] - s_orig[s]);
        const double w1 = (double)(s_orig[t] - s_orig[j]);
        const double den = __dadd_rn(w0, w1);                           // RAHT.py:317
        sched[base + p] = (uint32_t)j | ((uint32_t)s << 10) | ((uint32_t)l << 20);
        ab[base + p] = make_double2(__dsqrt_rn(__ddiv_rn(w0, den)), __dsqrt_rn(__ddiv_rn(w1, den)));   // :321-322
        if (w) w[s_orig[j]] = den;                                      // :325-328 (final value of w[i1])
    }
    if (tid == 0) {
        TileHdr *H = hdr + tile;
        H->n_rows = (uint16_t)n_rows;
        H->n_bf = (uint16_t)s_nbf;
        H->n_rounds = (uint16_t)s_nr;
        H->pad = 0;
    }
}

struct CarryPred {
    const uint8_t *flag;
    __device__ bool operator()(int64_t i) const { return flag[i] != 0; }
};
struct CarryEmit {
    const uint32_t *rows;
    const uint8_t *lev;
    uint32_t *rows_out;
    uint8_t *lev_out;
    __device__ void operator()(unsigned long long pos, int64_t i) const
    {
        rows_out[pos] = rows ? rows[i] : (uint32_t)i;
        lev_out[pos] = lev[i];
    }
};

size_t plan_smem_bytes(int R) { return (size_t)(R + 1) * 4 + (size_t)R * 4 + (size_t)(R + 1) + 16; }

}  // namespace

RAHT_API int raht_tree_levels(const int64_t *MC, int64_t n, uint8_t *level, int32_t *status, void *stream)
{
    RAHT_REQUIRE(n >= 0 && (n == 0 || (MC && level)), "raht_tree_levels: bad arguments");
    if (n == 0) return RAHT_OK;
    (raht_count_launches(1), levels_kernel)<<<grid_for(n, 256, 8), 256, 0, (cudaStream_t)stream>>>(MC, n, level, status);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API size_t raht_tree_level_workspace_bytes(int64_t n) { return sel_workspace_bytes(n < 0 ? 0 : n) + 256; }

RAHT_API int raht_tree_level(const uint8_t *level, int64_t n, int l, int64_t *list, uint8_t *flags, int64_t *weights,
                             int64_t *count, void *workspace, size_t workspace_bytes, void *stream)
{
    RAHT_REQUIRE(n >= 1 && level && list && count && workspace && l >= 0 && l < 64, "raht_tree_level: bad arguments");
    if (workspace_bytes < raht_tree_level_workspace_bytes(n)) {
        raht_set_error("raht_tree_level: workspace too small");
        return RAHT_E_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    LevelPred pred{level, l};
    LevelEmit emit{list};
    device_select(n, pred, emit, count, workspace, st);
    (raht_count_launches(1), level_fill_kernel)<<<grid_for(n, 256, 8), 256, 0, st>>>(level, n, l, list, count, flags, weights);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_tree_max_level(const uint8_t *level, int64_t n, int32_t *max_level, void *stream)
{
    RAHT_REQUIRE(n >= 0 && max_level && (n == 0 || level), "raht_tree_max_level: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    (raht_count_launches(1), max_level_init)<<<1, 1, 0, st>>>(max_level);
    if (n > 1) (raht_count_launches(1), max_level_kernel)<<<grid_for(n, 256, 8), 256, 0, st>>>(level, n, max_level);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API size_t raht_order_ragft_workspace_bytes(int64_t n)
{
    if (n < 0) n = 0;
    int64_t nb = (n + OR_TILE - 1) / OR_TILE;
    return align_up((size_t)(nb * OR_BUCKETS + 1) * 8, 256) + 256;
}

RAHT_API int raht_order_ragft(const uint8_t *level, int64_t n, int64_t *order, int32_t *position, void *workspace,
                              size_t workspace_bytes, void *stream)
{
    RAHT_REQUIRE(n >= 0 && (n == 0 || (level && workspace)), "raht_order_ragft: bad arguments");
    if (n == 0) return RAHT_OK;
    if (workspace_bytes < raht_order_ragft_workspace_bytes(n)) {
        raht_set_error("raht_order_ragft: workspace too small");
        return RAHT_E_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    int64_t nb = (n + OR_TILE - 1) / OR_TILE;
    unsigned long long *counts = (unsigned long long *)workspace;
    (raht_count_launches(1), order_count_kernel)<<<(unsigned)nb, OR_THREADS, 0, st>>>(level, n, nb, counts);
    (raht_count_launches(1), scan_counts_kernel)<<<1, 1024, 0, st>>>(counts, nb * OR_BUCKETS, nullptr);
    (raht_count_launches(1), order_scatter_kernel)<<<(unsigned)nb, OR_THREADS, 0, st>>>(level, n, nb, counts, order, position);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

// ---- plan create / destroy -------------------------------------------------------------------------
RAHT_API void raht_plan_destroy(raht_plan *plan)
{
    if (!plan) return;
    for (int k = 0; k < plan->n_stages; ++k) {
        PlanStage &S = plan->st[k];
        cudaFree(S.rows);
        cudaFree(S.to_next);
        cudaFree(S.sched);
        cudaFree(S.ab);
        cudaFree(S.hdr);
    }
    delete plan;
}

#define PLAN_CUDA(call)                                                                            \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess) {                                                                  \
            raht_set_error("%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));    \
            cudaFree(lev_cur_owned);                                                               \
            cudaFree(sel_ws);                                                                      \
            cudaFree(d_count);                                                                     \
            raht_plan_destroy(P);                                                                  \
            return RAHT_E_CUDA;                                                                    \
        }                                                                                          \
    } while (0)

RAHT_API int raht_plan_create(const uint8_t *level, int64_t n, int tile_rows, double *w, raht_plan **plan, void *stream)
{
    RAHT_REQUIRE(plan, "raht_plan_create: plan is null");
    *plan = nullptr;
    RAHT_REQUIRE(n >= 0 && n < ((int64_t)1 << 32) - 1, "raht_plan_create: n out of range");
    RAHT_REQUIRE(tile_rows >= 128 && tile_rows <= 1024 && (tile_rows % 2) == 0,
                 "raht_plan_create: tile_rows must be even and in [128, 1024]");
    RAHT_REQUIRE(n == 0 || level, "raht_plan_create: level is null");
    cudaStream_t st = (cudaStream_t)stream;
    raht_plan *P = new (std::nothrow) raht_plan();
    RAHT_REQUIRE(P, "raht_plan_create: out of host memory");
    P->n = n;
    P->tile_rows = tile_rows;
    P->n_stages = 0;
    P->bytes = 0;
    const int R = tile_rows;
    uint8_t *lev_cur_owned = nullptr;     // level array of the current stage when it is ours
    void *sel_ws = nullptr;
    int64_t *d_count = nullptr;
    if (n == 0) { *plan = P; return RAHT_OK; }

    PLAN_CUDA(cudaMalloc(&sel_ws, sel_workspace_bytes(n)));
    PLAN_CUDA(cudaMalloc(&d_count, 256));
    if (plan_smem_bytes(R) > 48 * 1024)
        PLAN_CUDA(cudaFuncSetAttribute(plan_stage_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)plan_smem_bytes(R)));

    const uint8_t *lev_cur = level;
    uint32_t *rows_cur = nullptr;
    int64_t n_cur = n;
    for (int k = 0;; ++k) {
        if (k >= PLAN_MAX_STAGES) {
            raht_set_error("raht_plan_create: too many stages");
            cudaFree(lev_cur_owned); cudaFree(sel_ws); cudaFree(d_count);
            raht_plan_destroy(P);
            return RAHT_E_INVALID;
        }
        PlanStage &S = P->st[k];
        S = PlanStage();
        P->n_stages = k + 1;
        S.n_entries = n_cur;
        S.n_tiles = (n_cur + R - 1) / R;
        S.rows = rows_cur;                                    // ownership moves into the plan
        rows_cur = nullptr;
        size_t slots = (size_t)S.n_tiles * R;
        PLAN_CUDA(cudaMalloc(&S.to_next, (size_t)n_cur));
        PLAN_CUDA(cudaMalloc(&S.sched, slots * sizeof(uint32_t)));
        PLAN_CUDA(cudaMalloc(&S.ab, slots * sizeof(double2)));
        PLAN_CUDA(cudaMalloc(&S.hdr, (size_t)S.n_tiles * sizeof(TileHdr)));
        P->bytes += (size_t)n_cur + slots * 20 + (size_t)S.n_tiles * sizeof(TileHdr) + (S.rows ? (size_t)n_cur * 4 : 0);
        (raht_count_launches(1), plan_stage_kernel)<<<(unsigned)S.n_tiles, PB_THREADS, plan_smem_bytes(R), st>>>(
            S.rows, lev_cur, n_cur, n, R, S.to_next, S.sched, S.ab, S.hdr, w);
        PLAN_CUDA(cudaGetLastError());
        if (S.n_tiles == 1) break;

        // compact the carried entries into the next stage (order preserved)
        size_t cap = (size_t)S.n_tiles * 128;                 // <= 63 prefix + 63 suffix records + root per tile
        if (cap > (size_t)n_cur) cap = (size_t)n_cur;
        uint32_t *rows_next = nullptr;
        uint8_t *lev_next = nullptr;
        PLAN_CUDA(cudaMalloc(&rows_next, cap * sizeof(uint32_t)));
        rows_cur = rows_next;
        cudaError_t e2 = cudaMalloc(&lev_next, cap);
        if (e2 != cudaSuccess) { cudaFree(rows_next); rows_cur = nullptr; PLAN_CUDA(e2); }
        CarryPred pred{S.to_next};
        CarryEmit emit{S.rows, lev_cur, rows_next, lev_next};
        device_select(n_cur, pred, emit, d_count, sel_ws, st);
        int64_t h_count = 0;
        cudaError_t e3 = cudaMemcpyAsync(&h_count, d_count, 8, cudaMemcpyDeviceToHost, st);
        if (e3 == cudaSuccess) e3 = cudaStreamSynchronize(st);
        if (e3 != cudaSuccess) { cudaFree(rows_next); cudaFree(lev_next); rows_cur = nullptr; PLAN_CUDA(e3); }
        cudaFree(lev_cur_owned);
        lev_cur_owned = lev_next;
        lev_cur = lev_next;
        if (h_count <= 0 || h_count >= n_cur) {
            raht_set_error("raht_plan_create: stage %d did not shrink (%lld -> %lld); is level[] from raht_tree_levels?", k,
                           (long long)n_cur, (long long)h_count);
            cudaFree(rows_next); cudaFree(lev_cur_owned); cudaFree(sel_ws); cudaFree(d_count);
            raht_plan_destroy(P);
            return RAHT_E_INPUT;
        }
        n_cur = h_count;
    }
    cudaError_t e4 = cudaStreamSynchronize(st);
    cudaFree(lev_cur_owned);
    lev_cur_owned = nullptr;
    cudaFree(sel_ws);
    sel_ws = nullptr;
    cudaFree(d_count);
    d_count = nullptr;
    PLAN_CUDA(e4);
    *plan = P;
    return RAHT_OK;
}

RAHT_API int raht_plan_info(const raht_plan *plan, int64_t *info_host, int cap)
{
    RAHT_REQUIRE(plan && info_host && cap >= 4, "raht_plan_info: bad arguments");
    info_host[0] = plan->n;
    info_host[1] = plan->tile_rows;
    info_host[2] = plan->n_stages;
    info_host[3] = (int64_t)plan->bytes;
    for (int k = 0; k < plan->n_stages && 4 + k < cap; ++k) info_host[4 + k] = plan->st[k].n_entries;
    return RAHT_OK;
}
