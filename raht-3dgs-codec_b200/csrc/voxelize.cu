// voxelize.cu -- voxelisation front end.
// Replaces python/voxelize_pc.py:25-59 (get_morton_code), :62-172 (voxelize_pc_batched) and the
// Morton blocks of python/RAHT_param.py:21-36, :201-212.  All HBM-bound streaming kernels: one
// coalesced pass each, grid-stride with the grid a multiple of the SM count.
#include "common.cuh"
#include <math.h>

// ---- Morton helpers: digit_i = z_i | y_i<<1 | x_i<<2 at bits 3i..3i+2 (voxelize_pc.py:46-57) ----
__host__ __device__ __forceinline__ uint64_t spread3(uint64_t x)
{
    x &= 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}
__host__ __device__ __forceinline__ uint64_t compact3(uint64_t x)
{
    x &= 0x1249249249249249ull;
    x = (x ^ (x >> 2)) & 0x10c30c30c30c30c3ull;
    x = (x ^ (x >> 4)) & 0x100f00f00f00f00full;
    x = (x ^ (x >> 8)) & 0x1f0000ff0000ffull;
    x = (x ^ (x >> 16)) & 0x1f00000000ffffull;
    x = (x ^ (x >> 32)) & 0x1fffffull;
    return x;
}
__host__ __device__ __forceinline__ uint64_t morton3(uint64_t x, uint64_t y, uint64_t z, int J)
{
    uint64_t m = ((uint64_t)1 << J) - 1;
    return spread3(z & m) | (spread3(y & m) << 1) | (spread3(x & m) << 2);
}

// ---- bounds -----------------------------------------------------------------------------------
// (the branch is on the SIGN BIT, not on v >= 0: -0.0f must take the path of the negative values, where its pattern
// 0x80000000 orders correctly; as a signed int it would be INT_MIN and win every minimum / lose every maximum)
__device__ __forceinline__ void atomic_min_f32(float *addr, float v)
{
    if (__float_as_int(v) >= 0) atomicMin((int *)addr, __float_as_int(v));
    else atomicMax((unsigned *)addr, __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_f32(float *addr, float v)
{
    if (__float_as_int(v) >= 0) atomicMax((int *)addr, __float_as_int(v));
    else atomicMin((unsigned *)addr, __float_as_uint(v));
}

__global__ void bounds_init_kernel(float *b)
{
    pdl_enter();
    if (threadIdx.x < 3) b[threadIdx.x] = INFINITY;
    else if (threadIdx.x < 6) b[threadIdx.x] = -INFINITY;
}

__global__ void __launch_bounds__(256) bounds3_kernel(const float *__restrict__ V, int64_t n, int64_t ld, float *bounds6)
{
    pdl_enter();
    float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float *p = V + i * ld;
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            float v = p[d];
            mn[d] = fminf(mn[d], v);
            mx[d] = fmaxf(mx[d], v);
        }
    }
    __shared__ float s[6][8];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        for (int o = 16; o; o >>= 1) {
            mn[d] = fminf(mn[d], __shfl_xor_sync(~0u, mn[d], o));
            mx[d] = fmaxf(mx[d], __shfl_xor_sync(~0u, mx[d], o));
        }
        if (lane_id() == 0) { s[d][threadIdx.x >> 5] = mn[d]; s[3 + d][threadIdx.x >> 5] = mx[d]; }
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        float r = s[threadIdx.x][0];
        for (int w = 1; w < 8; ++w) r = threadIdx.x < 3 ? fminf(r, s[threadIdx.x][w]) : fmaxf(r, s[threadIdx.x][w]);
        if (threadIdx.x < 3) atomic_min_f32(bounds6 + threadIdx.x, r);
        else atomic_max_f32(bounds6 + threadIdx.x, r);
    }
}

RAHT_API int raht_bounds3_f32(const float *V, int64_t n, int64_t ld, float *bounds6, void *stream)
{
    RAHT_REQUIRE(bounds6 && (V || n == 0) && ld >= 3 && n >= 0, "raht_bounds3_f32: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    (raht_count_launches(1), bounds_init_kernel)<<<1, 32, 0, st>>>(bounds6);
    if (n) RAHT_PDL_LAUNCH(PDL_FRONT, bounds3_kernel, grid_for(n, 256, 8), 256, 0, st, V, n, ld, bounds6);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

struct F3 { float v[3]; };

__global__ void voxel_params_kernel(const float *bounds6, int J, int use_vmin, F3 vmin_given, int use_width,
                                    double width_given, float *params4, double *params_d2)
{
    pdl_enter();
    if (threadIdx.x != 0) return;
    float vmin[3];
    for (int d = 0; d < 3; ++d) vmin[d] = use_vmin ? vmin_given.v[d] : bounds6[d];   // voxelize_pc.py:87-90
    double width;
    if (use_width) width = width_given;
    else {
        // width = (V - vmin).max().item(): fp32 subtraction is monotone, so the maximum over rows is
        // reached at the per-column maximum (voxelize_pc.py:92-95).
        float m = __fsub_rn(bounds6[3], vmin[0]);
        m = fmaxf(m, __fsub_rn(bounds6[4], vmin[1]));
        m = fmaxf(m, __fsub_rn(bounds6[5], vmin[2]));
        width = (double)m;
    }
    double voxel_size = width / (double)((int64_t)1 << J);                            // :97
    params4[0] = vmin[0]; params4[1] = vmin[1]; params4[2] = vmin[2];
    params4[3] = (float)voxel_size;     // Python scalar promoted to the tensor dtype
    params_d2[0] = width; params_d2[1] = voxel_size;
}

RAHT_API int raht_voxel_params(const float *bounds6, int J, int use_given_vmin, const float *vmin_host,
                               int use_given_width, double width_host, float *params4, double *params_d2,
                               void *stream)
{
    RAHT_REQUIRE(params4 && params_d2 && J >= 1 && J <= 21, "raht_voxel_params: bad arguments (J in [1,21])");
    RAHT_REQUIRE((use_given_vmin && use_given_width) || bounds6, "raht_voxel_params: bounds6 required");
    RAHT_REQUIRE(!use_given_vmin || vmin_host, "raht_voxel_params: vmin_host required");
    F3 vm = {{0, 0, 0}};
    if (use_given_vmin) { vm.v[0] = vmin_host[0]; vm.v[1] = vmin_host[1]; vm.v[2] = vmin_host[2]; }
    RAHT_PDL_LAUNCH(PDL_FRONT, voxel_params_kernel, 1, 32, 0, (cudaStream_t)stream, bounds6, J, use_given_vmin, vm, use_given_width, width_host, params4, params_d2);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

// ---- Morton keys ------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) morton_f32_kernel(const float *__restrict__ V, int64_t n, int64_t ld,
                                                         const float *__restrict__ params4, int J,
                                                         uint64_t *__restrict__ keys, int64_t *__restrict__ vint)
{
    pdl_enter();
    const float vx = params4[0], vy = params4[1], vz = params4[2], vs = params4[3];
    const int64_t hi = ((int64_t)1 << J) - 1;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float *p = V + i * ld;
        float v0[3] = {__fsub_rn(p[0], vx), __fsub_rn(p[1], vy), __fsub_rn(p[2], vz)};   // voxelize_pc.py:90
        int64_t q[3];
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            int64_t t = (int64_t)floorf(__fdiv_rn(v0[d], vs));                            // :98
            q[d] = t < 0 ? 0 : (t > hi ? hi : t);
        }
        keys[i] = morton3((uint64_t)q[0], (uint64_t)q[1], (uint64_t)q[2], J);            // :100
        if (vint) { vint[3 * i] = q[0]; vint[3 * i + 1] = q[1]; vint[3 * i + 2] = q[2]; }
    }
}

RAHT_API int raht_morton_keys_f32(const float *V, int64_t n, int64_t ld, const float *params4, int J,
                                  uint64_t *keys, int64_t *vint, void *stream)
{
    RAHT_REQUIRE(n >= 0 && ld >= 3 && J >= 1 && J <= 21 && params4 && (n == 0 || (V && keys)),
                 "raht_morton_keys_f32: bad arguments");
    if (n == 0) return RAHT_OK;
    RAHT_PDL_LAUNCH(PDL_FRONT, morton_f32_kernel, grid_for(n, 256, 8), 256, 0, (cudaStream_t)stream, V, n, ld, params4, J, keys, vint);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

__global__ void __launch_bounds__(256) morton_i64_kernel(const int64_t *__restrict__ V, int64_t n, int J,
                                                         int64_t *__restrict__ keys)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        keys[i] = (int64_t)morton3((uint64_t)V[3 * i], (uint64_t)V[3 * i + 1], (uint64_t)V[3 * i + 2], J);
}

RAHT_API int raht_morton_keys_i64(const int64_t *V, int64_t n, int J, int64_t *keys, void *stream)
{
    RAHT_REQUIRE(n >= 0 && J >= 1 && J <= 21 && (n == 0 || (V && keys)), "raht_morton_keys_i64: bad arguments");
    if (n == 0) return RAHT_OK;
    (raht_count_launches(1), morton_i64_kernel)<<<grid_for(n, 256, 8), 256, 0, (cudaStream_t)stream>>>(V, n, J, keys);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

struct D3 { double v[3]; };

__global__ void __launch_bounds__(256) morton_f64_kernel(const double *__restrict__ V, int64_t n, D3 minV, double Q,
                                                         int depth, int64_t *__restrict__ keys, int32_t *status)
{
    const int64_t hi = ((int64_t)1 << depth) - 1;
    bool bad = false;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t c[3];
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            c[d] = (int64_t)floor(__ddiv_rn(__dsub_rn(V[3 * i + d], minV.v[d]), Q));      // RAHT_param.py:22, :205
            bad |= (c[d] < 0) | (c[d] > hi);                                              // :26-27
        }
        keys[i] = (int64_t)morton3((uint64_t)c[0], (uint64_t)c[1], (uint64_t)c[2], depth);
    }
    if (__any_sync(~0u, bad) && lane_id() == 0 && status) atomicOr(status, RAHT_ST_OUT_OF_BOUNDS);
}

RAHT_API int raht_morton_keys_f64(const double *V, int64_t n, const double *minV_host, double width, int depth,
                                  int64_t *keys, int32_t *status, void *stream)
{
    RAHT_REQUIRE(n >= 0 && depth >= 1 && depth <= 21 && minV_host && (n == 0 || (V && keys)),
                 "raht_morton_keys_f64: bad arguments (depth in [1,21])");
    if (n == 0) return RAHT_OK;
    D3 mv = {{minV_host[0], minV_host[1], minV_host[2]}};
    double Q = width / (double)((int64_t)1 << depth);                                     // :21, :204
    (raht_count_launches(1), morton_f64_kernel)<<<grid_for(n, 256, 8), 256, 0, (cudaStream_t)stream>>>(V, n, mv, Q, depth, keys, status);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

__global__ void __launch_bounds__(256) morton_decode_kernel(const uint64_t *__restrict__ keys, int64_t n,
                                                            float *__restrict__ coords)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        uint64_t k = keys[i];
        coords[3 * i] = (float)compact3(k >> 2);
        coords[3 * i + 1] = (float)compact3(k >> 1);
        coords[3 * i + 2] = (float)compact3(k);
    }
}

RAHT_API int raht_morton_decode_f32(const uint64_t *keys, int64_t n, int J, float *coords, void *stream)
{
    (void)J;
    RAHT_REQUIRE(n >= 0 && (n == 0 || (keys && coords)), "raht_morton_decode_f32: bad arguments");
    if (n == 0) return RAHT_OK;
    (raht_count_launches(1), morton_decode_kernel)<<<grid_for(n, 256, 8), 256, 0, (cudaStream_t)stream>>>(keys, n, coords);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

// ---- voxel boundaries (voxelize_pc.py:114-120) ---------------------------------------------------
struct BoundaryPred {
    const uint64_t *k;
    __device__ bool operator()(int64_t i) const { return i == 0 || k[i] != k[i - 1]; }
};
struct BoundaryEmit {
    const uint64_t *k;
    int64_t *vi;
    int32_t *off;
    uint64_t *uk;
    __device__ void operator()(unsigned long long pos, int64_t i) const
    {
        if (vi) vi[pos] = i;
        if (off) off[pos] = (int32_t)i;
        if (uk) uk[pos] = k[i];
    }
};
__global__ void segments_tail_kernel(const int64_t *count, int64_t *nvox, int32_t *off, int64_t n)
{
    pdl_enter();
    int64_t c = *count;
    if (nvox) *nvox = c;
    if (off) off[c] = (int32_t)n;
}

RAHT_API size_t raht_voxel_segments_workspace_bytes(int64_t n) { return sel_workspace_bytes(n) + 256; }

RAHT_API int raht_voxel_segments(const uint64_t *sorted_keys, int64_t n, int64_t *voxel_indices,
                                 int32_t *cluster_offsets, uint64_t *unique_keys, int64_t *nvox, void *workspace,
                                 size_t workspace_bytes, void *stream)
{
    RAHT_REQUIRE(n >= 0 && (n == 0 || sorted_keys) && workspace, "raht_voxel_segments: bad arguments");
    if (workspace_bytes < raht_voxel_segments_workspace_bytes(n)) {
        raht_set_error("raht_voxel_segments: workspace too small");
        return RAHT_E_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    int64_t *count = (int64_t *)workspace;
    void *sel_ws = (char *)workspace + 256;
    BoundaryPred pred{sorted_keys};
    BoundaryEmit emit{sorted_keys, voxel_indices, cluster_offsets, unique_keys};
    device_select(n, pred, emit, count, sel_ws, st);
    RAHT_PDL_LAUNCH(PDL_FRONT, segments_tail_kernel, 1, 1, 0, st, (const int64_t *)count, nvox, cluster_offsets, n);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

// ---- PCsorted / DeltaV (voxelize_pc.py:103-111) ------------------------------------------------
__global__ void __launch_bounds__(256) gather_rows_kernel(const float *__restrict__ PC, int64_t n, int64_t ld,
                                                          const uint32_t *__restrict__ idx,
                                                          const float *__restrict__ params4,
                                                          float *__restrict__ PCsorted, float *__restrict__ DeltaV,
                                                          int64_t ld_delta)
{
    const int64_t total = n * ld;
    const float vs = params4 ? params4[3] : 1.0f;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        int64_t i = e / ld;
        int c = (int)(e - i * ld);
        float v = PC[(int64_t)idx[i] * ld + c];
        if (PCsorted) PCsorted[e] = v;
        if (DeltaV && c < 3) {
            float v0 = __fsub_rn(v, params4[c]);
            float vox = __fmul_rn(vs, floorf(__fdiv_rn(v0, vs)));
            DeltaV[i * ld_delta + c] = __fsub_rn(v0, vox);
        }
    }
}

RAHT_API int raht_gather_rows_f32(const float *PC, int64_t n, int64_t ld, const uint32_t *idx, const float *params4,
                                  float *PCsorted, float *DeltaV, int64_t ld_delta, void *stream)
{
    RAHT_REQUIRE(n >= 0 && ld >= 1 && (n == 0 || (PC && idx)) && (!DeltaV || (params4 && ld >= 3 && ld_delta >= 3)),
                 "raht_gather_rows_f32: bad arguments");
    if (n == 0) return RAHT_OK;
    (raht_count_launches(1), gather_rows_kernel)<<<grid_for(n * ld, 256, 8), 256, 0, (cudaStream_t)stream>>>(PC, n, ld, idx, params4, PCsorted,
                                                                                  DeltaV, ld_delta);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

// ---- per-voxel attribute mean (voxelize_pc.py:122-153) -----------------------------------------
__global__ void __launch_bounds__(256) attr_mean_kernel(const float *__restrict__ PC, int64_t n, int64_t ld, int d,
                                                        const uint32_t *__restrict__ idx,
                                                        const int64_t *__restrict__ voxel_indices, int64_t nvox,
                                                        float *__restrict__ Cvox, int64_t ld_out, int col_out,
                                                        float *__restrict__ DeltaC, int64_t ld_delta)
{
    const int64_t total = nvox * d;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        int64_t v = e / d;
        int c = (int)(e - v * d);
        int64_t s = voxel_indices[v], t = (v + 1 < nvox) ? voxel_indices[v + 1] : n;
        float acc = 0.0f;
        for (int64_t i = s; i < t; ++i) acc = __fadd_rn(acc, PC[(int64_t)idx[i] * ld + 3 + c]);   // scatter_add_, sorted order
        float mean = __fdiv_rn(acc, (float)(t - s));
        Cvox[v * ld_out + col_out + c] = mean;
        if (DeltaC)
            for (int64_t i = s; i < t; ++i)
                DeltaC[i * ld_delta + 3 + c] = __fsub_rn(PC[(int64_t)idx[i] * ld + 3 + c], mean);
    }
}

RAHT_API int raht_voxel_attr_mean_f32(const float *PC, int64_t n, int64_t ld, int d, const uint32_t *idx,
                                      const int64_t *voxel_indices, int64_t nvox, float *Cvox, int64_t ld_out,
                                      int col_out, float *DeltaC, int64_t ld_delta, void *stream)
{
    RAHT_REQUIRE(n >= 0 && nvox >= 0 && d >= 0 && ld >= 3 + d, "raht_voxel_attr_mean_f32: bad arguments");
    if (nvox == 0 || d == 0) return RAHT_OK;
    RAHT_REQUIRE(PC && idx && voxel_indices && Cvox, "raht_voxel_attr_mean_f32: null pointer");
    (raht_count_launches(1), attr_mean_kernel)<<<grid_for(nvox * d, 256, 8), 256, 0, (cudaStream_t)stream>>>(
        PC, n, ld, d, idx, voxel_indices, nvox, Cvox, ld_out, col_out, DeltaC, ld_delta);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}
