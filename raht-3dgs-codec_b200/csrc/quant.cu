// quant.cu -- uniform quantisation, coefficient reorder and channel marshalling (SURVEY 8a rows
// a9-a11, a13).  Replaces python/encode_3dgs.py:204 (floor(Coeff/step+0.5)), :210
// (index_select(order_RAGFT)), :215-222 (int32 cast + one Python list per channel), the mirror
// :255-268 (stack, *step, argsort(order) gather) and the PSNR reduction :298-310.
//
// One kernel each way: rows are gathered by order_RAGFT (whole 8*D-byte rows, coalesced), quantised
// in fp64 exactly as torch does (true division, add 0.5, floor), transposed through shared memory
// and written channel-major so that each channel is one contiguous int32 sequence for the entropy
// coder.  Algorithmic traffic: D*e + 8 B read, 4*D B written per row.
#include "common.cuh"
#include <atomic>
#include <math.h>

namespace {

constexpr int QR_THREADS = 256;
constexpr int QR_WARPS = QR_THREADS / 32;
constexpr int QR_P = 128;              // positions per CTA
constexpr int QR_PITCH = QR_P + 1;     // odd pitch: conflict-free transposition

// A warp takes the positions j = warp, warp + 8, ... of the CTA's 128 (QR_ROWS = 16 of them).  Their rows are known up
// front: lane i reads order[] for the i-th of them (one round trip instead of one per row), and -- NS > 0: D <= 32 NS,
// the row held in NS registers per lane -- the rows are gathered QR_BATCH at a time before the first is quantised, so
// a warp keeps QR_BATCH x D x 8 bytes in flight (one row at a time, the kernel moved 2.3 TB/s and was bound by the
// order -> row round trips, not by the bandwidth).  NS = 0: any D, a row at a time.
constexpr int QR_ROWS = QR_P / QR_WARPS;
constexpr int QR_BATCH = 4;

// RECIP: multiply by the correctly rounded reciprocal of the step instead of dividing -- what torch evaluates on CUDA for
// `tensor / python_scalar` (ATen div_true_kernel_cuda: a CPU-scalar divisor becomes a multiplication by 1 / b), i.e. the
// reference's floor(Coeff / step + 0.5) on its production device; identical for power-of-two steps.
template <typename Tel, int NS, int RECIP>
__global__ void __launch_bounds__(QR_THREADS)
quant_reorder_kernel(const Tel *__restrict__ T, int64_t n, int D, const int64_t *__restrict__ order,
                     const double *__restrict__ steps, int32_t *__restrict__ q, int64_t ld, int32_t *status)
{
    extern __shared__ __align__(16) unsigned char qr_smem[];
    pdl_enter();
    int32_t *s_q = (int32_t *)qr_smem;                       // [D][QR_PITCH]
    double *s_step = (double *)(s_q + (size_t)D * QR_PITCH + ((D * QR_PITCH) & 1));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t p0 = (int64_t)blockIdx.x * QR_P;
    const int np = (int)min((int64_t)QR_P, n - p0);
    int64_t my_row = -1;                                     // row of the warp's lane-th position
    if (lane < QR_ROWS && warp + QR_WARPS * lane < np) {
        const int64_t p = p0 + warp + QR_WARPS * lane;
        my_row = order ? order[p] : p;
    }
    for (int c = tid; c < D; c += QR_THREADS) s_step[c] = RECIP ? __ddiv_rn(1.0, steps[c]) : steps[c];
    __syncthreads();
    bool ovf = false;
    if constexpr (NS > 0) {
        double st[NS];
#pragma unroll
        for (int k = 0; k < NS; ++k) st[k] = lane + 32 * k < D ? s_step[lane + 32 * k] : 1.0;
#pragma unroll
        for (int i0 = 0; i0 < QR_ROWS; i0 += QR_BATCH) {
            Tel v[QR_BATCH][NS];
#pragma unroll
            for (int r = 0; r < QR_BATCH; ++r) {
                const int64_t row = __shfl_sync(~0u, my_row, i0 + r);
#pragma unroll
                for (int k = 0; k < NS; ++k)
                    if (row >= 0 && lane + 32 * k < D) v[r][k] = T[(size_t)row * D + lane + 32 * k];
            }
#pragma unroll
            for (int r = 0; r < QR_BATCH; ++r) {
                const int j = warp + QR_WARPS * (i0 + r);
                if (j >= np) continue;                                                    // warp-uniform
#pragma unroll
                for (int k = 0; k < NS; ++k) {
                    const int c = lane + 32 * k;
                    if (c < D) {
                        const double sc = RECIP ? __dmul_rn((double)v[r][k], st[k]) : __ddiv_rn((double)v[r][k], st[k]);
                        double qq = floor(__dadd_rn(sc, 0.5));                                    // encode_3dgs.py:204
                        ovf |= !(qq >= -2147483648.0 && qq <= 2147483647.0);
                        s_q[c * QR_PITCH + j] = (int32_t)qq;                              // :215 .to(torch.int32)
                    }
                }
            }
        }
    } else {
        for (int i = 0; i < QR_ROWS; ++i) {
            const int j = warp + QR_WARPS * i;
            const int64_t row = __shfl_sync(~0u, my_row, i);
            if (j >= np) continue;
            const Tel *src = T + (size_t)row * D;
            for (int c = lane; c < D; c += 32) {
                double v = (double)src[c];
                double qq = floor(__dadd_rn(RECIP ? __dmul_rn(v, s_step[c]) : __ddiv_rn(v, s_step[c]), 0.5));
                ovf |= !(qq >= -2147483648.0 && qq <= 2147483647.0);
                s_q[c * QR_PITCH + j] = (int32_t)qq;
            }
        }
    }
    if (__any_sync(~0u, ovf) && lane == 0 && status) atomicOr(status, RAHT_ST_INT32_OVERFLOW);
    __syncthreads();
    const int total = D * QR_P;
    for (int i = tid; i < total; i += QR_THREADS) {
        int c = i / QR_P, j = i - c * QR_P;
        if (j < np) q[(size_t)c * ld + p0 + j] = s_q[c * QR_PITCH + j];
    }
}

template <typename Tel>
__global__ void __launch_bounds__(QR_THREADS)
dequant_unreorder_kernel(const int32_t *__restrict__ q, int64_t ld, int64_t n, int D, const int64_t *__restrict__ order,
                         const double *__restrict__ steps, Tel *__restrict__ T)
{
    extern __shared__ __align__(16) unsigned char qr_smem[];
    int32_t *s_q = (int32_t *)qr_smem;
    double *s_step = (double *)(s_q + (size_t)D * QR_PITCH + ((D * QR_PITCH) & 1));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t p0 = (int64_t)blockIdx.x * QR_P;
    const int np = (int)min((int64_t)QR_P, n - p0);
    int64_t my_row = -1;                                     // row of the warp's lane-th position (one round trip for all 16)
    if (lane < QR_ROWS && warp + QR_WARPS * lane < np) {
        const int64_t p = p0 + warp + QR_WARPS * lane;
        my_row = order ? order[p] : p;
    }
    for (int c = tid; c < D; c += QR_THREADS) s_step[c] = steps[c];
    const int total = D * QR_P;
    constexpr int LD_BATCH = 8;                              // symbol loads in flight per thread
    for (int i0 = tid; i0 < total; i0 += LD_BATCH * QR_THREADS) {
        int32_t v[LD_BATCH];
#pragma unroll
        for (int u = 0; u < LD_BATCH; ++u) {
            const int i = i0 + u * QR_THREADS;
            const int c = i / QR_P, j = i - c * QR_P;
            if (i < total && j < np) v[u] = q[(size_t)c * ld + p0 + j];
        }
#pragma unroll
        for (int u = 0; u < LD_BATCH; ++u) {
            const int i = i0 + u * QR_THREADS;
            const int c = i / QR_P, j = i - c * QR_P;
            if (i < total && j < np) s_q[c * QR_PITCH + j] = v[u];
        }
    }
    __syncthreads();
    for (int i = 0; i < QR_ROWS; ++i) {
        const int j = warp + QR_WARPS * i;
        const int64_t row = __shfl_sync(~0u, my_row, i);
        if (j >= np) continue;                                                    // warp-uniform
        Tel *dst = T + (size_t)row * D;
        for (int c = lane; c < D; c += 32)
            dst[c] = (Tel)__dmul_rn((double)s_q[c * QR_PITCH + j], s_step[c]);    // encode_3dgs.py:261
    }
}

__global__ void __launch_bounds__(256) quantize_kernel(const double *__restrict__ T, int64_t count, double step,
                                                       double *__restrict__ q)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x)
        q[i] = floor(__dadd_rn(__ddiv_rn(T[i], step), 0.5));
}
__global__ void __launch_bounds__(256) dequantize_kernel(const double *__restrict__ q, int64_t count, double step,
                                                         double *__restrict__ T)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x)
        T[i] = __dmul_rn(q[i], step);
}

// ---- sum of squared differences -----------------------------------------------------------------
constexpr int SSE_BLOCKS_MAX = 1184;      // 148 SMs x 8

template <typename TA, typename TB>
__global__ void __launch_bounds__(256) sse_partial_kernel(const TA *__restrict__ A, const TB *__restrict__ B,
                                                          int64_t n, int D, int c0, int c1, double *partial)
{
    const int wc = c1 - c0;
    const int64_t total = n * wc;
    double acc = 0.0;
    if (wc == D) {                                   // all columns: the matrices are two flat arrays
        for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
            const double d = (double)A[e] - (double)B[e];
            acc += d * d;
        }
    } else {
        for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
            int64_t r = e / wc;
            int c = c0 + (int)(e - r * wc);
            const double d = (double)A[r * D + c] - (double)B[r * D + c];
            acc += d * d;
        }
    }
    __shared__ double s[8];
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(~0u, acc, o);
    if (lane_id() == 0) s[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += s[w];
        partial[blockIdx.x] = t;
    }
}
__global__ void __launch_bounds__(256) sse_final_kernel(const double *partial, int nb, double *out)
{
    double acc = 0.0;
    for (int i = threadIdx.x; i < nb; i += 256) acc += partial[i];
    __shared__ double s[8];
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(~0u, acc, o);
    if (lane_id() == 0) s[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += s[w];
        *out = t;
    }
}

// ---- per-channel minimum and maximum (coefficient range for the per-attribute steps, encode_3dgs_debug.py:353-355) ----
constexpr int MM_BLOCKS_MAX = 592;        // 148 SMs x 4
constexpr int MM_ROWS = 4;                // rows in flight per CTA pass (blockDim.y)

template <typename Tel>
__global__ void __launch_bounds__(256) minmax_partial_kernel(const Tel *__restrict__ T, int64_t n, int D, double *__restrict__ partial)
{
    extern __shared__ double mm_s[];                             // [MM_ROWS][2][D]
    const int cx = threadIdx.x, ry = threadIdx.y, cw = blockDim.x;
    for (int c0 = 0; c0 < D; c0 += cw) {
        const int c = c0 + cx;
        double lo = INFINITY, hi = -INFINITY;
        if (c < D)
            for (int64_t r = (int64_t)blockIdx.x * MM_ROWS + ry; r < n; r += (int64_t)gridDim.x * MM_ROWS) {
                const double v = (double)T[(size_t)r * D + c];
                lo = fmin(lo, v);
                hi = fmax(hi, v);
            }
        if (c < D) { mm_s[(ry * 2 + 0) * D + c] = lo; mm_s[(ry * 2 + 1) * D + c] = hi; }
    }
    __syncthreads();
    for (int c = threadIdx.y * blockDim.x + threadIdx.x; c < D; c += blockDim.x * blockDim.y) {
        double lo = INFINITY, hi = -INFINITY;
        for (int y = 0; y < MM_ROWS; ++y) { lo = fmin(lo, mm_s[(y * 2 + 0) * D + c]); hi = fmax(hi, mm_s[(y * 2 + 1) * D + c]); }
        partial[((size_t)blockIdx.x * 2 + 0) * D + c] = lo;
        partial[((size_t)blockIdx.x * 2 + 1) * D + c] = hi;
    }
}
__global__ void __launch_bounds__(256) minmax_final_kernel(const double *__restrict__ partial, int nb, int D, double *__restrict__ out)
{
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < D; c += gridDim.x * blockDim.x) {
        double lo = INFINITY, hi = -INFINITY;
        for (int b = 0; b < nb; ++b) { lo = fmin(lo, partial[((size_t)b * 2 + 0) * D + c]); hi = fmax(hi, partial[((size_t)b * 2 + 1) * D + c]); }
        out[c] = lo;
        out[D + c] = hi;
    }
}

size_t qr_smem_bytes(int D) { return ((size_t)D * QR_PITCH + 1) * 4 + (size_t)D * 8 + 16; }

}  // namespace

template <typename Tel, int RECIP>
static int quant_reorder_launch(const Tel *T, int64_t n, int D, const int64_t *order, const double *steps, int32_t *q, int64_t ld,
                                int32_t *status, size_t smem, cudaStream_t st)
{
    unsigned grid = (unsigned)((n + QR_P - 1) / QR_P);
    auto kern = D <= 32 ? quant_reorder_kernel<Tel, 1, RECIP> : D <= 64 ? quant_reorder_kernel<Tel, 2, RECIP> : quant_reorder_kernel<Tel, 0, RECIP>;
    RAHT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    if (raht_use_pdl(PDL_QUANT)) RAHT_CUDA(pdl_launch(kern, dim3(grid), dim3(QR_THREADS), smem, st, T, n, D, order, steps, q, ld, status));
    else kern<<<grid, QR_THREADS, smem, st>>>(T, n, D, order, steps, q, ld, status);
    return RAHT_OK;
}

RAHT_API int raht_quant_reorder(const void *T, int t_dtype, int64_t n, int D, const int64_t *order, const double *steps,
                                int32_t *q, int64_t ld, int32_t *status, int flags, void *stream)
{
    RAHT_REQUIRE(n >= 0 && D >= 1 && ld >= n && steps, "raht_quant_reorder: bad arguments");
    RAHT_REQUIRE((flags & ~RAHT_QUANT_RECIPROCAL) == 0, "raht_quant_reorder: unknown flags");
    if (n == 0) return RAHT_OK;
    RAHT_REQUIRE(T && q, "raht_quant_reorder: null pointer");
    size_t smem = qr_smem_bytes(D);
    RAHT_REQUIRE(smem <= 227 * 1024, "raht_quant_reorder: D too large");
    cudaStream_t st = (cudaStream_t)stream;
    const bool recip = (flags & RAHT_QUANT_RECIPROCAL) != 0;
    raht_count_launches(1);
    int rc;
    if (t_dtype == RAHT_F64) {
        rc = recip ? quant_reorder_launch<double, 1>((const double *)T, n, D, order, steps, q, ld, status, smem, st)
                   : quant_reorder_launch<double, 0>((const double *)T, n, D, order, steps, q, ld, status, smem, st);
    } else if (t_dtype == RAHT_F32) {
        rc = recip ? quant_reorder_launch<float, 1>((const float *)T, n, D, order, steps, q, ld, status, smem, st)
                   : quant_reorder_launch<float, 0>((const float *)T, n, D, order, steps, q, ld, status, smem, st);
    } else {
        raht_set_error("raht_quant_reorder: bad dtype");
        return RAHT_E_INVALID;
    }
    if (rc != RAHT_OK) return rc;
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_dequant_unreorder(const int32_t *q, int64_t ld, int64_t n, int D, const int64_t *order,
                                    const double *steps, void *T, int t_dtype, void *stream)
{
    RAHT_REQUIRE(n >= 0 && D >= 1 && ld >= n && steps, "raht_dequant_unreorder: bad arguments");
    if (n == 0) return RAHT_OK;
    RAHT_REQUIRE(T && q, "raht_dequant_unreorder: null pointer");
    size_t smem = qr_smem_bytes(D);
    RAHT_REQUIRE(smem <= 227 * 1024, "raht_dequant_unreorder: D too large");
    cudaStream_t st = (cudaStream_t)stream;
    unsigned grid = (unsigned)((n + QR_P - 1) / QR_P);
    raht_count_launches(1);
    if (t_dtype == RAHT_F64) {
        RAHT_CUDA(cudaFuncSetAttribute(dequant_unreorder_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        dequant_unreorder_kernel<double><<<grid, QR_THREADS, smem, st>>>(q, ld, n, D, order, steps, (double *)T);
    } else if (t_dtype == RAHT_F32) {
        RAHT_CUDA(cudaFuncSetAttribute(dequant_unreorder_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        dequant_unreorder_kernel<float><<<grid, QR_THREADS, smem, st>>>(q, ld, n, D, order, steps, (float *)T);
    } else {
        raht_set_error("raht_dequant_unreorder: bad dtype");
        return RAHT_E_INVALID;
    }
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_quantize_f64(const double *T, int64_t count, double step, double *q, void *stream)
{
    RAHT_REQUIRE(count >= 0 && (count == 0 || (T && q)), "raht_quantize_f64: bad arguments");
    if (count == 0) return RAHT_OK;
    (raht_count_launches(1), quantize_kernel)<<<grid_for(count, 256, 8), 256, 0, (cudaStream_t)stream>>>(T, count, step, q);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_dequantize_f64(const double *q, int64_t count, double step, double *T, void *stream)
{
    RAHT_REQUIRE(count >= 0 && (count == 0 || (T && q)), "raht_dequantize_f64: bad arguments");
    if (count == 0) return RAHT_OK;
    (raht_count_launches(1), dequantize_kernel)<<<grid_for(count, 256, 8), 256, 0, (cudaStream_t)stream>>>(q, count, step, T);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API size_t raht_sse_workspace_bytes(int64_t n, int D)
{
    (void)n; (void)D;
    return (size_t)SSE_BLOCKS_MAX * 8 + 256;
}

RAHT_API int raht_sse(const void *A, int a_dtype, const void *B, int b_dtype, int64_t n, int D, int c0, int c1, double *sse,
                      void *workspace, size_t workspace_bytes, void *stream)
{
    RAHT_REQUIRE(n >= 0 && D >= 1 && c0 >= 0 && c1 <= D && c0 < c1 && sse && workspace, "raht_sse: bad arguments");
    RAHT_REQUIRE((a_dtype == RAHT_F32 || a_dtype == RAHT_F64) && (b_dtype == RAHT_F32 || b_dtype == RAHT_F64), "raht_sse: bad dtype");
    if (workspace_bytes < raht_sse_workspace_bytes(n, D)) {
        raht_set_error("raht_sse: workspace too small");
        return RAHT_E_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    int nb = grid_for(n * (c1 - c0), 256, 8);
    if (nb > SSE_BLOCKS_MAX) nb = SSE_BLOCKS_MAX;
    if (n == 0) nb = 1;
    double *part = (double *)workspace;
    raht_count_launches(2);
    if (a_dtype == RAHT_F64 && b_dtype == RAHT_F64)
        sse_partial_kernel<double, double><<<nb, 256, 0, st>>>((const double *)A, (const double *)B, n, D, c0, c1, part);
    else if (a_dtype == RAHT_F32 && b_dtype == RAHT_F64)
        sse_partial_kernel<float, double><<<nb, 256, 0, st>>>((const float *)A, (const double *)B, n, D, c0, c1, part);
    else if (a_dtype == RAHT_F64 && b_dtype == RAHT_F32)
        sse_partial_kernel<double, float><<<nb, 256, 0, st>>>((const double *)A, (const float *)B, n, D, c0, c1, part);
    else
        sse_partial_kernel<float, float><<<nb, 256, 0, st>>>((const float *)A, (const float *)B, n, D, c0, c1, part);
    sse_final_kernel<<<1, 256, 0, st>>>((const double *)workspace, nb, sse);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_sse_f64(const double *A, const double *B, int64_t n, int D, int c0, int c1, double *sse, void *workspace,
                          size_t workspace_bytes, void *stream)
{
    return raht_sse(A, RAHT_F64, B, RAHT_F64, n, D, c0, c1, sse, workspace, workspace_bytes, stream);
}

RAHT_API size_t raht_channel_minmax_workspace_bytes(int64_t n, int D)
{
    (void)n;
    return (size_t)MM_BLOCKS_MAX * 2 * (size_t)(D > 0 ? D : 1) * 8 + 256;
}

RAHT_API int raht_channel_minmax(const void *T, int t_dtype, int64_t n, int D, double *minmax, void *workspace, size_t workspace_bytes,
                                 void *stream)
{
    RAHT_REQUIRE(n >= 1 && D >= 1 && T && minmax && workspace, "raht_channel_minmax: bad arguments (n must be >= 1)");
    if (workspace_bytes < raht_channel_minmax_workspace_bytes(n, D)) {
        raht_set_error("raht_channel_minmax: workspace too small");
        return RAHT_E_WORKSPACE;
    }
    const size_t smem = (size_t)MM_ROWS * 2 * D * 8;
    RAHT_REQUIRE(smem <= 48 * 1024, "raht_channel_minmax: D too large");
    cudaStream_t st = (cudaStream_t)stream;
    int nb = (int)((n + MM_ROWS - 1) / MM_ROWS);
    if (nb > MM_BLOCKS_MAX) nb = MM_BLOCKS_MAX;
    const dim3 block(64, MM_ROWS);
    raht_count_launches(2);
    if (t_dtype == RAHT_F64) minmax_partial_kernel<double><<<nb, block, smem, st>>>((const double *)T, n, D, (double *)workspace);
    else if (t_dtype == RAHT_F32) minmax_partial_kernel<float><<<nb, block, smem, st>>>((const float *)T, n, D, (double *)workspace);
    else {
        raht_set_error("raht_channel_minmax: bad dtype");
        return RAHT_E_INVALID;
    }
    minmax_final_kernel<<<(D + 255) / 256, 256, 0, st>>>((const double *)workspace, nb, D, minmax);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

// ---- no-allocation variant ------------------------------------------------------------------------------------------
// raht_channel_minmax_static: the same two kernels as raht_channel_minmax, with the per-CTA partial minima / maxima in a
// file-scope device array instead of a caller workspace, for callers that cannot allocate (code under CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Every GPU that loads the library holds the array (592 KiB) until exit.  All calls on a device
// share it: calls that overlap on the GPU (different streams, no ordering) race on it.  Nothing in it has to start at
// zero -- the first kernel overwrites every entry [0, nb) x 2 x D that the second one reads -- so no clear is issued.
#define RAHT_MINMAX_STATIC_MAX_D_ 64
static __device__ __align__(256) double g_minmax_scratch[(size_t)MM_BLOCKS_MAX * 2 * RAHT_MINMAX_STATIC_MAX_D_];

namespace {
constexpr int STATIC_MAX_DEVICES = 64;
std::atomic<double *> g_minmax_scratch_addr[STATIC_MAX_DEVICES];

// g_minmax_scratch is, in host code, a host-side stand-in: the device address is asked for once per device, with
// that device current (the caller's), and kept.
int minmax_scratch_for_current_device(double **out)
{
    int dev = -1;
    RAHT_CUDA(cudaGetDevice(&dev));
    RAHT_REQUIRE(dev >= 0 && dev < STATIC_MAX_DEVICES, "raht_channel_minmax_static: device ordinal out of range");
    double *p = g_minmax_scratch_addr[dev].load(std::memory_order_acquire);
    if (!p) {
        void *sym = nullptr;
        RAHT_CUDA(cudaGetSymbolAddress(&sym, g_minmax_scratch));
        RAHT_REQUIRE(sym != nullptr, "raht_channel_minmax_static: cudaGetSymbolAddress returned null");
        p = (double *)sym;
        g_minmax_scratch_addr[dev].store(p, std::memory_order_release);
    }
    *out = p;
    return RAHT_OK;
}
}  // namespace

RAHT_API int raht_channel_minmax_static_max_channels(void) { return RAHT_MINMAX_STATIC_MAX_D_; }

RAHT_API int raht_channel_minmax_static(const void *T, int t_dtype, int64_t n, int D, double *minmax, void *stream)
{
    RAHT_REQUIRE(n >= 1 && D >= 1 && T && minmax, "raht_channel_minmax_static: bad arguments (n must be >= 1)");
    if (D > RAHT_MINMAX_STATIC_MAX_D_) {
        raht_set_error("raht_channel_minmax_static: D exceeds the static scratch (64 channels at most); use raht_channel_minmax with a workspace");
        return RAHT_E_WORKSPACE;
    }
    RAHT_REQUIRE(t_dtype == RAHT_F64 || t_dtype == RAHT_F32, "raht_channel_minmax_static: bad dtype");
    double *scratch = nullptr;
    int rc = minmax_scratch_for_current_device(&scratch);
    if (rc != RAHT_OK) return rc;
    const size_t smem = (size_t)MM_ROWS * 2 * D * 8;                       // <= 4 KiB at D = 64
    cudaStream_t st = (cudaStream_t)stream;
    int nb = (int)((n + MM_ROWS - 1) / MM_ROWS);
    if (nb > MM_BLOCKS_MAX) nb = MM_BLOCKS_MAX;
    const dim3 block(64, MM_ROWS);
    raht_count_launches(2);
    if (t_dtype == RAHT_F64) minmax_partial_kernel<double><<<nb, block, smem, st>>>((const double *)T, n, D, scratch);
    else minmax_partial_kernel<float><<<nb, block, smem, st>>>((const float *)T, n, D, scratch);
    minmax_final_kernel<<<(D + 255) / 256, 256, 0, st>>>(scratch, nb, D, minmax);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}
