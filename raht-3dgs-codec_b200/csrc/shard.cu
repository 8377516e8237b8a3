// shard.cu -- the top of the RAHT tree across Morton-prefix shards (the reference has no
// multi-GPU path, this is the part of python/RAHT.py:253-336 / python/iRAHT.py:40-114 that joins the
// shards).  A scene is cut into P contiguous ranges of its Morton-sorted voxels (one per GPU); every
// range is transformed on its own as if it were a whole cloud, which runs every butterfly whose rows
// lie inside the range and leaves the range's low-pass row (its first row, weight = rows of the
// range).  The butterflies that remain join those P rows: row i is merged, at the highest bit in
// which the first Morton code of range i differs from the last code before it, with the nearest
// earlier range of a higher level -- the same max-Cartesian tree as plan.cuh, with the weights taken
// from the ranges' global row offsets instead of 1 per row.  Same arithmetic as transform.cu
// (separately rounded products and sums), so the coefficients are bit-identical to the single-GPU
// transform of the whole scene.
#include "common.cuh"

namespace {

constexpr int TOP_MAX = 1024;

template <typename Tel> struct TopArith;
template <> struct TopArith<double> {
    static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
    static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
    static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
};
template <> struct TopArith<float> {
    static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
    static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
    static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
};

// level[i] (i >= 1) = highest bit in which prev_last[i] (the last Morton code before range i) and first[i]
// differ; offsets[P + 1] = global index of each range's first row, offsets[P] = rows of the scene.
template <typename Tel>
__global__ void __launch_bounds__(256) shard_top_kernel(const int64_t *__restrict__ first, const int64_t *__restrict__ prev_last,
                                                        const int64_t *__restrict__ offsets, int P, Tel *__restrict__ rows, int D,
                                                        int inverse, uint8_t *__restrict__ level_out, int32_t *status)
{
    __shared__ uint8_t lev[TOP_MAX];
    __shared__ int16_t left[TOP_MAX], right[TOP_MAX];
    __shared__ double ca[TOP_MAX], cb[TOP_MAX];
    const int tid = threadIdx.x;
    for (int i = tid; i < P; i += blockDim.x) {
        if (i == 0) { lev[0] = 255; continue; }
        const int64_t a = prev_last[i], b = first[i];
        if (!(a < b) && status) atomicOr(status, RAHT_ST_NOT_SORTED_UNIQUE);
        const uint64_t x = (uint64_t)a ^ (uint64_t)b;
        lev[i] = x ? (uint8_t)(63 - __clzll((long long)x)) : (uint8_t)0;
    }
    __syncthreads();
    for (int i = tid; i < P; i += blockDim.x) {
        if (level_out) level_out[i] = lev[i];
        if (i == 0) continue;
        int s = 0, t = P;
        for (int j = i - 1; j >= 0; --j) if (lev[j] > lev[i]) { s = j; break; }
        for (int j = i + 1; j < P; ++j) if (lev[j] > lev[i]) { t = j; break; }
        left[i] = (int16_t)s;
        right[i] = (int16_t)t;
        const double w0 = (double)(offsets[i] - offsets[s]), w1 = (double)(offsets[t] - offsets[i]);
        const double den = __dadd_rn(w0, w1);                                   // RAHT.py:317
        ca[i] = __dsqrt_rn(__ddiv_rn(w0, den));                                 // :321
        cb[i] = __dsqrt_rn(__ddiv_rn(w1, den));                                 // :322
    }
    __syncthreads();
    using AR = TopArith<Tel>;
    // every thread owns channels tid, tid + 256, ...: the butterflies of a channel run in level order without barriers
    for (int c = tid; c < D; c += blockDim.x) {
        for (int ll = 0; ll < 63; ++ll) {
            const int l = inverse ? 62 - ll : ll;
            for (int i = 1; i < P; ++i) {
                if (lev[i] != l) continue;
                Tel *p0 = rows + (size_t)left[i] * D + c, *p1 = rows + (size_t)i * D + c;
                const Tel a = (Tel)ca[i], b = (Tel)cb[i], x0 = *p0, x1 = *p1;
                if (!inverse) {
                    *p0 = AR::add(AR::mul(a, x0), AR::mul(b, x1));             // RAHT.py:331
                    *p1 = AR::add(AR::mul(-b, x0), AR::mul(a, x1));            // RAHT.py:332
                } else {
                    *p0 = AR::sub(AR::mul(a, x0), AR::mul(b, x1));             // iRAHT.py:108
                    *p1 = AR::add(AR::mul(b, x0), AR::mul(a, x1));             // iRAHT.py:109
                }
            }
        }
    }
}

// The same join over CELLS (aligned Morton-prefix subtrees), some of which may be empty: range i holds counts[i] rows,
// first[i] / last[i] are its first and last Morton code.  Empty cells take no part.  level_out[i]: level of the cell's
// first row in the whole scene (255 for the first non-empty cell, 0 for empty cells).
template <typename Tel>
__global__ void __launch_bounds__(256) shard_top_cells_kernel(const int64_t *__restrict__ first, const int64_t *__restrict__ last,
                                                              const int64_t *__restrict__ counts, int C, Tel *__restrict__ rows, int D,
                                                              int inverse, uint8_t *__restrict__ level_out, int32_t *status)
{
    __shared__ uint8_t lev[TOP_MAX], live[TOP_MAX];
    __shared__ int16_t left[TOP_MAX], prevne[TOP_MAX];
    __shared__ double ca[TOP_MAX], cb[TOP_MAX];
    __shared__ long long off[TOP_MAX + 1];
    const int tid = threadIdx.x;
    if (tid == 0) {                                        // C <= 1024: offsets and the previous non-empty cell, serially
        long long run = 0;
        int p = -1;
        for (int i = 0; i < C; ++i) {
            off[i] = run;
            prevne[i] = (int16_t)p;
            const long long c = counts[i];
            live[i] = c > 0;
            if (c > 0) { run += c; p = i; }
        }
        off[C] = run;
    }
    __syncthreads();
    for (int i = tid; i < C; i += blockDim.x) {
        uint8_t l = 0;
        if (live[i]) {
            if (prevne[i] < 0) l = 255;
            else {
                const int64_t a = last[prevne[i]], b = first[i];
                if (!(a < b) && status) atomicOr(status, RAHT_ST_NOT_SORTED_UNIQUE);
                const uint64_t x = (uint64_t)a ^ (uint64_t)b;
                l = x ? (uint8_t)(63 - __clzll((long long)x)) : (uint8_t)0;
            }
        }
        lev[i] = l;
    }
    __syncthreads();
    for (int i = tid; i < C; i += blockDim.x) {
        if (level_out) level_out[i] = lev[i];
        left[i] = -1;
        if (!live[i] || lev[i] == 255) continue;
        int s = -1, t = C;
        for (int j = i - 1; j >= 0; --j) if (live[j] && lev[j] > lev[i]) { s = j; break; }
        for (int j = i + 1; j < C; ++j) if (live[j] && lev[j] > lev[i]) { t = j; break; }
        left[i] = (int16_t)s;                              // (s >= 0: the first non-empty cell has level 255)
        const double w0 = (double)(off[i] - off[s]), w1 = (double)(off[t] - off[i]);
        const double den = __dadd_rn(w0, w1);                                   // RAHT.py:317
        ca[i] = __dsqrt_rn(__ddiv_rn(w0, den));                                 // :321
        cb[i] = __dsqrt_rn(__ddiv_rn(w1, den));                                 // :322
    }
    __syncthreads();
    using AR = TopArith<Tel>;
    for (int c = tid; c < D; c += blockDim.x) {
        for (int ll = 0; ll < 63; ++ll) {
            const int l = inverse ? 62 - ll : ll;
            for (int i = 1; i < C; ++i) {
                if (left[i] < 0 || lev[i] != l) continue;
                Tel *p0 = rows + (size_t)left[i] * D + c, *p1 = rows + (size_t)i * D + c;
                const Tel a = (Tel)ca[i], b = (Tel)cb[i], x0 = *p0, x1 = *p1;
                if (!inverse) {
                    *p0 = AR::add(AR::mul(a, x0), AR::mul(b, x1));             // RAHT.py:331
                    *p1 = AR::add(AR::mul(-b, x0), AR::mul(a, x1));            // RAHT.py:332
                } else {
                    *p0 = AR::sub(AR::mul(a, x0), AR::mul(b, x1));             // iRAHT.py:108
                    *p1 = AR::add(AR::mul(b, x0), AR::mul(a, x1));             // iRAHT.py:109
                }
            }
        }
    }
}

// Gaussians as packed rows [means3 quats4 scales3 opacity1 colors] for the repartition exchange: out[i] = row perm[i]
// (perm == NULL: identity); and back.  A warp per row.
__global__ void __launch_bounds__(256) pack_rows_kernel(const uint32_t *__restrict__ perm, int64_t n, const float *__restrict__ means,
                                                        const float *__restrict__ quats, const float *__restrict__ scales,
                                                        const float *__restrict__ opac, const float *__restrict__ colors, int cd,
                                                        float *__restrict__ out)
{
    const int lane = threadIdx.x & 31;
    const int D = 11 + cd;
    for (int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); i < n; i += (int64_t)gridDim.x * 8) {
        const size_t r = perm ? (size_t)perm[i] : (size_t)i;
        float *dst = out + (size_t)i * D;
        for (int c = lane; c < D; c += 32) {
            float v;
            if (c < 3) v = means[r * 3 + c];
            else if (c < 7) v = quats[r * 4 + c - 3];
            else if (c < 10) v = scales[r * 3 + c - 7];
            else if (c == 10) v = opac[r];
            else v = colors[r * cd + c - 11];
            dst[c] = v;
        }
    }
}
__global__ void __launch_bounds__(256) unpack_rows_kernel(const float *__restrict__ in, int64_t n, int cd, float *__restrict__ means,
                                                          float *__restrict__ quats, float *__restrict__ scales, float *__restrict__ opac,
                                                          float *__restrict__ colors)
{
    const int lane = threadIdx.x & 31;
    const int D = 11 + cd;
    for (int64_t i = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); i < n; i += (int64_t)gridDim.x * 8) {
        const float *src = in + (size_t)i * D;
        for (int c = lane; c < D; c += 32) {
            const float v = src[c];
            if (c < 3) means[(size_t)i * 3 + c] = v;
            else if (c < 7) quats[(size_t)i * 4 + c - 3] = v;
            else if (c < 10) scales[(size_t)i * 3 + c - 7] = v;
            else if (c == 10) opac[i] = v;
            else colors[(size_t)i * cd + c - 11] = v;
        }
    }
}

// Histogram of (v[i] >> shift) over a few bins (cells of the sharded partition, owner ranks): per-CTA counters in
// shared memory, a warp adds one count per distinct bin (match_any) -- a global scatter_add over 2 .. 1024 bins serialises
// 25 M atomics on a handful of addresses (measured 5-10 ms per histogram in the first sharded version).
constexpr int HIST_MAX_BINS = 4096;
__global__ void __launch_bounds__(256) histogram_kernel(const int64_t *__restrict__ v, int64_t n, int shift, int nbins,
                                                        unsigned long long *__restrict__ out)
{
    extern __shared__ unsigned int h[];
    for (int b = threadIdx.x; b < nbins; b += blockDim.x) h[b] = 0u;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    for (int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x - lane); i0 < n; i0 += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = i0 + lane;
        const bool ok = i < n;
        int b = ok ? (int)((uint64_t)v[i] >> shift) : -1;
        if (b >= nbins) b = nbins - 1;
        const unsigned peers = __match_any_sync(~0u, b);
        if (ok && lane == __ffs((int)peers) - 1) atomicAdd(&h[b], (unsigned)__popc(peers));
    }
    __syncthreads();
    for (int b = threadIdx.x; b < nbins; b += blockDim.x)
        if (h[b]) atomicAdd(out + b, (unsigned long long)h[b]);
}

}  // namespace

RAHT_API int raht_shard_top(const int64_t *first_keys, const int64_t *prev_last_keys, const int64_t *offsets, int P,
                            void *rows, int dtype, int D, int inverse, uint8_t *level_out, int32_t *status, void *stream)
{
    RAHT_REQUIRE(P >= 0 && P <= TOP_MAX && D >= 1, "raht_shard_top: P must be in [0, 1024]");
    if (P == 0) return RAHT_OK;
    RAHT_REQUIRE(first_keys && prev_last_keys && offsets && rows, "raht_shard_top: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    raht_count_launches(1);
    if (dtype == RAHT_F64)
        shard_top_kernel<double><<<1, 256, 0, st>>>(first_keys, prev_last_keys, offsets, P, (double *)rows, D, inverse, level_out, status);
    else if (dtype == RAHT_F32)
        shard_top_kernel<float><<<1, 256, 0, st>>>(first_keys, prev_last_keys, offsets, P, (float *)rows, D, inverse, level_out, status);
    else {
        raht_set_error("raht_shard_top: bad dtype");
        return RAHT_E_INVALID;
    }
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_shard_top_cells(const int64_t *first_keys, const int64_t *last_keys, const int64_t *counts, int C, void *rows,
                                  int dtype, int D, int inverse, uint8_t *level_out, int32_t *status, void *stream)
{
    RAHT_REQUIRE(C >= 0 && C <= TOP_MAX && D >= 1, "raht_shard_top_cells: C must be in [0, 1024]");
    if (C == 0) return RAHT_OK;
    RAHT_REQUIRE(first_keys && last_keys && counts && rows, "raht_shard_top_cells: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    raht_count_launches(1);
    if (dtype == RAHT_F64)
        shard_top_cells_kernel<double><<<1, 256, 0, st>>>(first_keys, last_keys, counts, C, (double *)rows, D, inverse, level_out, status);
    else if (dtype == RAHT_F32)
        shard_top_cells_kernel<float><<<1, 256, 0, st>>>(first_keys, last_keys, counts, C, (float *)rows, D, inverse, level_out, status);
    else {
        raht_set_error("raht_shard_top_cells: bad dtype");
        return RAHT_E_INVALID;
    }
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_pack_rows(const uint32_t *perm, int64_t n, const float *means, const float *quats, const float *scales,
                            const float *opacities, const float *colors, int color_dim, float *out, void *stream)
{
    RAHT_REQUIRE(n >= 0 && color_dim >= 0, "raht_pack_rows: bad arguments");
    if (n == 0) return RAHT_OK;
    RAHT_REQUIRE(means && quats && scales && opacities && (colors || color_dim == 0) && out, "raht_pack_rows: null pointer");
    raht_count_launches(1);
    pack_rows_kernel<<<grid_for(n, 8, 8), 256, 0, (cudaStream_t)stream>>>(perm, n, means, quats, scales, opacities, colors, color_dim, out);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_unpack_rows(const float *rows, int64_t n, int color_dim, float *means, float *quats, float *scales, float *opacities,
                              float *colors, void *stream)
{
    RAHT_REQUIRE(n >= 0 && color_dim >= 0, "raht_unpack_rows: bad arguments");
    if (n == 0) return RAHT_OK;
    RAHT_REQUIRE(rows && means && quats && scales && opacities && (colors || color_dim == 0), "raht_unpack_rows: null pointer");
    raht_count_launches(1);
    unpack_rows_kernel<<<grid_for(n, 8, 8), 256, 0, (cudaStream_t)stream>>>(rows, n, color_dim, means, quats, scales, opacities, colors);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_histogram_i64(const int64_t *values, int64_t n, int shift, int nbins, int64_t *counts, void *stream)
{
    RAHT_REQUIRE(n >= 0 && shift >= 0 && shift < 64 && nbins >= 1 && nbins <= HIST_MAX_BINS && counts, "raht_histogram_i64: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    RAHT_CUDA(cudaMemsetAsync(counts, 0, (size_t)nbins * 8, st));
    if (n == 0) return RAHT_OK;
    RAHT_REQUIRE(values, "raht_histogram_i64: null pointer");
    raht_count_launches(1);
    histogram_kernel<<<grid_for(n, 256, 4), 256, (size_t)nbins * 4, st>>>(values, n, shift, nbins, (unsigned long long *)counts);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}
