// merge.cu -- segmented Gaussian merge.
// Replaces cuda/merge_cluster.cu:2-111 (merge_weighted_mean_kernel, one thread per voxel with
// strided gathers and a 48x re-read of the opacities) and cuda/merge_cluster_wrapper.cu:11-116.
//
// Mapping: one thread per (voxel, output channel); the 11+color_dim channels of a voxel sit in
// consecutive lanes, so the gathers of a member's colour row are coalesced and the packed output row
// [means3|quats4|scales3|opacity1|colors] is written contiguously.  Every channel sums its members in
// segment order with separately rounded multiply and add, so the result is a deterministic function
// of the inputs (the reference's --use_fast_math build is not).
#include "common.cuh"
#include <math.h>

namespace {

constexpr int MG_UNROLL = 4;      // (voxel, channel) elements per thread, loads batched level by level

// Channel c of voxel [s, t) given its first member g0 (already fetched: opacity o0, and value x0 / quaternion q0).
// Sums run in segment order with separately rounded multiply and add, starting from 0 as the reference does.
__device__ __forceinline__ float merge_element(int c, int s, int t, float o0, float x0, float4 q0,
                                               const int32_t *__restrict__ cidx, const float *__restrict__ means,
                                               const float *__restrict__ quats, const float *__restrict__ scales,
                                               const float *__restrict__ opac, const float *__restrict__ colors, int color_dim,
                                               int by_opacity)
{
    if (t <= s) return 0.0f;                                             // merge_cluster.cu:27-29
    const float w0 = by_opacity ? o0 : 1.0f;                             // :42
    if (c >= 3 && c < 7) {
        // quaternion: all four weighted sums are needed for the norm (:74-87)
        float q[4] = {__fadd_rn(0.f, __fmul_rn(q0.x, w0)), __fadd_rn(0.f, __fmul_rn(q0.y, w0)),
                      __fadd_rn(0.f, __fmul_rn(q0.z, w0)), __fadd_rn(0.f, __fmul_rn(q0.w, w0))};
        for (int i = s + 1; i < t; ++i) {
            const int g = cidx[i];
            const float w = by_opacity ? opac[g] : 1.0f;
            const float4 qq = *reinterpret_cast<const float4 *>(quats + (size_t)g * 4);
            q[0] = __fadd_rn(q[0], __fmul_rn(qq.x, w));
            q[1] = __fadd_rn(q[1], __fmul_rn(qq.y, w));
            q[2] = __fadd_rn(q[2], __fmul_rn(qq.z, w));
            q[3] = __fadd_rn(q[3], __fmul_rn(qq.w, w));
        }
        const float nn = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(q[0], q[0]), __fmul_rn(q[1], q[1])), __fmul_rn(q[2], q[2])),
                                   __fmul_rn(q[3], q[3]));
        const float qn = __fsqrt_rn(nn);                                 // :75-76
        if (qn > 0.0f) return __fdiv_rn(q[c - 3], qn);
        return (c == 6) ? 1.0f : 0.0f;                                   // fallback (0,0,0,1) :82-86
    }
    if (c == 10) {
        float acc = __fadd_rn(0.0f, o0);
        for (int i = s + 1; i < t; ++i) acc = __fadd_rn(acc, opac[cidx[i]]);
        return fminf(acc, 1.0f);                                         // :94
    }
    const float *src;
    int ld, col;
    if (c < 3) { src = means; ld = 3; col = c; }
    else if (c < 10) { src = scales; ld = 3; col = c - 7; }
    else { src = colors; ld = color_dim; col = c - 11; }
    float acc = __fadd_rn(0.0f, __fmul_rn(x0, w0)), wsum = __fadd_rn(0.0f, w0);
    for (int i = s + 1; i < t; ++i) {
        const int g = cidx[i];
        const float w = by_opacity ? opac[g] : 1.0f;
        acc = __fadd_rn(acc, __fmul_rn(src[(size_t)g * ld + col], w));
        wsum = __fadd_rn(wsum, w);
    }
    if (c < 10) {
        if (wsum == 0.0f) wsum = 1.0f;                                   // :65-67
        return __fdiv_rn(acc, wsum);
    }
    return (wsum > 0.0f) ? __fdiv_rn(acc, wsum) : 0.0f;                  // :108-109
}

// Thread (x, y) of a 64 x 4 block: channel c = x (+64, ... when there are more than 64 channels) of voxel
// blockIdx * 4 + y, MG_UNROLL voxels per thread a grid stride apart.  No integer division anywhere; the offset /
// member-index / opacity loads of a voxel are the same address for its 59 threads (one broadcast transaction), the
// colour row and the packed output row are contiguous across x.
constexpr int MG_TX = 64, MG_TY = 4;

template <typename OutT>
__global__ void __launch_bounds__(MG_TX *MG_TY)
merge_kernel(const int32_t *__restrict__ cidx, const int32_t *__restrict__ coff, int64_t nclu,
             const float *__restrict__ means, const float *__restrict__ quats, const float *__restrict__ scales,
             const float *__restrict__ opac, const float *__restrict__ colors, int color_dim, int by_opacity,
             float *__restrict__ m_means, float *__restrict__ m_quats, float *__restrict__ m_scales,
             float *__restrict__ m_opac, float *__restrict__ m_colors, OutT *__restrict__ packed)
{
    const int OC = 11 + color_dim;
    const int64_t vstride = (int64_t)gridDim.x * MG_TY;
    for (int64_t v0 = (int64_t)blockIdx.x * MG_TY + threadIdx.y; v0 < nclu; v0 += MG_UNROLL * vstride) {
        int s[MG_UNROLL], t[MG_UNROLL], g[MG_UNROLL];
        float o0[MG_UNROLL];
        // the dependent loads (offsets -> member index -> member data) are issued level by level for all voxels
#pragma unroll
        for (int u = 0; u < MG_UNROLL; ++u) {
            const int64_t v = v0 + u * vstride;
            const bool ok = v < nclu;
            s[u] = ok ? coff[v] : 0;
            t[u] = ok ? coff[v + 1] : 0;
        }
#pragma unroll
        for (int u = 0; u < MG_UNROLL; ++u) g[u] = t[u] > s[u] ? cidx[s[u]] : -1;
#pragma unroll
        for (int u = 0; u < MG_UNROLL; ++u) o0[u] = g[u] >= 0 ? opac[g[u]] : 0.f;
        for (int cc = threadIdx.x; cc < OC; cc += MG_TX) {
            float x0[MG_UNROLL];
            float4 q0[MG_UNROLL];
#pragma unroll
            for (int u = 0; u < MG_UNROLL; ++u) {
                x0[u] = 0.f; q0[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (g[u] >= 0) {
                    if (cc >= 3 && cc < 7) q0[u] = *reinterpret_cast<const float4 *>(quats + (size_t)g[u] * 4);
                    else if (cc < 3) x0[u] = means[(size_t)g[u] * 3 + cc];
                    else if (cc < 10) x0[u] = scales[(size_t)g[u] * 3 + (cc - 7)];
                    else if (cc > 10) x0[u] = colors[(size_t)g[u] * color_dim + (cc - 11)];
                }
            }
#pragma unroll
            for (int u = 0; u < MG_UNROLL; ++u) {
                const int64_t vv = v0 + u * vstride;
                if (vv >= nclu) continue;
                const float out = merge_element(cc, s[u], t[u], o0[u], x0[u], q0[u], cidx, means, quats, scales, opac, colors,
                                                color_dim, by_opacity);
                if (cc < 3) { if (m_means) m_means[vv * 3 + cc] = out; }
                else if (cc < 7) { if (m_quats) m_quats[vv * 4 + (cc - 3)] = out; }
                else if (cc < 10) { if (m_scales) m_scales[vv * 3 + (cc - 7)] = out; }
                else if (cc == 10) { if (m_opac) m_opac[vv] = out; }
                else { if (m_colors) m_colors[vv * (int64_t)color_dim + (cc - 11)] = out; }
                if (packed) packed[vv * OC + cc] = (OutT)out;
            }
        }
    }
}

}  // namespace

RAHT_API int raht_merge_segments(const int32_t *cluster_indices, const int32_t *cluster_offsets, int64_t num_clusters,
                                 const float *means, const float *quats, const float *scales, const float *opacities,
                                 const float *colors, int color_dim, int weight_by_opacity, float *m_means,
                                 float *m_quats, float *m_scales, float *m_opacities, float *m_colors, void *packed,
                                 int packed_dtype, void *stream)
{
    RAHT_REQUIRE(num_clusters >= 0 && color_dim >= 0, "raht_merge_segments: bad sizes");
    if (num_clusters == 0) return RAHT_OK;
    RAHT_REQUIRE(cluster_indices && cluster_offsets && means && quats && scales && opacities && (colors || color_dim == 0),
                 "raht_merge_segments: null input pointer");
    RAHT_REQUIRE(((uintptr_t)quats & 15) == 0, "raht_merge_segments: quats must be 16-byte aligned");
    RAHT_REQUIRE(packed_dtype == RAHT_F32 || packed_dtype == RAHT_F64, "raht_merge_segments: bad packed_dtype");
    cudaStream_t st = (cudaStream_t)stream;
    int64_t total = num_clusters * (11 + color_dim);
    int grid = grid_for((num_clusters + MG_UNROLL - 1) / MG_UNROLL, MG_TY, 6);
    const dim3 block(MG_TX, MG_TY);
    raht_count_launches(1);
    if (packed && packed_dtype == RAHT_F64)
        merge_kernel<double><<<grid, block, 0, st>>>(cluster_indices, cluster_offsets, num_clusters, means, quats, scales,
                                                   opacities, colors, color_dim, weight_by_opacity, m_means, m_quats,
                                                   m_scales, m_opacities, m_colors, (double *)packed);
    else
        merge_kernel<float><<<grid, block, 0, st>>>(cluster_indices, cluster_offsets, num_clusters, means, quats, scales,
                                                  opacities, colors, color_dim, weight_by_opacity, m_means, m_quats,
                                                  m_scales, m_opacities, m_colors, (float *)packed);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}
