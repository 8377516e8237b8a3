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
#include <stdlib.h>

namespace {

constexpr int MG_UNROLL = 4;      // voxels per group: their row gathers are in flight together (two groups buffered)
constexpr int MG_WARPS = 8;

// What a lane needs to know about its output channel c = 32 k + lane, fixed for the whole kernel: where the channel
// lives in the reference's five arrays (base pointer + row stride) and how its sum is finished.
enum { CH_MEAN = 0, CH_QUAT = 1, CH_OPAC = 2, CH_COLOR = 3, CH_NONE = 4 };   // CH_MEAN also covers the scales (same rule)
struct LaneChannel {
    const float *base;
    int stride;
    int kind;
};
__device__ __forceinline__ LaneChannel lane_channel(int c, int OC, const float *means, const float *quats, const float *scales,
                                                    const float *opac, const float *colors, int color_dim)
{
    LaneChannel L;
    L.base = colors + (c - 11); L.stride = color_dim; L.kind = CH_COLOR;
    if (c == 10) { L.base = opac; L.stride = 1; L.kind = CH_OPAC; }
    if (c < 10) { L.base = scales + (c - 7); L.stride = 3; L.kind = CH_MEAN; }
    if (c < 7) { L.base = quats + (c - 3); L.stride = 4; L.kind = CH_QUAT; }
    if (c < 3) { L.base = means + c; L.stride = 3; L.kind = CH_MEAN; }
    if (c >= OC) { L.base = means; L.stride = 0; L.kind = CH_NONE; }
    return L;
}

// A warp takes blocks of 32 consecutive voxels.  Lane l reads the segment bounds and the first member index of voxel l
// of the block (two coalesced round trips per 32 voxels; they used to be two dependent, warp-uniform round trips per
// group of four), then the voxels are processed MG_UNROLL at a time, lane = output channel (slot k holds channel
// 32 k + lane, NS slots): the member's colour row and the packed output row are contiguous across the lanes.  The row
// gathers of the next group are issued before the current group is finished (two register buffers), so the only
// dependent round trips left per block are the two index loads.  Every channel sums its members in segment order with
// separately rounded multiply and add, starting from 0 as the reference does (merge_cluster.cu:36-72,97-106), so the
// result is a deterministic function of the inputs.
// SEP: also write the five separate output arrays of the reference interface (the codec only takes the packed rows).
template <int NS>
__device__ __forceinline__ void merge_load_group(const LaneChannel (&ch)[NS], int gl, int j0, float (&x)[MG_UNROLL][NS])
{
#pragma unroll
    for (int u = 0; u < MG_UNROLL; ++u) {
        const int g = __shfl_sync(~0u, gl, j0 + u);
#pragma unroll
        for (int k = 0; k < NS; ++k)
            x[u][k] = (g >= 0 && ch[k].kind != CH_NONE) ? ch[k].base[(int64_t)g * ch[k].stride] : 0.f;
    }
}

template <typename OutT, int NS, bool SEP>
__device__ __forceinline__ void merge_finish_group(const LaneChannel (&ch)[NS], const int32_t *__restrict__ cidx, int sl, int tl, int j0,
                                                   int64_t vb, int64_t nclu, const float (&x)[MG_UNROLL][NS], int OC, int color_dim,
                                                   int by_opacity, float *__restrict__ m_means, float *__restrict__ m_quats,
                                                   float *__restrict__ m_scales, float *__restrict__ m_opac,
                                                   float *__restrict__ m_colors, OutT *__restrict__ packed)
{
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int u = 0; u < MG_UNROLL; ++u) {
        const int64_t v = vb + j0 + u;
        const int su = __shfl_sync(~0u, sl, j0 + u), tu = __shfl_sync(~0u, tl, j0 + u);
        if (v >= nclu) continue;                                         // warp-uniform
        float acc[NS], wsum = 0.f;
#pragma unroll
        for (int k = 0; k < NS; ++k) acc[k] = 0.f;
        float xm[NS];
#pragma unroll
        for (int k = 0; k < NS; ++k) xm[k] = x[u][k];
        for (int i = su; i < tu; ++i) {                                  // warp-uniform trip count (1 for most voxels)
            if (i > su) {
                const int64_t gi = cidx[i];
#pragma unroll
                for (int k = 0; k < NS; ++k) xm[k] = ch[k].kind != CH_NONE ? ch[k].base[gi * ch[k].stride] : 0.f;
            }
            const float o = __shfl_sync(~0u, xm[0], 10);                 // the member's opacity
            const float w = by_opacity ? o : 1.0f;                       // :42
            wsum = __fadd_rn(wsum, w);
#pragma unroll
            for (int k = 0; k < NS; ++k) {
                const float wk = ch[k].kind == CH_OPAC ? 1.0f : w;       // the opacity channel sums the raw opacities (:60)
                acc[k] = __fadd_rn(acc[k], __fmul_rn(xm[k], wk));
            }
        }
        // quaternion norm from lanes 3..6 (:74-76)
        const float q0 = __shfl_sync(~0u, acc[0], 3), q1 = __shfl_sync(~0u, acc[0], 4);
        const float q2 = __shfl_sync(~0u, acc[0], 5), q3 = __shfl_sync(~0u, acc[0], 6);
        const float qn = __fsqrt_rn(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(q0, q0), __fmul_rn(q1, q1)), __fmul_rn(q2, q2)),
                                              __fmul_rn(q3, q3)));
        const bool empty = tu <= su;                                     // :27-29
#pragma unroll
        for (int k = 0; k < NS; ++k) {
            const int kind = ch[k].kind;
            if (kind == CH_NONE) continue;
            const int c = 32 * k + lane;
            // one division per channel: by the quaternion norm (:77-87), or by the weight sum -- means and scales take
            // 1 for an all-zero weight (:65-72), colours give 0 then (:108-109)
            const float den = kind == CH_QUAT ? qn : ((kind == CH_MEAN && wsum == 0.0f) ? 1.0f : wsum);
            float out = __fdiv_rn(acc[k], den);
            if (kind == CH_QUAT && !(qn > 0.0f)) out = c == 6 ? 1.0f : 0.0f;
            if (kind == CH_COLOR && !(wsum > 0.0f)) out = 0.0f;
            if (kind == CH_OPAC) out = fminf(acc[k], 1.0f);                                          // :94
            if (empty) out = 0.0f;
            if constexpr (SEP) {
                if (c < 3) { if (m_means) m_means[v * 3 + c] = out; }
                else if (c < 7) { if (m_quats) m_quats[v * 4 + (c - 3)] = out; }
                else if (c < 10) { if (m_scales) m_scales[v * 3 + (c - 7)] = out; }
                else if (c == 10) { if (m_opac) m_opac[v] = out; }
                else { if (m_colors) m_colors[v * (int64_t)color_dim + (c - 11)] = out; }
            }
            if (packed) packed[v * OC + c] = (OutT)out;
        }
    }
}

// The same for a group whose MG_UNROLL voxels all exist and have exactly ONE member (every group of a scene whose voxels
// are finer than its Gaussians: 99.98 % of the voxels of the 1 M / 6 M bench scenes).  The arithmetic is the general
// path's with its member loop run once -- 0 + w, 0 + x * w, the same divisions and selects -- so the results are
// bit-identical; what goes away is the variable-trip loop, the per-voxel bounds shuffles, the kind tests of the slots
// above the first (channels >= 32 are colours) and the divergence scaffolding around them: the general path executed
// 170 warp instructions per voxel at 59 channels (profiles/r02_ncu_full_v1.txt: issue-bound at 68 %).
template <typename OutT, int NS, bool SEP>
__device__ __forceinline__ void merge_finish_group_single(const LaneChannel (&ch)[NS], int64_t v0, const float (&x)[MG_UNROLL][NS], int OC,
                                                          int color_dim, int by_opacity, float *__restrict__ m_means,
                                                          float *__restrict__ m_quats, float *__restrict__ m_scales,
                                                          float *__restrict__ m_opac, float *__restrict__ m_colors,
                                                          OutT *__restrict__ packed)
{
    const int lane = threadIdx.x & 31;
    const int kind0 = ch[0].kind;
#pragma unroll
    for (int u = 0; u < MG_UNROLL; ++u) {
        const int64_t v = v0 + u;
        const float o = __shfl_sync(~0u, x[u][0], 10);                   // the member's opacity
        const float w = by_opacity ? o : 1.0f;                           // :42
        const float wsum = __fadd_rn(0.f, w);
        float acc[NS];
        acc[0] = __fadd_rn(0.f, __fmul_rn(x[u][0], kind0 == CH_OPAC ? 1.0f : w));
#pragma unroll
        for (int k = 1; k < NS; ++k) acc[k] = __fadd_rn(0.f, __fmul_rn(x[u][k], w));
        // quaternion norm from lanes 3..6 (:74-76)
        const float q0 = __shfl_sync(~0u, acc[0], 3), q1 = __shfl_sync(~0u, acc[0], 4);
        const float q2 = __shfl_sync(~0u, acc[0], 5), q3 = __shfl_sync(~0u, acc[0], 6);
        const float qn = __fsqrt_rn(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(q0, q0), __fmul_rn(q1, q1)), __fmul_rn(q2, q2)),
                                              __fmul_rn(q3, q3)));
        const bool wpos = wsum > 0.0f;
        OutT *prow = packed ? packed + v * OC : nullptr;
        if (kind0 != CH_NONE) {                                          // slot 0: channels 0 .. 31, every kind
            const float den = kind0 == CH_QUAT ? qn : ((kind0 == CH_MEAN && wsum == 0.0f) ? 1.0f : wsum);
            float out = __fdiv_rn(acc[0], den);
            if (kind0 == CH_QUAT && !(qn > 0.0f)) out = lane == 6 ? 1.0f : 0.0f;
            if (kind0 == CH_COLOR && !wpos) out = 0.0f;
            if (kind0 == CH_OPAC) out = fminf(acc[0], 1.0f);                                         // :94
            if constexpr (SEP) {
                const int c = lane;
                if (c < 3) { if (m_means) m_means[v * 3 + c] = out; }
                else if (c < 7) { if (m_quats) m_quats[v * 4 + (c - 3)] = out; }
                else if (c < 10) { if (m_scales) m_scales[v * 3 + (c - 7)] = out; }
                else if (c == 10) { if (m_opac) m_opac[v] = out; }
                else { if (m_colors) m_colors[v * (int64_t)color_dim + (c - 11)] = out; }
            }
            if (prow) prow[lane] = (OutT)out;
        }
#pragma unroll
        for (int k = 1; k < NS; ++k) {                                   // channels >= 32: colours (or nothing)
            if (ch[k].kind == CH_NONE) continue;
            const int c = 32 * k + lane;
            float out = __fdiv_rn(acc[k], wsum);
            if (!wpos) out = 0.0f;
            if constexpr (SEP) { if (m_colors) m_colors[v * (int64_t)color_dim + (c - 11)] = out; }
            if (prow) prow[c] = (OutT)out;
        }
    }
}

template <typename OutT, int NS, bool SEP>
__global__ void __launch_bounds__(32 * MG_WARPS, NS <= 2 ? 4 : 1)
merge_kernel(const int32_t *__restrict__ cidx, const int32_t *__restrict__ coff, int64_t nclu,
             const float *__restrict__ means, const float *__restrict__ quats, const float *__restrict__ scales,
             const float *__restrict__ opac, const float *__restrict__ colors, int color_dim, int by_opacity,
             float *__restrict__ m_means, float *__restrict__ m_quats, float *__restrict__ m_scales,
             float *__restrict__ m_opac, float *__restrict__ m_colors, OutT *__restrict__ packed)
{
    static_assert(32 % (2 * MG_UNROLL) == 0, "a block of 32 voxels is a whole number of group pairs");
    const int OC = 11 + color_dim;
    const int lane = threadIdx.x & 31;
    LaneChannel ch[NS];
#pragma unroll
    for (int k = 0; k < NS; ++k) ch[k] = lane_channel(32 * k + lane, OC, means, quats, scales, opac, colors, color_dim);
    const int64_t nwarps = (int64_t)gridDim.x * MG_WARPS;
    const int64_t nblk = (nclu + 31) >> 5;
    for (int64_t b = (int64_t)blockIdx.x * MG_WARPS + (threadIdx.x >> 5); b < nblk; b += nwarps) {
        const int64_t vb = b << 5;
        int sl = 0, tl = 0;
        if (vb + lane < nclu) {
            sl = coff[vb + lane];
            tl = coff[vb + lane + 1];
        }
        const int gl = tl > sl ? cidx[sl] : -1;
        const int nv = (int)min((int64_t)32, nclu - vb);
        // voxels of the block that exist and have exactly one member; a group of MG_UNROLL such voxels takes the lean path
        const unsigned single = __ballot_sync(~0u, vb + lane < nclu && tl - sl == 1);
        constexpr unsigned GROUP_MASK = (1u << MG_UNROLL) - 1u;
        float xa[MG_UNROLL][NS], xb[MG_UNROLL][NS];
        merge_load_group<NS>(ch, gl, 0, xa);
        for (int j0 = 0; j0 < nv; j0 += 2 * MG_UNROLL) {                 // warp-uniform
            const bool second = j0 + MG_UNROLL < nv;
            if (second) merge_load_group<NS>(ch, gl, j0 + MG_UNROLL, xb);
            if (((single >> j0) & GROUP_MASK) == GROUP_MASK)
                merge_finish_group_single<OutT, NS, SEP>(ch, vb + j0, xa, OC, color_dim, by_opacity, m_means, m_quats, m_scales, m_opac,
                                                         m_colors, packed);
            else
                merge_finish_group<OutT, NS, SEP>(ch, cidx, sl, tl, j0, vb, nclu, xa, OC, color_dim, by_opacity, m_means, m_quats,
                                                  m_scales, m_opac, m_colors, packed);
            if (second) {
                if (j0 + 2 * MG_UNROLL < nv) merge_load_group<NS>(ch, gl, j0 + 2 * MG_UNROLL, xa);
                if (((single >> (j0 + MG_UNROLL)) & GROUP_MASK) == GROUP_MASK)
                    merge_finish_group_single<OutT, NS, SEP>(ch, vb + j0 + MG_UNROLL, xb, OC, color_dim, by_opacity, m_means, m_quats,
                                                             m_scales, m_opac, m_colors, packed);
                else
                    merge_finish_group<OutT, NS, SEP>(ch, cidx, sl, tl, j0 + MG_UNROLL, vb, nclu, xb, OC, color_dim, by_opacity, m_means,
                                                      m_quats, m_scales, m_opac, m_colors, packed);
            }
        }
        // (requesting the next block's bounds and member indices a block ahead was measured: 73 registers, no gain)
    }
}

template <typename OutT, bool SEP>
void launch_merge_s(int slots, int grid, cudaStream_t st, const int32_t *ci, const int32_t *co, int64_t n, const float *me,
                    const float *qu, const float *sc, const float *op, const float *cl, int cd, int byo, float *m1, float *m2,
                    float *m3, float *m4, float *m5, OutT *pk)
{
    const int T = 32 * MG_WARPS;
    if (slots <= 1) merge_kernel<OutT, 1, SEP><<<grid, T, 0, st>>>(ci, co, n, me, qu, sc, op, cl, cd, byo, m1, m2, m3, m4, m5, pk);
    else if (slots <= 2) merge_kernel<OutT, 2, SEP><<<grid, T, 0, st>>>(ci, co, n, me, qu, sc, op, cl, cd, byo, m1, m2, m3, m4, m5, pk);
    else if (slots <= 4) merge_kernel<OutT, 4, SEP><<<grid, T, 0, st>>>(ci, co, n, me, qu, sc, op, cl, cd, byo, m1, m2, m3, m4, m5, pk);
    else merge_kernel<OutT, 8, SEP><<<grid, T, 0, st>>>(ci, co, n, me, qu, sc, op, cl, cd, byo, m1, m2, m3, m4, m5, pk);
}

template <typename OutT>
void launch_merge(int slots, int grid, cudaStream_t st, const int32_t *ci, const int32_t *co, int64_t n, const float *me,
                  const float *qu, const float *sc, const float *op, const float *cl, int cd, int byo, float *m1, float *m2,
                  float *m3, float *m4, float *m5, OutT *pk)
{
    if (m1 || m2 || m3 || m4 || m5) launch_merge_s<OutT, true>(slots, grid, st, ci, co, n, me, qu, sc, op, cl, cd, byo, m1, m2, m3, m4, m5, pk);
    else launch_merge_s<OutT, false>(slots, grid, st, ci, co, n, me, qu, sc, op, cl, cd, byo, m1, m2, m3, m4, m5, pk);
}

}  // namespace

static thread_local bool g_merge_share = false;
void raht_merge_share_gpu() { g_merge_share = true; }          // (internal, common.cuh) applies to this thread's next raht_merge_segments

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
    RAHT_REQUIRE(color_dim <= 245, "raht_merge_segments: color_dim must be <= 245");
    cudaStream_t st = (cudaStream_t)stream;
    const int slots = (11 + color_dim + 31) / 32;
    // One persistent wave (4 CTAs of 64 registers per SM hold every register of the GPU) is the fastest merge on its own,
    // but nothing else gets onto an SM until it ends.  When the caller builds the tree on the side stream at the same time
    // (raht_merge_share_gpu, scene.cu) the grid is up to 8 waves of short-lived CTAs instead, and the tree's kernels take
    // the slots they free: merge 0.234 -> 0.303 ms, but the tree build (0.105 ms exposed before) disappears under it.
    // RAHT_MERGE_WAVES overrides the 8.
    static const int env_waves = getenv("RAHT_MERGE_WAVES") ? atoi(getenv("RAHT_MERGE_WAVES")) : 0;
    const int waves = !g_merge_share ? 1 : (env_waves > 0 ? env_waves : 8);
    g_merge_share = false;
    const int grid = grid_for((num_clusters + 31) / 32, MG_WARPS, 4 * waves);          // a warp per block of 32 voxels
    raht_count_launches(1);
    if (packed && packed_dtype == RAHT_F64)
        launch_merge<double>(slots, grid, st, cluster_indices, cluster_offsets, num_clusters, means, quats, scales, opacities,
                             colors, color_dim, weight_by_opacity, m_means, m_quats, m_scales, m_opacities, m_colors,
                             (double *)packed);
    else
        launch_merge<float>(slots, grid, st, cluster_indices, cluster_offsets, num_clusters, means, quats, scales, opacities,
                            colors, color_dim, weight_by_opacity, m_means, m_quats, m_scales, m_opacities, m_colors,
                            (float *)packed);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}
