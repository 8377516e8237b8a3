// scene.cu -- scene-level entry points: the whole encode (voxelise -> merge -> RAHT_param -> RAHT -> quantise ->
// reorder -> RLGR) and the whole decode (RAHT_param -> RLGR^-1 -> dequantise -> inverse reorder -> iRAHT) of
// python/test_voxelize_3dgs.py:198-265 + python/encode_3dgs.py:144-276, issued by THREE host calls instead of ~25.
//
// The kernels and their order are exactly those of the per-operator entry points (this file only calls them); what
// changes is the host side: one call launches a whole leg of the pipeline back to back from C, out of one caller-owned
// workspace arena, where the Python operators paid an allocator round trip and a ctypes transition per kernel (round 1:
// 70 launches and ~0.4 ms of launch gaps per 1 M-Gaussian step).  The only host synchronisation left in an encode is the
// one the data forces: the voxel count must reach the host before the second leg can be sized.
//
//   raht_scene_front    Gaussian centres -> Morton keys -> stable radix sort -> voxel boundaries; nvox to pinned memory
//   raht_scene_encode   merge -> tree (levels, order_RAGFT, plan) -> forward transform -> quantise + reorder -> RLGR
//   raht_scene_decode   tree from the voxel keys -> RLGR decode -> dequantise + un-reorder -> inverse transform -> coordinates
//
// Stage events (optional): `events[k]` are cudaEvent_t handles recorded behind the k-th stage of the call, so that a
// caller can still time the stages (bench.py) although they no longer return to the host in between.
#include "common.cuh"
#include <new>
#include <stdlib.h>

namespace {

struct SceneArena {
    char *base;
    size_t off, cap;
    bool dry;                                     // size computation only
    void *take(size_t bytes)
    {
        off = align_up(off, 256);
        void *p = dry ? nullptr : (void *)(base + off);
        off += bytes;
        return p;
    }
};

inline void record(void **events, int k, cudaStream_t st)
{
    if (events && events[k]) cudaEventRecord((cudaEvent_t)events[k], st);
}

struct StepsArg {
    double v[64];
};
__global__ void steps_fill_kernel(StepsArg a, int D, double *out)
{
    if ((int)threadIdx.x < D) out[threadIdx.x] = a.v[threadIdx.x];
}

int upload_steps(const double *steps_host, int D, double *dst, cudaStream_t st)
{
    if (D <= 64) {
        StepsArg a;
        for (int c = 0; c < 64; ++c) a.v[c] = c < D ? steps_host[c] : 1.0;
        raht_count_launches(1);
        steps_fill_kernel<<<1, 64, 0, st>>>(a, D, dst);
        RAHT_LAUNCH_CHECK();
    } else {
        RAHT_CUDA(cudaMemcpyAsync(dst, steps_host, (size_t)D * 8, cudaMemcpyHostToDevice, st));
    }
    return RAHT_OK;
}

// ---- front -----------------------------------------------------------------------------------------------------
struct FrontBufs {
    float *bounds6;
    uint64_t *keys, *keys_sorted;
    int64_t *voxel_indices;
    void *sort_ws, *seg_ws;
    size_t sort_ws_bytes, seg_ws_bytes;
};
size_t front_layout(SceneArena &A, int64_t n, int J, FrontBufs &B)
{
    B.bounds6 = (float *)A.take(64);
    B.keys = (uint64_t *)A.take((size_t)n * 8);
    B.keys_sorted = (uint64_t *)A.take((size_t)n * 8);
    B.voxel_indices = (int64_t *)A.take((size_t)n * 8);
    B.sort_ws_bytes = raht_radix_sort_workspace_bytes(n, 0, 3 * J);
    B.sort_ws = A.take(B.sort_ws_bytes);
    B.seg_ws_bytes = raht_voxel_segments_workspace_bytes(n);
    B.seg_ws = A.take(B.seg_ws_bytes);
    return align_up(A.off, 256);
}

// ---- encode / decode ---------------------------------------------------------------------------------------------
// Two workspaces: the TREE (level, order_RAGFT, its inverse, w and the transform plan -- what a caller may want to keep
// next to the coded scene) and the SCRATCH (coefficients, symbols, coder workspace -- dead when the call's work is done).
// layout[] handed back to the caller: byte offsets into the tree workspace (LAY_LEVEL .. LAY_PLAN_BYTES) and into the
// scratch (the rest), so that it can look at the intermediate results without copies.
enum { LAY_LEVEL = 0, LAY_ORDER, LAY_POSITION, LAY_W, LAY_PLAN, LAY_PLAN_BYTES, LAY_PACKED, LAY_T, LAY_Q, LAY_Q_LD, LAY_STEPS, LAY_COUNT };

struct CodecBufs {
    uint8_t *level;
    int64_t *order;
    int32_t *position;
    double *w;
    void *plan_ws;
    size_t plan_ws_bytes;
    void *order_ws;
    size_t order_ws_bytes;
    void *packed;
    void *T;
    int32_t *q;
    int64_t ld;
    double *steps;
    void *rlgr_ws;
    size_t rlgr_ws_bytes;
};
size_t tree_layout(SceneArena &A, int64_t nvox, int tile_rows, CodecBufs &B, int64_t *layout)
{
    size_t o;
    o = align_up(A.off, 256); B.level = (uint8_t *)A.take((size_t)nvox);
    if (layout) layout[LAY_LEVEL] = (int64_t)o;
    o = align_up(A.off, 256); B.order = (int64_t *)A.take((size_t)nvox * 8);
    if (layout) layout[LAY_ORDER] = (int64_t)o;
    o = align_up(A.off, 256); B.position = (int32_t *)A.take((size_t)nvox * 4);
    if (layout) layout[LAY_POSITION] = (int64_t)o;
    o = align_up(A.off, 256); B.w = (double *)A.take((size_t)nvox * 8);
    if (layout) layout[LAY_W] = (int64_t)o;
    B.plan_ws_bytes = raht_plan_workspace_bytes(nvox, tile_rows);
    o = align_up(A.off, 256); B.plan_ws = A.take(B.plan_ws_bytes);
    if (layout) { layout[LAY_PLAN] = (int64_t)o; layout[LAY_PLAN_BYTES] = (int64_t)B.plan_ws_bytes; }
    return align_up(A.off, 256);
}
size_t scratch_layout(SceneArena &A, int64_t nvox, int D, int t_dtype, bool encoder, bool own_packed, CodecBufs &B, int64_t *layout)
{
    const size_t esz = t_dtype == RAHT_F64 ? 8 : 4;
    B.ld = (nvox + 31) / 32 * 32;                                    // rows of q start on 128-byte lines
    size_t o;
    o = align_up(A.off, 256); B.steps = (double *)A.take((size_t)D * 8);
    if (layout) layout[LAY_STEPS] = (int64_t)o;
    B.order_ws_bytes = raht_order_ragft_workspace_bytes(nvox);
    B.order_ws = A.take(B.order_ws_bytes);
    B.packed = nullptr;
    if (layout) layout[LAY_PACKED] = -1;
    if (own_packed) {
        o = align_up(A.off, 256); B.packed = A.take((size_t)nvox * D * 4);
        if (layout) layout[LAY_PACKED] = (int64_t)o;
    }
    o = align_up(A.off, 256); B.T = A.take((size_t)nvox * D * esz);
    if (layout) layout[LAY_T] = (int64_t)o;
    o = align_up(A.off, 256); B.q = (int32_t *)A.take((size_t)D * B.ld * 4);
    if (layout) { layout[LAY_Q] = (int64_t)o; layout[LAY_Q_LD] = B.ld; }
    B.rlgr_ws = nullptr;
    B.rlgr_ws_bytes = 0;
    if (encoder) {
        B.rlgr_ws_bytes = raht_rlgr_encode_workspace_bytes(nvox, D);
        B.rlgr_ws = A.take(B.rlgr_ws_bytes);
    }
    return align_up(A.off, 256);
}

bool ws_ok(const void *p, size_t have, size_t need) { return p && have >= need && ((uintptr_t)p & 255) == 0; }

int build_tree(const uint64_t *keys, int64_t nvox, int tile_rows, CodecBufs &B, int32_t *status, raht_plan **plan, cudaStream_t st)
{
    RAHT_TRY(raht_tree_levels((const int64_t *)keys, nvox, B.level, status, st));
    RAHT_TRY(raht_order_ragft(B.level, nvox, B.order, B.position, B.order_ws, B.order_ws_bytes, st));
    RAHT_TRY(raht_plan_create(B.level, nvox, tile_rows, B.w, B.plan_ws, B.plan_ws_bytes, plan, st));
    return RAHT_OK;
}

}  // namespace

RAHT_API size_t raht_scene_front_workspace_bytes(int64_t n, int J)
{
    if (n < 0 || J < 1 || J > 21) return 0;
    SceneArena A{nullptr, 0, 0, true};
    FrontBufs B;
    return front_layout(A, n, J, B) + 256;
}

RAHT_API int raht_scene_front(const float *means, int64_t n, int J, int use_given, const float *vmin_host, double width_host,
                              float *params4, double *params_d2, uint64_t *unique_keys, uint32_t *sort_idx,
                              int32_t *cluster_offsets, int64_t *nvox, void *front_pinned_host, int32_t *status,
                              void *workspace, size_t workspace_bytes, void **events, void *stream)
{
    RAHT_REQUIRE(n >= 1 && J >= 1 && J <= 21 && means && params4 && params_d2 && unique_keys && sort_idx && cluster_offsets && nvox,
                 "raht_scene_front: bad arguments");
    RAHT_REQUIRE(!use_given || vmin_host, "raht_scene_front: vmin missing");
    SceneArena A{(char *)workspace, 0, workspace_bytes, false};
    FrontBufs B;
    const size_t need = front_layout(A, n, J, B);
    if (!workspace || workspace_bytes < need || ((uintptr_t)workspace & 255) != 0) {
        raht_set_error("raht_scene_front: workspace too small or not 256-byte aligned (%zu < %zu)", workspace_bytes, need);
        return RAHT_E_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (use_given) {
        RAHT_CUDA(cudaMemsetAsync(B.bounds6, 0, 24, st));
        RAHT_TRY(raht_voxel_params(B.bounds6, J, 1, vmin_host, 1, width_host, params4, params_d2, st));
    } else {
        RAHT_TRY(raht_bounds3_f32(means, n, 3, B.bounds6, st));
        RAHT_TRY(raht_voxel_params(B.bounds6, J, 0, nullptr, 0, 0.0, params4, params_d2, st));
    }
    RAHT_TRY(raht_morton_keys_f32(means, n, 3, params4, J, B.keys, nullptr, st));
    record(events, 0, st);                                                                       // morton
    RAHT_TRY(raht_radix_sort_pairs(B.keys, B.keys_sorted, nullptr, sort_idx, n, 0, 3 * J, status, B.sort_ws, B.sort_ws_bytes, st));
    record(events, 1, st);                                                                       // sort
    RAHT_TRY(raht_voxel_segments(B.keys_sorted, n, B.voxel_indices, cluster_offsets, unique_keys, nvox, B.seg_ws, B.seg_ws_bytes, st));
    if (front_pinned_host) {                                         // [0, 8): nvox; [8, 24): width, voxel_size
        RAHT_TRY(raht_publish_host(nvox, front_pinned_host, 8, st));
        RAHT_TRY(raht_publish_host(params_d2, (char *)front_pinned_host + 8, 16, st));
    }
    record(events, 2, st);                                                                       // segments
    return RAHT_OK;
}

RAHT_API size_t raht_scene_tree_workspace_bytes(int64_t nvox, int tile_rows)
{
    if (nvox < 1) return 0;
    SceneArena A{nullptr, 0, 0, true};
    CodecBufs B;
    return tree_layout(A, nvox, tile_rows, B, nullptr) + 256;
}

RAHT_API size_t raht_scene_encode_workspace_bytes(int64_t nvox, int D, int t_dtype, int own_packed)
{
    if (nvox < 1 || D < 1 || (t_dtype != RAHT_F64 && t_dtype != RAHT_F32)) return 0;
    SceneArena A{nullptr, 0, 0, true};
    CodecBufs B;
    return scratch_layout(A, nvox, D, t_dtype, true, own_packed != 0, B, nullptr) + 256;
}

RAHT_API int raht_scene_encode(const uint32_t *sort_idx, const int32_t *cluster_offsets, int64_t nvox, const float *means,
                               const float *quats, const float *scales, const float *opacities, const float *colors,
                               int color_dim, const uint64_t *keys, const double *steps_host, int t_dtype, int tile_rows,
                               void *packed_out, uint8_t *streams, int64_t slot_bytes, int64_t *nbytes, void *sync_index,
                               int32_t *status, void *results_pinned_host, raht_plan **plan, int64_t *layout_host,
                               void *tree_workspace, size_t tree_workspace_bytes, void *workspace, size_t workspace_bytes,
                               void **events, void *stream)
{
    RAHT_REQUIRE(nvox >= 1 && sort_idx && cluster_offsets && means && quats && scales && opacities && colors && keys && steps_host &&
                     streams && nbytes && status && plan,
                 "raht_scene_encode: bad arguments");
    RAHT_REQUIRE(t_dtype == RAHT_F64 || t_dtype == RAHT_F32, "raht_scene_encode: bad dtype");
    const int D = 11 + color_dim;
    SceneArena AT{(char *)tree_workspace, 0, tree_workspace_bytes, false}, A{(char *)workspace, 0, workspace_bytes, false};
    CodecBufs B;
    int64_t layout[LAY_COUNT];
    const size_t need_t = tree_layout(AT, nvox, tile_rows, B, layout);
    const size_t need = scratch_layout(A, nvox, D, t_dtype, true, packed_out == nullptr, B, layout);
    if (!ws_ok(tree_workspace, tree_workspace_bytes, need_t) || !ws_ok(workspace, workspace_bytes, need)) {
        raht_set_error("raht_scene_encode: a workspace is too small or not 256-byte aligned (tree %zu < %zu or scratch %zu < %zu)",
                       tree_workspace_bytes, need_t, workspace_bytes, need);
        return RAHT_E_WORKSPACE;
    }
    if (layout_host) for (int k = 0; k < LAY_COUNT; ++k) layout_host[k] = layout[k];
    cudaStream_t st = (cudaStream_t)stream;
    void *packed = packed_out ? packed_out : B.packed;
    *plan = nullptr;
    RAHT_TRY(upload_steps(steps_host, D, B.steps, st));
    // The tree is built on the device's side stream next to the merge (both need only the front leg's results); the merge
    // grid holds every register of the GPU, so the tree's kernels mostly run in its tail: raht_param 0.129 -> 0.108 ms.
    // RAHT_ENC_SIDE=0: one stream.
    static const int enc_side = getenv("RAHT_ENC_SIDE") ? atoi(getenv("RAHT_ENC_SIDE")) : 1;
    cudaStream_t side = enc_side ? raht_side_fork(st) : nullptr;
    if (side) raht_merge_share_gpu();
    RAHT_TRY(raht_merge_segments((const int32_t *)sort_idx, cluster_offsets, nvox, means, quats, scales, opacities, colors, color_dim, 1,
                                 nullptr, nullptr, nullptr, nullptr, nullptr, packed, RAHT_F32, st));
    record(events, 0, st);                                                                       // merge
    RAHT_TRY(build_tree(keys, nvox, tile_rows, B, status, plan, side ? side : st));
    if (side) RAHT_TRY(raht_side_join(st));
    record(events, 1, st);                                                                       // raht_param
    RAHT_TRY(raht_forward(*plan, packed, RAHT_F32, D, B.T, t_dtype, 0, st));
    record(events, 2, st);                                                                       // raht
    RAHT_TRY(raht_quant_reorder(B.T, t_dtype, nvox, D, B.order, B.steps, B.q, B.ld, status, 0, st));
    record(events, 3, st);                                                                       // quant_reorder
    RAHT_TRY(raht_rlgr_encode(B.q, nvox, B.ld, D, 1, streams, slot_bytes, nbytes, sync_index, status, B.rlgr_ws, B.rlgr_ws_bytes, st));
    if (results_pinned_host) {                                       // [0, 8 D): stream lengths; [8 D, 8 D + 4): status word
        RAHT_TRY(raht_publish_host(nbytes, results_pinned_host, (size_t)D * 8, st));
        RAHT_TRY(raht_publish_host(status, (char *)results_pinned_host + (size_t)D * 8, 4, st));
    }
    record(events, 4, st);                                                                       // rlgr_encode
    return RAHT_OK;
}

RAHT_API size_t raht_scene_decode_workspace_bytes(int64_t nvox, int D, int t_dtype)
{
    if (nvox < 1 || D < 1 || (t_dtype != RAHT_F64 && t_dtype != RAHT_F32)) return 0;
    SceneArena A{nullptr, 0, 0, true};
    CodecBufs B;
    return scratch_layout(A, nvox, D, t_dtype, false, false, B, nullptr) + 256;
}

RAHT_API int raht_scene_decode(const uint64_t *keys, int64_t nvox, int J, const uint8_t *streams, int64_t slot_bytes,
                               const int64_t *nbytes, const void *sync_index, int D, const double *steps_host, int t_dtype,
                               int tile_rows, void *attributes, float *coords, int32_t *status, void *status_pinned_host,
                               raht_plan **plan, int64_t *layout_host, void *tree_workspace, size_t tree_workspace_bytes,
                               void *workspace, size_t workspace_bytes, void **events, void *stream)
{
    RAHT_REQUIRE(nvox >= 1 && keys && streams && steps_host && attributes && status && plan && D >= 1, "raht_scene_decode: bad arguments");
    RAHT_REQUIRE(t_dtype == RAHT_F64 || t_dtype == RAHT_F32, "raht_scene_decode: bad dtype");
    SceneArena AT{(char *)tree_workspace, 0, tree_workspace_bytes, false}, A{(char *)workspace, 0, workspace_bytes, false};
    CodecBufs B;
    int64_t layout[LAY_COUNT];
    const size_t need_t = tree_layout(AT, nvox, tile_rows, B, layout);
    const size_t need = scratch_layout(A, nvox, D, t_dtype, false, false, B, layout);
    if (!ws_ok(tree_workspace, tree_workspace_bytes, need_t) || !ws_ok(workspace, workspace_bytes, need)) {
        raht_set_error("raht_scene_decode: a workspace is too small or not 256-byte aligned (tree %zu < %zu or scratch %zu < %zu)",
                       tree_workspace_bytes, need_t, workspace_bytes, need);
        return RAHT_E_WORKSPACE;
    }
    if (layout_host) for (int k = 0; k < LAY_COUNT; ++k) layout_host[k] = layout[k];
    cudaStream_t st = (cudaStream_t)stream;
    *plan = nullptr;
    RAHT_TRY(upload_steps(steps_host, D, B.steps, st));
    // The tree (levels, order, plan: latency-bound, needs only the keys) is built on the device's side stream while the
    // entropy decoder runs on the caller's; the stage events then read ~0 for raht_param_dec and both legs for rlgr_decode
    // (0.144 + 0.356 ms one after the other, 0.48 ms side by side: the decoder's one wave of CTAs leaves the plan little
    // room until it drains).  RAHT_DEC_SIDE=0: one stream.
    static const int dec_side = getenv("RAHT_DEC_SIDE") ? atoi(getenv("RAHT_DEC_SIDE")) : 1;
    cudaStream_t side = dec_side ? raht_side_fork(st) : nullptr;
    RAHT_TRY(build_tree(keys, nvox, tile_rows, B, status, plan, side ? side : st));
    record(events, 0, st);                                                                       // raht_param_dec
    if (side) raht_rlgr_decode_share_gpu();
    RAHT_TRY(raht_rlgr_decode(streams, slot_bytes, nbytes, sync_index, nvox, B.ld, D, 1, B.q, st));
    if (side) RAHT_TRY(raht_side_join(st));
    record(events, 1, st);                                                                       // rlgr_decode
    RAHT_TRY(raht_dequant_unreorder(B.q, B.ld, nvox, D, B.order, B.steps, B.T, t_dtype, st));
    record(events, 2, st);                                                                       // dequant_unreorder
    RAHT_TRY(raht_inverse(*plan, B.T, t_dtype, D, attributes, 0, st));
    record(events, 3, st);                                                                       // iraht
    if (coords) RAHT_TRY(raht_morton_decode_f32(keys, nvox, J, coords, st));
    if (status_pinned_host) RAHT_TRY(raht_publish_host(status, status_pinned_host, 4, st));
    record(events, 4, st);                                                                       // geometry
    return RAHT_OK;
}
