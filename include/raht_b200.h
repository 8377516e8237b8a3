/*
 * raht_b200.h -- C ABI of libraht_b200.so: the B200 (sm_100a) implementation of the RAHT-3DGS
 * encode/decode hot path of Daaohame/RAHT-3DGS-codec (reference commit ef45e2e).
 *
 * This is the drop-in boundary.  Everything the reference's Python operator layer needs for the
 * path  voxelise -> merge -> RAHT_param -> RAHT -> quantise -> reorder -> RLGR  (and the inverse)
 * is reachable through the entry points below: plain pointers and sizes, no torch types.
 * The reference has no C FFI of its own for this path (its natives are a torch-extension and a
 * pybind11 module); each entry point cites the reference interface it replaces (file:line relative
 * to the reference repository) so a maintainer can bind it with ctypes.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller unless its name ends in _host;
 *   - every call is stream-ordered on `stream` (a cudaStream_t passed as void*); only the calls
 *     documented as "synchronises" wait for the stream (they must learn a size on the host);
 *   - scratch memory is caller-provided: ask raht_*_workspace_bytes() first (the transform plan is
 *     the one library-owned object: raht_plan_create / raht_plan_destroy);
 *   - return value: 0 on success, a negative RAHT_E_* code otherwise; raht_last_error() returns
 *     a thread-local message for the last failing call;
 *   - inputs are never modified; outputs must not alias inputs.
 */
#ifndef RAHT_B200_H
#define RAHT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RAHT_B200_ABI_VERSION 18

enum {
    RAHT_OK = 0,
    RAHT_E_INVALID = -1,   /* bad argument (null pointer, size, depth, alignment)          */
    RAHT_E_WORKSPACE = -2, /* workspace too small                                          */
    RAHT_E_CUDA = -3,      /* a CUDA runtime call or launch failed                         */
    RAHT_E_NODEVICE = -4,  /* no sm_100 device / wrong architecture                        */
    RAHT_E_INPUT = -5      /* device-side validation failed (see status bits)              */
};

/* dtype tags for attribute matrices */
enum { RAHT_F32 = 0, RAHT_F64 = 1 };

/* device-side status bits (written with atomicOr into a caller-zeroed int32 word) */
enum {
    RAHT_ST_OUT_OF_BOUNDS = 1,     /* RAHT_param.py:26-27 ValueError("RAHT_param:OutOfBounds")  */
    RAHT_ST_NOT_SORTED_UNIQUE = 2, /* rows not Morton-sorted/unique (undefined in the reference) */
    RAHT_ST_INT32_OVERFLOW = 4,    /* |q| >= 2^31 on the int32 cast, encode_3dgs.py:215         */
    RAHT_ST_STREAM_OVERFLOW = 8,   /* RLGR stream did not fit its slot                          */
    RAHT_ST_SORT_STALL = 16        /* radix-sort look-back gave up (internal error)             */
};

int raht_abi_version(void);
/* kernels launched by this process through the library so far (bench.py reports the per-step delta) */
long long raht_launch_count(void);
const char *raht_last_error(void);
/* number of SMs / compute capability of the current device (major*10+minor); <0 on error */
int raht_device_info(int *sm_count, int *cc);
/* `bytes` (<= 1 MiB) of device memory stored into PINNED host memory by a kernel on `stream` (no copy engine:
 * a small result does not queue behind another stream's bulk device-to-host copy); valid on the host after
 * the stream (or an event recorded behind the call) has been synchronised.                                */
int raht_publish_host(const void *src, void *pinned_host, size_t bytes, void *stream);

/* ------------------------------------------------------------------------------------------------
 * (a1,a2) Voxelisation front end.   Replaces python/voxelize_pc.py:25-59 (get_morton_code) and
 * :62-172 (voxelize_pc_batched): min / max, fp32 quantisation, Morton keys, torch.sort (:101),
 * boundary detection (:114-120).
 * ------------------------------------------------------------------------------------------------ */

/* bounds6[0..2] = min over rows of V[:,0..2], bounds6[3..5] = max.  V is (n, ld) fp32, ld >= 3.
 * voxelize_pc.py:87-95 (vmin = V.min(0); width = (V - vmin).max()).  bounds6 needs no init.     */
int raht_bounds3_f32(const float *V, int64_t n, int64_t ld, float *bounds6, void *stream);

/* Device-side completion of voxelize_pc.py:87-97 from bounds6: if use_given_* != 0 the caller's
 * vmin_host[3] / width_host are used instead of the data's.  Writes params4 = {vmin x,y,z,
 * (float)voxel_size} and params_d2 = {width, voxel_size} (doubles).                               */
int raht_voxel_params(const float *bounds6, int J, int use_given_vmin, const float *vmin_host,
                      int use_given_width, double width_host, float *params4, double *params_d2,
                      void *stream);

/* keys[i] = Morton( clamp(floor((V[i]-vmin)/voxel_size), 0, 2^J-1) ), fp32 arithmetic exactly as
 * torch does for a float32 tensor (voxelize_pc.py:90,98,100).  params4 from raht_voxel_params.
 * vint (n,3) int64 optional: the clamped integer coordinates (V0_integer).                        */
int raht_morton_keys_f32(const float *V, int64_t n, int64_t ld, const float *params4, int J,
                         uint64_t *keys, int64_t *vint, void *stream);

/* get_morton_code(V int64 (n,3), J)  voxelize_pc.py:25-59 */
int raht_morton_keys_i64(const int64_t *V, int64_t n, int J, int64_t *keys, void *stream);

/* RAHT_param.py:21-36 / :201-212: Vint = floor((V - minV)/Q), Q = width/2^depth, fp64; sets
 * RAHT_ST_OUT_OF_BOUNDS in *status when an index leaves [0, 2^depth-1].                           */
int raht_morton_keys_f64(const double *V, int64_t n, const double *minV_host, double width, int depth,
                         int64_t *keys, int32_t *status, void *stream);

/* inverse of the Morton interleave: coords (n,3) as fp32 (PCvox, voxelize_pc.py:158-160) */
int raht_morton_decode_f32(const uint64_t *keys, int64_t n, int J, float *coords, void *stream);

/* Onesweep LSD radix sort of (key, payload) pairs, stable, bits [begin_bit, end_bit) of the key.
 * vals_in == NULL sorts the identity permutation (payload = original index).  Replaces
 * torch.sort at voxelize_pc.py:101 (stable: ties keep original order).  status: caller-zeroed
 * int32 word (RAHT_ST_SORT_STALL).                                                                 */
size_t raht_radix_sort_workspace_bytes(int64_t n, int begin_bit, int end_bit);
int raht_radix_sort_pairs(const uint64_t *keys_in, uint64_t *keys_out, const uint32_t *vals_in,
                          uint32_t *vals_out, int64_t n, int begin_bit, int end_bit, int32_t *status,
                          void *workspace, size_t workspace_bytes, void *stream);

/* voxel_indices[v] = first sorted position of voxel v (voxelize_pc.py:114-118); *nvox (device).
 * Also writes cluster_offsets int32 (nvox+1 entries, test_voxelize_3dgs.py:225-233) when non-NULL
 * and the unique keys (nvox) when non-NULL.                                                        */
size_t raht_voxel_segments_workspace_bytes(int64_t n);
int raht_voxel_segments(const uint64_t *sorted_keys, int64_t n, int64_t *voxel_indices,
                        int32_t *cluster_offsets, uint64_t *unique_keys, int64_t *nvox,
                        void *workspace, size_t workspace_bytes, void *stream);

/* Row gather + voxelisation residual: PCsorted = PC[idx], DeltaV = V0 - voxel_size*floor(V0/voxel_size)
 * (voxelize_pc.py:103-111).  Either output may be NULL.  PC is (n, ld) fp32; PCsorted (n, ld);
 * DeltaV (n, ld_delta) columns 0..2.                                                              */
int raht_gather_rows_f32(const float *PC, int64_t n, int64_t ld, const uint32_t *idx,
                         const float *params4, float *PCsorted, float *DeltaV, int64_t ld_delta,
                         void *stream);

/* Per-voxel attribute mean (voxelize_pc.py:122-153): Cvox[v] = sum of C[idx[s..e)] in sorted order
 * / count; PC is (n, ld) fp32 with the d attributes in columns 3..3+d.  Cvox (nvox, ld_out) written
 * at columns col_out..col_out+d.  DeltaC (n, ld_delta) at columns 3..3+d optional.                 */
int raht_voxel_attr_mean_f32(const float *PC, int64_t n, int64_t ld, int d, const uint32_t *idx,
                             const int64_t *voxel_indices, int64_t nvox, float *Cvox, int64_t ld_out,
                             int col_out, float *DeltaC, int64_t ld_delta, void *stream);

/* ------------------------------------------------------------------------------------------------
 * (a4) Segmented Gaussian merge.  Replaces cuda/merge_cluster.cu:2-111 and its host wrapper
 * cuda/merge_cluster_wrapper.cu:11-116 (merge_clusters_cuda).  fp32, IEEE (no fast-math).
 * Any of the five outputs may be NULL.  packed (optional): row-major (num_clusters, 11+color_dim)
 * matrix [means3|quats4|scales3|opacity1|colors] in packed_dtype, the layout the transform eats.
 * Empty segments produce zero rows (the reference leaves torch::zeros, merge_cluster.cu:27-29).
 * ------------------------------------------------------------------------------------------------ */
int raht_merge_segments(const int32_t *cluster_indices, const int32_t *cluster_offsets,
                        int64_t num_clusters, const float *means, const float *quats,
                        const float *scales, const float *opacities, const float *colors,
                        int color_dim, int weight_by_opacity, float *m_means, float *m_quats,
                        float *m_scales, float *m_opacities, float *m_colors, void *packed,
                        int packed_dtype, void *stream);

/* ------------------------------------------------------------------------------------------------
 * (a5,a6) Tree construction.  Replaces python/RAHT_param.py:6-83 (RAHT_param) and :191-279
 * (RAHT_param_reorder_fast).
 *
 * Compact device tree: for every row i of the Morton-sorted unique codes, level[i] = bit level at
 * which row i is merged as a right sibling = index of the highest bit in which MC[i-1] and MC[i]
 * differ; level[0] = 255 (root).  Everything the reference returns (List/Flags/weights per level,
 * order_RAGFT, w) is a function of level[] alone.
 * ------------------------------------------------------------------------------------------------ */
int raht_tree_levels(const int64_t *MC, int64_t n, uint8_t *level, int32_t *status, void *stream);

/* List[l] / Flags[l] / weights[l] of RAHT_param.py:226-277 for one level; *count (device int64). */
size_t raht_tree_level_workspace_bytes(int64_t n);
int raht_tree_level(const uint8_t *level, int64_t n, int l, int64_t *list, uint8_t *flags,
                    int64_t *weights, int64_t *count, void *workspace, size_t workspace_bytes,
                    void *stream);

/* max over rows>=1 of level[] (-1 when n<=1) -> *max_level (device int32): len(List) = max+1.     */
int raht_tree_max_level(const uint8_t *level, int64_t n, int32_t *max_level, void *stream);

/* order_RAGFT (RAHT_param.py:219-274): the root (the row of level 255: row 0), then rows by descending octree level (level/3),
 * ascending row index inside a group.  order (n x int64, optional); position (n x int32, optional)
 * is the inverse permutation (the argsort of encode_3dgs.py:267).                                  */
size_t raht_order_ragft_workspace_bytes(int64_t n);
int raht_order_ragft(const uint8_t *level, int64_t n, int64_t *order, int32_t *position,
                     void *workspace, size_t workspace_bytes, void *stream);

/* Transform plan: the butterflies of the tree scheduled for the fused shared-memory kernels.
 * Stage 0 cuts the rows into tiles of tile_rows Morton-contiguous rows; every butterfly whose
 * row span lies inside one tile runs there, the others ("deferred", at most 129 per tile) go, 8
 * tiles' worth at a time, into a tile of the next stage, and so on until one tile remains.  The
 * hierarchy is static: stage count, grids and buffer sizes depend on (n, tile_rows) only, so the
 * call neither synchronises nor allocates on the device -- the plan lives in the caller's
 * workspace (256-byte aligned, raht_plan_workspace_bytes), which must outlive it.
 * tile_rows in [32, 1024], even.  w (n doubles, optional): the second output of RAHT2_optimized
 * (RAHT.py:286,325-328).                                                                         */
typedef struct raht_plan raht_plan;
size_t raht_plan_workspace_bytes(int64_t n, int tile_rows);
int raht_plan_create(const uint8_t *level, int64_t n, int tile_rows, double *w, void *workspace,
                     size_t workspace_bytes, raht_plan **plan, void *stream);
void raht_plan_destroy(raht_plan *plan);   /* frees the host handle only */
/* info[0]=n, [1]=tile_rows, [2]=stages, [3]=workspace bytes used, [4+k]=tiles in stage k      */
int raht_plan_info(const raht_plan *plan, int64_t *info_host, int cap);

/* ------------------------------------------------------------------------------------------------
 * (a7,a8) Transform.  Replaces python/RAHT.py:253-336 (RAHT2_optimized) and python/iRAHT.py:40-114
 * (inverse_RAHT_optimized).  With t_dtype = RAHT_F64: IEEE fp64, no FMA contraction, bit-exact with
 * the reference on CPU.  t_dtype = RAHT_F32 is the build-specific fast mode.
 * C is (n, D) row-major of c_dtype; T (n, D) row-major of t_dtype.  flags: bit 0 = do not use TMA.
 * ------------------------------------------------------------------------------------------------ */
int raht_forward(const raht_plan *plan, const void *C, int c_dtype, int D, void *T, int t_dtype,
                 int flags, void *stream);
int raht_inverse(const raht_plan *plan, const void *T, int t_dtype, int D, void *C, int flags,
                 void *stream);

/* Top of the tree across Morton-prefix shards (no counterpart in the reference, which is
 * single-GPU: this is the part of RAHT.py:253-336 / iRAHT.py:40-114 that joins the shards).  The
 * Morton-sorted voxels of a scene are cut into P contiguous ranges; each range is transformed on its
 * own (raht_forward on its rows), which leaves its low-pass in its first row.  rows (P, D) holds those
 * first rows; first_keys[i] = first Morton code of range i, prev_last_keys[i] = last code before range
 * i (i >= 1), offsets[P+1] = global row index of each range's first row (offsets[P] = rows of the
 * scene).  Runs the P-1 butterflies that join the ranges in place (inverse != 0: their inverses, in
 * reverse order) with the arithmetic of raht_forward / raht_inverse: together with the per-range
 * transforms the coefficients are bit-identical to the transform of the whole scene.
 * level_out (P bytes, optional): level of each range's first row in the whole scene (255 for range
 * 0) -- what raht_order_ragft needs for those rows.  One CTA; P <= 1024.                          */
int raht_shard_top(const int64_t *first_keys, const int64_t *prev_last_keys, const int64_t *offsets,
                   int P, void *rows, int dtype, int D, int inverse, uint8_t *level_out,
                   int32_t *status, void *stream);

/* The same join over up to 1024 CELLS (aligned Morton-prefix subtrees, any of which may be empty): range i holds
 * counts[i] rows, first_keys[i] / last_keys[i] are its first and last Morton code (ignored when counts[i] == 0).
 * A rank that owns several cells transforms them as a FOREST (level 255 on the first row of every cell, see
 * raht_plan_create) and contributes one row per cell.  level_out[C] as above (0 for empty cells).            */
int raht_shard_top_cells(const int64_t *first_keys, const int64_t *last_keys, const int64_t *counts,
                         int C, void *rows, int dtype, int D, int inverse, uint8_t *level_out,
                         int32_t *status, void *stream);

/* Gaussians as packed fp32 rows [means3 | quats4 | scales3 | opacity | colors] for the repartition exchange of
 * the sharded encoder (one collective instead of one per attribute array): out[i] = the row of Gaussian perm[i]
 * (perm == NULL: identity), and the inverse.                                                                */
int raht_pack_rows(const uint32_t *perm, int64_t n, const float *means, const float *quats,
                   const float *scales, const float *opacities, const float *colors, int color_dim,
                   float *out, void *stream);
int raht_unpack_rows(const float *rows, int64_t n, int color_dim, float *means, float *quats,
                     float *scales, float *opacities, float *colors, void *stream);

/* counts[b] (nbins <= 4096 int64, overwritten) = how many values[i] >> shift equal b (larger ones count in the last
 * bin): the cell histogram and the per-owner counts of the sharded partition.                               */
int raht_histogram_i64(const int64_t *values, int64_t n, int shift, int nbins, int64_t *counts,
                       void *stream);

/* Fused forms of raht_forward + raht_quant_reorder and of raht_dequant_unreorder + raht_inverse (same results, bit for
 * bit): the rows a stage-0 tile of the plan finishes -- all but the ones it defers to the upper stages -- are quantised
 * straight out of shared memory (or dequantised straight into it), so the fp64 coefficient matrix never makes its round
 * trip through HBM.  pos[n] (device int32): position of row r in the coded order (inverse permutation of order_RAGFT,
 * raht_order_ragft's second output); steps[D] (device); q: channel-major int32 (D, ld).  T: scratch (n, D) of t_dtype;
 * afterwards only the rows deferred by stage 0 hold coefficients.                                                      */
int raht_forward_quant(const raht_plan *plan, const void *C, int c_dtype, int D, void *T, int t_dtype,
                       const int32_t *pos, const double *steps, int32_t *q, int64_t ld, int32_t *status,
                       int flags, void *stream);
int raht_dequant_inverse(const raht_plan *plan, const int32_t *q, int64_t ld, const int32_t *pos,
                         const double *steps, int D, void *T, int t_dtype, void *C, int flags, void *stream);

/* ------------------------------------------------------------------------------------------------
 * (a9,a10,a11) Quantise + reorder + channel marshalling.  Replaces encode_3dgs.py:204 (floor(T/step
 * +0.5)), :210 (index_select(order_RAGFT)), :215-222 (int32 cast, per-channel sequences) and the
 * mirror :255-268.  steps: D doubles (device) -- one per channel (encode_3dgs_debug.py:376-381);
 * q is channel-major int32: q[c*ld + p] = quantised T[order[p], c].  order == NULL = identity.
 * flags: 0 = true division, what torch computes on the CPU (the parity convention);
 * RAHT_QUANT_RECIPROCAL = T * (1/step), what torch computes on CUDA for tensor / python_scalar (the
 * reference's production device, encode_3dgs.py:22,204); identical for power-of-two steps.
 * ------------------------------------------------------------------------------------------------ */
#define RAHT_QUANT_RECIPROCAL 1
int raht_quant_reorder(const void *T, int t_dtype, int64_t n, int D, const int64_t *order,
                       const double *steps, int32_t *q, int64_t ld, int32_t *status, int flags,
                       void *stream);
int raht_dequant_unreorder(const int32_t *q, int64_t ld, int64_t n, int D, const int64_t *order,
                           const double *steps, void *T, int t_dtype, void *stream);
/* row-major variants without marshalling: Coeff_enc = floor(T/step+0.5) (fp64 out), and q*step */
int raht_quantize_f64(const double *T, int64_t count, double step, double *q, void *stream);
int raht_dequantize_f64(const double *q, int64_t count, double step, double *T, void *stream);

/* ------------------------------------------------------------------------------------------------
 * (a12) RLGR entropy coder.  Replaces python/PyRLGR/src/libs/rlgr/membuf.cpp:340-423 (rlgrWrite),
 * :258-338 (rlgrRead), :46-56 (close), :59-62 (get_buffer) as bound by bindings.cpp:34-57.
 * One independent stream per channel, byte-identical to the reference's.
 * q: channel-major int32 (nch, ld); streams: nch slots of slot_bytes (multiple of 8) bytes;
 * nbytes[c]: closed stream length (exact even when the stream did not fit: RAHT_ST_STREAM_OVERFLOW).
 *
 * sync_index (optional, device, nch * raht_rlgr_sync_entries(n) entries of 8 bytes): side
 * information the encoder can emit next to the reference bitstream -- for every range of 256
 * symbols [256 r, 256 r + 256), raht_rlgr_sync_entries(n) = ceil(n / 256) per channel, the bit offset
 * and coder state (k_P, k_RP) before the range's first event and the position of that event's
 * non-zero symbol inside the range:
 *   word 0 = bit offset (low 32), word 1 = offset bits 32..38 | k_P << 7 | k_RP << 14 | first << 22
 *   (first = 1023: the range holds only zeros).
 * The streams do not depend on it.  raht_rlgr_decode with the index decodes every range in
 * parallel; with sync_index == NULL (a stream written by the reference coder) it decodes one
 * channel per thread, as the format forces.  q is fully written either way.
 * ------------------------------------------------------------------------------------------------ */
int64_t raht_rlgr_sync_entries(int64_t n);
size_t raht_rlgr_encode_workspace_bytes(int64_t n, int nch);
int raht_rlgr_encode(const int32_t *q, int64_t n, int64_t ld, int nch, int flag_signed,
                     uint8_t *streams, int64_t slot_bytes, int64_t *nbytes, void *sync_index,
                     int32_t *status, void *workspace, size_t workspace_bytes, void *stream);
/* diagnostics of the last encode on `workspace`: out6_host = {chunk solves in the speculation pass,
 * their fixed-point rounds, chunks re-solved by the resolve pass, slowest chunk of the speculation
 * pass as SM cycles << 24 | channel << 8 | rounds, rounds spent in the re-solves, chunks settled
 * through the transition tables}.  Synchronises.                                                     */
int raht_rlgr_encode_stats(const void *workspace, int64_t n, int nch, unsigned long long *out6_host,
                           void *stream);
int raht_rlgr_decode(const uint8_t *streams, int64_t slot_bytes, const int64_t *nbytes,
                     const void *sync_index, int64_t n, int64_t ld, int nch, int flag_signed,
                     int32_t *q, void *stream);
/* Sync index of streams that came without one (written by the reference coder, membuf.cpp:340-423): one walk of
 * every stream (a warp per channel: bit positions and coder states only, the stream staged through shared memory
 * ahead of the parsing lane), writing the entries raht_rlgr_encode would have written -- entry r points before the
 * event that holds the first non-zero symbol of range r.  raht_rlgr_decode with that index is then parallel over
 * ranges; the walk is the serial part the format forces (rlgrRead, membuf.cpp:258-338, carries k_P / k_RP from
 * symbol to symbol).  sync_index: nch * raht_rlgr_sync_entries(n) entries of 8 bytes, fully written.            */
int raht_rlgr_build_index(const uint8_t *streams, int64_t slot_bytes, const int64_t *nbytes, int64_t n,
                          int nch, void *sync_index, void *stream);

/* (f3) per-channel range of an (n, D) matrix: minmax[0..D) = column minima, minmax[D..2D) = column maxima (device
 * doubles).  Feeds the per-attribute quantisation steps of encode_3dgs_debug.py:326-381 (range of the coefficients of
 * an attribute group / levels allotted to it); the rule itself is host arithmetic (raht_b200/codec.py).             */
size_t raht_channel_minmax_workspace_bytes(int64_t n, int D);
int raht_channel_minmax(const void *T, int t_dtype, int64_t n, int D, double *minmax, void *workspace,
                        size_t workspace_bytes, void *stream);
/* The same result WITHOUT any allocation and without a caller workspace.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  The per-CTA partial results live in a file-scope device array of the library
 * (592 KiB, held by every GPU that loads the library until exit), which limits D to
 * raht_channel_minmax_static_max_channels() = 64; larger D returns RAHT_E_WORKSPACE with a message.  The call makes no
 * cudaMalloc / cudaFree and no allocating library call (the first call on each device looks the array's device address
 * up with cudaGetSymbolAddress, with that device current).  Every call on a device shares the array: calls that may
 * overlap on the GPU (different streams without ordering) race; calls on one stream are safe.  Nothing in the array
 * needs to start at zero, the previous call's contents are overwritten before they are read.                        */
int raht_channel_minmax_static_max_channels(void);
int raht_channel_minmax_static(const void *T, int t_dtype, int64_t n, int D, double *minmax, void *stream);

/* (a13) sum over rows of (A-B)^2 in fp64 for columns [c0,c1) of two (n, D) matrices -> *sse
 * (device double, overwritten), encode_3dgs.py:298-310.                                            */
int raht_sse_f64(const double *A, const double *B, int64_t n, int D, int c0, int c1, double *sse,
                 void *workspace, size_t workspace_bytes, void *stream);
size_t raht_sse_workspace_bytes(int64_t n, int D);
/* the same with either matrix in fp32 or fp64 (RAHT_F32 / RAHT_F64): the merged attributes are fp32, the
 * reconstruction fp64 (the reference up-casts before subtracting, encode_3dgs.py:298).                */
int raht_sse(const void *A, int a_dtype, const void *B, int b_dtype, int64_t n, int D, int c0, int c1,
             double *sse, void *workspace, size_t workspace_bytes, void *stream);

/* ------------------------------------------------------------------------------------------------
 * (f4) Render evaluator as CUDA compute: the reference renders original and merged / decoded Gaussians from random
 * views with the third-party gsplat.rasterization and compares the images (python/quality_eval.py:283-353
 * render_gaussians, :373-520 try_render_comparison).  gsplat is not vendored in the reference: these entry points
 * restate the published 3DGS / gsplat "classic" rasterisation with gsplat's defaults (near plane 0.01, eps2d 0.3,
 * 3-sigma extent, 16 x 16 tiles, alpha >= 1/255, alpha cap 0.999, transmittance cut 1e-4, colour =
 * max(SH(view direction) + 0.5, 0), quaternions normalised as quality_eval.py:335 does); parity unpinned (checked
 * against oracle/splat_oracle.py).  One view per call sequence, fp32:
 *   raht_splat_project     camera transform, near / frustum cull, EWA projection, conic, tile rectangle and SH colour
 *                          of every Gaussian into the workspace; *n_isect (device) = tiles touched in total.
 *                          viewmat16_host: row-major 4x4 world-to-camera matrix (camera looks down +z), HOST memory;
 *                          colors: (n, color_dim), coefficient-major [k][rgb] (colors.reshape(-1, K, 3), :319-321);
 *                          sh_degree -1: the first three entries are RGB (:324-326).
 *   raht_splat_tile_keys   key = tile << 32 | depth bits, value = Gaussian, m = *n_isect entries (the caller reads it,
 *                          allocates, and sorts with raht_radix_sort_pairs over bits [0, 32 + tile bits))
 *   raht_splat_tile_ranges ranges[2 t], ranges[2 t + 1] = first / one-past-last sorted entry of tile t (n_tiles pairs)
 *   raht_splat_rasterize   front-to-back alpha blending per 16 x 16 tile over background3_host -> image (height,
 *                          width, 3) fp32
 * ------------------------------------------------------------------------------------------------ */
size_t raht_splat_workspace_bytes(int64_t n);
int raht_splat_project(const float *means, const float *quats, const float *scales, const float *opacities,
                       const float *colors, int color_dim, int sh_degree, int64_t n,
                       const float *viewmat16_host, float fx, float fy, float cx, float cy, int width,
                       int height, void *workspace, size_t workspace_bytes, int64_t *n_isect, void *stream);
int raht_splat_tile_keys(void *workspace, size_t workspace_bytes, int64_t n, int width, int height,
                         uint64_t *keys, int32_t *vals, int64_t m, void *stream);
int raht_splat_tile_ranges(const uint64_t *sorted_keys, int64_t m, int n_tiles, int32_t *ranges,
                           void *stream);
int raht_splat_rasterize(void *workspace, size_t workspace_bytes, int64_t n, const int32_t *sorted_vals,
                         const int32_t *ranges, int width, int height, const float *background3_host,
                         float *image, void *stream);

/* ------------------------------------------------------------------------------------------------
 * (f1, f2) Container support: geometry symbols and stream packing.  The reference never serialises
 * its bitstreams (python/encode_3dgs.py:247 takes len(buf) and drops the bytes) and ships the voxel
 * positions uncompressed in a PLY (python/test_voxelize_3dgs.py:284-288); the container format is
 * defined by this build (raht_b200/container.py; checker: oracle/container_oracle.py).
 *
 * raht_geometry_symbols: keys[n] sorted unique Morton codes (>= 0) -> symbols[3][ld] int32:
 *   d[0] = key[0], d[i] = key[i] - key[i-1] - 1; symbols[j][i] = (d[i] >> 21 j) & 0x1fffff.
 *   They are coded with raht_rlgr_encode(..., nch = 3, flag_signed = 0, ...).
 *   status |= RAHT_ST_NOT_SORTED_UNIQUE / RAHT_ST_OUT_OF_BOUNDS (negative key).
 * raht_geometry_keys: the inverse (64-bit inclusive scan).
 * raht_streams_pack: the nch streams of an encoder call (slot c at streams + c * slot_bytes,
 *   nbytes[c] bytes) laid end to end in `out`, each padded with zero bytes to a multiple of 8;
 *   offsets[nch + 1] (device) receives the start of every stream and the total.  A total above
 *   out_capacity sets RAHT_ST_STREAM_OVERFLOW and writes nothing past the capacity.
 * raht_streams_unpack: the inverse, offsets[nch + 1] as written by the packer (device memory).
 * ------------------------------------------------------------------------------------------------ */
int raht_geometry_symbols(const int64_t *keys, int64_t n, int64_t ld, int32_t *symbols,
                          int32_t *status, void *stream);
size_t raht_geometry_keys_workspace_bytes(int64_t n);
int raht_geometry_keys(const int32_t *symbols, int64_t n, int64_t ld, int64_t *keys,
                       void *workspace, size_t workspace_bytes, void *stream);
int raht_streams_pack(const uint8_t *streams, int64_t slot_bytes, const int64_t *nbytes, int nch,
                      uint8_t *out, int64_t out_capacity, int64_t *offsets, int32_t *status,
                      void *stream);
int raht_streams_unpack(const uint8_t *packed, const int64_t *offsets, int nch, uint8_t *streams,
                        int64_t slot_bytes, void *stream);

/* (g) Point-cloud path of python/encode_ply.py: utils.rgb_to_yuv (python/utils.py:4-33), BT.709 full range,
 * rgb (n, 3) fp32 or fp64 in [0, 255] -> yuv (n, 3) fp64 in [0, 255], clipped.  Floating-point row: within 1e-10 of
 * the reference's matmul (whose summation order is the BLAS's).                                              */
int raht_rgb_to_yuv(const void *rgb, int dtype, int64_t n, double *yuv, void *stream);

/* (f1) Scene I/O: the vertex records of the reference's 3DGS PLY files built / split on the device.
 * Replaces the per-vertex Python write loop of python/quality_eval.py:18-117 (save_ply) and the numpy
 * structured-dtype parse of python/data_util.py:272-382 (read_compressed_3dgs_ply).  A record is 14 + color_dim
 * little-endian fp32: x y z | nx ny nz (zeros) | colours | opacity | scale_0..2 | rot_0..3.
 * unpack: v_int (n, 3) int64 = positions truncated (optional), means (n, 3) fp32 (optional), attributes
 * (n, 8 + color_dim) fp32 = [rot 4 | scale 3 | opacity | colours] -- the reader's column order.              */
int raht_ply_pack_vertices(const float *means, const float *quats, const float *scales,
                           const float *opacities, const float *colors, int color_dim, int64_t n,
                           float *records, void *stream);
int raht_ply_unpack_vertices(const float *records, int color_dim, int64_t n, int64_t *v_int,
                             float *means, float *attributes, void *stream);

/* ------------------------------------------------------------------------------------------------
 * Scene-level calls: one leg of the pipeline per host call.  Same kernels, same order and same
 * results as the per-operator entry points above (python/test_voxelize_3dgs.py:198-265 for the
 * front, python/encode_3dgs.py:144-276 for encode / decode); what changes is the host side -- the
 * launches of a leg are issued back to back from C out of ONE caller-owned workspace (256-byte
 * aligned device memory), instead of one allocator round trip and one FFI transition per kernel.
 * The only synchronisation an encode needs is the caller's, between raht_scene_front and
 * raht_scene_encode: the voxel count sizes the second leg.  Nothing here synchronises.
 *
 * events (optional): cudaEvent_t handles, recorded behind the stages of the call in the order
 *   front:  morton, sort, segments            encode: merge, raht_param, raht, quant_reorder, rlgr_encode
 *   decode: raht_param_dec, rlgr_decode, dequant_unreorder, iraht, geometry
 * (NULL entries are skipped), so that the stages can still be timed.
 * *_pinned_host (optional): PINNED host memory that receives the small results without a copy
 *   engine (raht_publish_host); valid once the stream has been synchronised.
 * Two workspaces per encode / decode: the TREE workspace (raht_scene_tree_workspace_bytes) holds what a
 *   caller may keep next to the coded scene -- level, order_RAGFT, its inverse, w and the plan that
 *   *plan points into (raht_plan_destroy frees the host handle only); the scratch workspace holds the
 *   coefficients, the symbols and the coder's workspace and is dead when the call's work is done.
 * layout_host (optional, 11 x int64): byte offsets of level (nvox x u8), order (nvox x i64), position
 *   (nvox x i32), w (nvox x f64), the plan workspace and its size IN THE TREE WORKSPACE; of packed
 *   (nvox x D f32; -1 when the caller supplied packed_out), T (nvox x D of t_dtype), q (D x ld int32),
 *   ld itself and steps (D x f64) IN THE SCRATCH -- every intermediate can be looked at without a copy.
 * ------------------------------------------------------------------------------------------------ */
size_t raht_scene_front_workspace_bytes(int64_t n, int J);
/* means (n,3) fp32 -> params4 / params_d2 (raht_voxel_params; use_given != 0: quantise against the
 * caller's vmin_host[3] / width_host), unique_keys (n x u64, first *nvox valid), sort_idx (n x u32:
 * the stable sort permutation = cluster_indices), cluster_offsets ((n + 1) x i32, first *nvox + 1
 * valid), *nvox (device int64) and, when given, front_pinned_host[0..8) = nvox, [8..24) = params_d2. */
int raht_scene_front(const float *means, int64_t n, int J, int use_given, const float *vmin_host,
                     double width_host, float *params4, double *params_d2, uint64_t *unique_keys,
                     uint32_t *sort_idx, int32_t *cluster_offsets, int64_t *nvox,
                     void *front_pinned_host, int32_t *status, void *workspace, size_t workspace_bytes,
                     void **events, void *stream);
size_t raht_scene_tree_workspace_bytes(int64_t nvox, int tile_rows);
size_t raht_scene_encode_workspace_bytes(int64_t nvox, int D, int t_dtype, int own_packed);
/* D = 11 + color_dim channels [means3 quats4 scales3 opacity1 colors]; steps_host: D doubles (host).
 * packed_out (nvox x D fp32, optional): the merged attributes (else kept in the workspace).
 * streams / slot_bytes / nbytes / sync_index / status as raht_rlgr_encode.  results_pinned_host
 * (8 D + 4 bytes): the D stream lengths, then the status word.                                     */
int raht_scene_encode(const uint32_t *sort_idx, const int32_t *cluster_offsets, int64_t nvox,
                      const float *means, const float *quats, const float *scales,
                      const float *opacities, const float *colors, int color_dim,
                      const uint64_t *keys, const double *steps_host, int t_dtype, int tile_rows,
                      void *packed_out, uint8_t *streams, int64_t slot_bytes, int64_t *nbytes,
                      void *sync_index, int32_t *status, void *results_pinned_host, raht_plan **plan,
                      int64_t *layout_host, void *tree_workspace, size_t tree_workspace_bytes,
                      void *workspace, size_t workspace_bytes, void **events, void *stream);
size_t raht_scene_decode_workspace_bytes(int64_t nvox, int D, int t_dtype);
/* keys (nvox sorted unique Morton codes) + the D streams -> attributes (nvox x D of t_dtype) and,
 * when coords != NULL, the voxel coordinates (nvox x 3 fp32).  The tree is rebuilt from the keys.  */
int raht_scene_decode(const uint64_t *keys, int64_t nvox, int J, const uint8_t *streams,
                      int64_t slot_bytes, const int64_t *nbytes, const void *sync_index, int D,
                      const double *steps_host, int t_dtype, int tile_rows, void *attributes,
                      float *coords, int32_t *status, void *status_pinned_host, raht_plan **plan,
                      int64_t *layout_host, void *tree_workspace, size_t tree_workspace_bytes,
                      void *workspace, size_t workspace_bytes, void **events, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* RAHT_B200_H */
