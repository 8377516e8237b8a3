/*
 * raht_oracle.c -- CPU ORACLE for the RAHT-3DGS hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This file is a plain-C restatement of the reference's algorithms (Daaohame/RAHT-3DGS-codec,
 * commit ef45e2e) for the hot path of BASELINE.json:north_star.  It is the CHECKER that the
 * CUDA product path is compared against.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it.  The product path never links, imports or
 * calls anything in oracle/.
 *
 * Parity pinning: the reference ships no golden vectors.  This restatement
 * is pinned (i) against the reference's own Python/C++ functions executed in the build container
 * (tests/test_oracle_vs_reference.py, skipped when /root/reference is absent), (ii) against
 * fixtures generated from those functions (tests/golden/, script tests/golden/make_golden.py),
 * (iii) against the known-answer vectors of SURVEY.md appendix B.
 *
 * Every function follows the reference's control flow (per-level list compaction, per-level
 * gather/butterfly/scatter, symbol-serial RLGR) -- deliberately NOT the closed forms the CUDA
 * path uses -- so that the two are independent derivations.
 *
 * Floating point: compile with -ffp-contract=off (no FMA contraction), no -ffast-math.
 * All file:line citations are relative to /root/reference/.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------------------------ */
/* (a1) Morton code.  python/voxelize_pc.py:25-59 (same math RAHT_param.py:30-36, :208-212).  */
/* digit_i = z_i + 2*y_i + 4*x_i placed at bits 3i..3i+2.                                      */
/* ------------------------------------------------------------------------------------------ */
static inline int64_t morton3(int64_t x, int64_t y, int64_t z, int J)
{
    int64_t mc = 0;
    for (int i = 1; i <= J; ++i) {
        int64_t bx = (x >> (i - 1)) & 1, by = (y >> (i - 1)) & 1, bz = (z >> (i - 1)) & 1;
        int64_t digit = bz + (by << 1) + (bx << 2);
        mc |= digit << (3 * (i - 1));
    }
    return mc;
}

ORC_API void orc_morton(const int64_t *V, int64_t n, int J, int64_t *MC)
{
    for (int64_t i = 0; i < n; ++i)
        MC[i] = morton3(V[3 * i], V[3 * i + 1], V[3 * i + 2], J);
}

/* RAHT_param.py:21-27 / :201-212: Vint = floor((V - minV)/Q) in fp64, Q = width / 2^depth.
 * Returns 0, or -1 when any index is outside [0, 2^depth-1] (ValueError in RAHT_param.py:26-27). */
ORC_API int orc_morton_from_f64(const double *V, int64_t n, const double *minV, double width,
                                int depth, int64_t *MC)
{
    double Q = width / (double)((int64_t)1 << depth);
    int bad = 0;
    int64_t hi = ((int64_t)1 << depth) - 1;
    for (int64_t i = 0; i < n; ++i) {
        int64_t c[3];
        for (int d = 0; d < 3; ++d) {
            c[d] = (int64_t)floor((V[3 * i + d] - minV[d]) / Q);
            if (c[d] < 0 || c[d] > hi) bad = 1;
        }
        MC[i] = morton3(c[0], c[1], c[2], depth);
    }
    return bad ? -1 : 0;
}

/* ------------------------------------------------------------------------------------------ */
/* (a2) voxelize_pc_batched.  python/voxelize_pc.py:62-172.                                    */
/* All arithmetic in fp32 as torch does for a float32 PC: V0 = V - vmin; width = max(V0) as a  */
/* Python float (double); voxel_size = width / 2**J (double); V0 / voxel_size is a float32     */
/* tensor divided by a Python scalar -> the scalar is rounded to float32 and the division is   */
/* an IEEE fp32 division (torch CPU).  torch.sort is replaced by a STABLE sort. */
/* Outputs: Vint (n,3) int64, MC (n), sort_idx (n), voxel_indices (<= n; count returned),      */
/* vmin_out[3], *width_out, *voxel_size_out.  DeltaV (n,3) fp32 optional (may be NULL).        */
/* ------------------------------------------------------------------------------------------ */
static void stable_sort_u64(const int64_t *keys, int64_t n, int64_t *sorted, int64_t *idx)
{
    /* LSD radix sort, 8 bits per pass: stable. keys are non-negative (< 2^63). */
    int64_t *k0 = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n ? n : 1));
    int64_t *k1 = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n ? n : 1));
    int64_t *i0 = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n ? n : 1));
    int64_t *i1 = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n ? n : 1));
    for (int64_t i = 0; i < n; ++i) { k0[i] = keys[i]; i0[i] = i; }
    for (int pass = 0; pass < 8; ++pass) {
        int64_t cnt[257];
        memset(cnt, 0, sizeof(cnt));
        int sh = 8 * pass;
        for (int64_t i = 0; i < n; ++i) cnt[((uint64_t)k0[i] >> sh & 0xff) + 1]++;
        for (int b = 0; b < 256; ++b) cnt[b + 1] += cnt[b];
        for (int64_t i = 0; i < n; ++i) {
            int64_t p = cnt[(uint64_t)k0[i] >> sh & 0xff]++;
            k1[p] = k0[i]; i1[p] = i0[i];
        }
        int64_t *t = k0; k0 = k1; k1 = t;
        t = i0; i0 = i1; i1 = t;
    }
    for (int64_t i = 0; i < n; ++i) { sorted[i] = k0[i]; idx[i] = i0[i]; }
    free(k0); free(k1); free(i0); free(i1);
}

ORC_API int64_t orc_voxelize(const float *V, int64_t n, int J,
                             const float *vmin_in /* NULL => min(V) */,
                             const double *width_in /* NULL => max(V0) */,
                             int64_t *Vint, int64_t *MC, int64_t *MC_sorted, int64_t *sort_idx,
                             int64_t *voxel_indices, float *DeltaV,
                             float *vmin_out, double *width_out, double *voxel_size_out)
{
    float vmin[3];
    if (vmin_in) { vmin[0] = vmin_in[0]; vmin[1] = vmin_in[1]; vmin[2] = vmin_in[2]; }
    else {
        for (int d = 0; d < 3; ++d) {                    /* voxelize_pc.py:87-88 */
            float m = V[d];
            for (int64_t i = 1; i < n; ++i) if (V[3 * i + d] < m) m = V[3 * i + d];
            vmin[d] = m;
        }
    }
    double width;
    if (width_in) width = *width_in;
    else {                                               /* :92-95  width = V0.max().item() */
        float mx = V[0] - vmin[0];
        for (int64_t i = 0; i < n; ++i)
            for (int d = 0; d < 3; ++d) {
                float v0 = V[3 * i + d] - vmin[d];
                if (v0 > mx) mx = v0;
            }
        width = (double)mx;
    }
    double voxel_size = width / (double)((int64_t)1 << J);   /* :97 */
    float vs = (float)voxel_size;                             /* scalar promoted to the tensor dtype */
    int64_t hi = ((int64_t)1 << J) - 1;
    for (int64_t i = 0; i < n; ++i) {                         /* :98 */
        for (int d = 0; d < 3; ++d) {
            float v0 = V[3 * i + d] - vmin[d];
            int64_t q = (int64_t)floorf(v0 / vs);
            if (q < 0) q = 0;
            if (q > hi) q = hi;
            Vint[3 * i + d] = q;
        }
        MC[i] = morton3(Vint[3 * i], Vint[3 * i + 1], Vint[3 * i + 2], J);   /* :100 */
    }
    stable_sort_u64(MC, n, MC_sorted, sort_idx);              /* :101 (stable) */
    if (DeltaV) {                                             /* :110-111 */
        for (int64_t i = 0; i < n; ++i)
            for (int d = 0; d < 3; ++d) {
                float v0 = V[3 * sort_idx[i] + d] - vmin[d];
                float vox = vs * floorf(v0 / vs);
                DeltaV[3 * i + d] = v0 - vox;
            }
    }
    int64_t nvox = 0;                                         /* :114-120 */
    for (int64_t i = 0; i < n; ++i)
        if (i == 0 || MC_sorted[i] != MC_sorted[i - 1]) voxel_indices[nvox++] = i;
    vmin_out[0] = vmin[0]; vmin_out[1] = vmin[1]; vmin_out[2] = vmin[2];
    *width_out = width; *voxel_size_out = voxel_size;
    return nvox;
}

/* voxelize_pc.py:122-153: per-voxel attribute mean (fp32, scatter_add in sorted order, then / count). */
ORC_API void orc_voxel_attr_mean(const float *C /* (n,d) original order */, int64_t n, int d,
                                 const int64_t *sort_idx, const int64_t *voxel_indices, int64_t nvox,
                                 float *Cvox /* (nvox,d) */)
{
    for (int64_t v = 0; v < nvox; ++v) {
        int64_t s = voxel_indices[v], e = (v + 1 < nvox) ? voxel_indices[v + 1] : n;
        float cnt = (float)(e - s);
        for (int c = 0; c < d; ++c) {
            float acc = 0.0f;
            for (int64_t i = s; i < e; ++i) acc += C[sort_idx[i] * d + c];
            Cvox[v * d + c] = acc / cnt;
        }
    }
}

/* ------------------------------------------------------------------------------------------ */
/* (a4) merge_weighted_mean_kernel.  cuda/merge_cluster.cu:2-111, one loop iteration per        */
/* cluster_id in place of one CUDA thread.  fp32, mul then add separately rounded.             */
/* ------------------------------------------------------------------------------------------ */
ORC_API void orc_merge_clusters(const int32_t *cluster_indices, const int32_t *cluster_offsets,
                                int64_t num_clusters,
                                const float *means, const float *quats, const float *scales,
                                const float *opacities, const float *colors, int color_dim,
                                int weight_by_opacity,
                                float *m_means, float *m_quats, float *m_scales,
                                float *m_opac, float *m_colors)
{
#pragma omp parallel for schedule(static)
    for (int64_t cid = 0; cid < num_clusters; ++cid) {
        int start = cluster_offsets[cid], end = cluster_offsets[cid + 1];
        if (end - start == 0) continue;                                   /* :27-29 */
        float ma[3] = {0, 0, 0}, qa[4] = {0, 0, 0, 0}, sa[3] = {0, 0, 0};
        float opacity_sum = 0.0f, total_weight = 0.0f;
        for (int i = start; i < end; ++i) {                               /* :38-62 */
            int idx = cluster_indices[i];
            float w = weight_by_opacity ? opacities[idx] : 1.0f;
            total_weight += w;
            for (int k = 0; k < 3; ++k) { float p = means[idx * 3 + k] * w; ma[k] += p; }
            for (int k = 0; k < 4; ++k) { float p = quats[idx * 4 + k] * w; qa[k] += p; }
            for (int k = 0; k < 3; ++k) { float p = scales[idx * 3 + k] * w; sa[k] += p; }
            opacity_sum += opacities[idx];
        }
        if (total_weight == 0.0f) total_weight = 1.0f;                    /* :65-67 */
        for (int k = 0; k < 3; ++k) m_means[cid * 3 + k] = ma[k] / total_weight;
        float s0 = qa[0] * qa[0], s1 = qa[1] * qa[1], s2 = qa[2] * qa[2], s3 = qa[3] * qa[3];
        float qn = sqrtf(((s0 + s1) + s2) + s3);                          /* :75-76 */
        if (qn > 0.0f) for (int k = 0; k < 4; ++k) m_quats[cid * 4 + k] = qa[k] / qn;
        else { m_quats[cid * 4] = 0; m_quats[cid * 4 + 1] = 0; m_quats[cid * 4 + 2] = 0; m_quats[cid * 4 + 3] = 1.0f; }
        for (int k = 0; k < 3; ++k) m_scales[cid * 3 + k] = sa[k] / total_weight;
        m_opac[cid] = fminf(opacity_sum, 1.0f);                           /* :94 */
        for (int c = 0; c < color_dim; ++c) {                             /* :97-110 */
            float acc = 0.0f, wsum = 0.0f;
            for (int i = start; i < end; ++i) {
                int idx = cluster_indices[i];
                float w = weight_by_opacity ? opacities[idx] : 1.0f;
                float p = colors[(int64_t)idx * color_dim + c] * w;
                acc += p; wsum += w;
            }
            m_colors[cid * (int64_t)color_dim + c] = (wsum > 0.0f) ? acc / wsum : 0.0f;
        }
    }
}

/* ------------------------------------------------------------------------------------------ */
/* (a5/a6) RAHT_param / RAHT_param_reorder_fast.  python/RAHT_param.py:6-83, :191-279.         */
/* Per-level loop with list compaction, exactly as the reference (incl. its N<=2 order quirk). */
/* ------------------------------------------------------------------------------------------ */
typedef struct {
    int64_t n;
    int nlevels;             /* len(Flags) == len(weights) */
    int nlists;              /* len(List) (== nlevels in every reachable case) */
    int64_t *len;            /* per level list length */
    int64_t **list;
    uint8_t **flags;
    int64_t **weights;
    int64_t *order;          /* order_RAGFT */
    int64_t order_len;       /* -1 => None (N==1) */
} orc_tree;

ORC_API orc_tree *orc_tree_build(const int64_t *MC, int64_t n, int depth)
{
    orc_tree *t = (orc_tree *)calloc(1, sizeof(orc_tree));
    int Nbits = 3 * depth;
    t->n = n;
    t->len = (int64_t *)calloc(66, sizeof(int64_t));
    t->list = (int64_t **)calloc(66, sizeof(int64_t *));
    t->flags = (uint8_t **)calloc(66, sizeof(uint8_t *));
    t->weights = (int64_t **)calloc(66, sizeof(int64_t *));
    t->order_len = -1;

    int64_t L = n;
    int64_t *cur = (int64_t *)malloc(sizeof(int64_t) * (size_t)(n ? n : 1));
    for (int64_t i = 0; i < n; ++i) cur[i] = i;            /* RAHT_param.py:214 */
    t->list[0] = cur; t->len[0] = L; t->nlists = 1;

    /* RA-GFT bookkeeping (:219-222): ac_list of index arrays, appended in coarse<-fine order */
    uint8_t *indices = (uint8_t *)calloc((size_t)(n ? n : 1), 1);
    uint8_t *pre = (uint8_t *)calloc((size_t)(n ? n : 1), 1);
    int64_t **ac = (int64_t **)calloc(80, sizeof(int64_t *));
    int64_t *ac_len = (int64_t *)calloc(80, sizeof(int64_t));
    int nac = 0;

    for (int j = 1; j < 65; ++j) {
        int lev = j - 1;
        int64_t *w = (int64_t *)malloc(sizeof(int64_t) * (size_t)L);     /* :228-230 */
        for (int64_t k = 0; k < L; ++k) w[k] = ((k + 1 < L) ? cur[k + 1] : n) - cur[k];
        t->weights[lev] = w;
        uint8_t *fl = (uint8_t *)calloc((size_t)L, 1);
        t->flags[lev] = fl;
        t->nlevels = lev + 1;
        if (L == 1) break;                                                /* :233-236 */
        uint64_t mask = ((uint64_t)1 << Nbits) - ((uint64_t)1 << j);      /* :224 */
        for (int64_t k = 0; k + 1 < L; ++k)                               /* :238-243 */
            fl[k] = ((((uint64_t)MC[cur[k]] ^ (uint64_t)MC[cur[k + 1]]) & mask) == 0);
        int64_t Lt = 0;                                                   /* :246-248 */
        int64_t *tmp = (int64_t *)malloc(sizeof(int64_t) * (size_t)L);
        for (int64_t k = 0; k < L; ++k) if (!(k > 0 && fl[k - 1])) tmp[Lt++] = cur[k];

        if (j % 3 == 0 && j > 2) {                                        /* :251-262 */
            memset(indices, 0, (size_t)n);
            for (int64_t k = 0; k < Lt; ++k) indices[tmp[k]] = 1;
            int64_t cnt = 0;
            int64_t *rem = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
            if (j == 3) { for (int64_t i = 0; i < n; ++i) if (!indices[i]) rem[cnt++] = i; }
            else        { for (int64_t i = 0; i < n; ++i) if (indices[i] ^ pre[i]) rem[cnt++] = i; }
            ac[nac] = rem; ac_len[nac++] = cnt;
            memcpy(pre, indices, (size_t)n);
        }
        if (Lt == 1 || j >= Nbits) {                                      /* :265-274 */
            memset(indices, 0, (size_t)n);
            for (int64_t k = 0; k < Lt; ++k) indices[tmp[k]] = 1;
            int64_t cnt = 0;
            int64_t *rem = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
            for (int64_t i = 0; i < n; ++i) if (indices[i] ^ pre[i]) rem[cnt++] = i;
            ac[nac] = rem; ac_len[nac++] = cnt;
            ac[nac] = tmp; ac_len[nac++] = Lt;
            int64_t tot = 0;
            for (int a = 0; a < nac; ++a) tot += ac_len[a];
            t->order = (int64_t *)malloc(sizeof(int64_t) * (size_t)(tot ? tot : 1));
            t->order_len = 0;
            for (int a = nac - 1; a >= 0; --a) {
                memcpy(t->order + t->order_len, ac[a], sizeof(int64_t) * (size_t)ac_len[a]);
                t->order_len += ac_len[a];
            }
            break;
        }
        cur = tmp; L = Lt;                                                /* :276-277 */
        t->list[t->nlists] = cur; t->len[t->nlists] = L; t->nlists++;
    }
    for (int a = 0; a < nac; ++a) free(ac[a]);   /* note: when we broke out, ac[nac-1]==tmp is freed here */
    free(ac); free(ac_len); free(indices); free(pre);
    return t;
}

ORC_API void orc_tree_free(orc_tree *t)
{
    if (!t) return;
    for (int l = 0; l < 66; ++l) { free(t->list[l]); free(t->flags[l]); free(t->weights[l]); }
    free(t->len); free(t->list); free(t->flags); free(t->weights); free(t->order); free(t);
}
ORC_API int orc_tree_nlevels(const orc_tree *t) { return t->nlevels; }
ORC_API int orc_tree_nlists(const orc_tree *t) { return t->nlists; }
ORC_API int64_t orc_tree_level_len(const orc_tree *t, int l) { return t->len[l]; }
ORC_API void orc_tree_level_copy(const orc_tree *t, int l, int64_t *list, uint8_t *flags, int64_t *weights)
{
    size_t L = (size_t)t->len[l];
    if (list) memcpy(list, t->list[l], L * sizeof(int64_t));
    if (flags) memcpy(flags, t->flags[l], L);
    if (weights) memcpy(weights, t->weights[l], L * sizeof(int64_t));
}
ORC_API int64_t orc_tree_order_len(const orc_tree *t) { return t->order_len; }
ORC_API void orc_tree_order_copy(const orc_tree *t, int64_t *order)
{
    if (t->order_len > 0) memcpy(order, t->order, sizeof(int64_t) * (size_t)t->order_len);
}

/* ------------------------------------------------------------------------------------------ */
/* (a7) RAHT2_optimized.  python/RAHT.py:253-336.  fp64; a = sqrt(w0/(w0+w1)), b = sqrt(w1/..); */
/* T[i0] = a*x0 + b*x1 ; T[i1] = (-b)*x0 + a*x1, each product and the sum separately rounded.  */
/* ------------------------------------------------------------------------------------------ */
ORC_API void orc_raht_forward(const orc_tree *t, const double *C, int D, double *T, double *w)
{
    int64_t n = t->n;
    memcpy(T, C, sizeof(double) * (size_t)n * (size_t)D);         /* :285 */
    for (int64_t i = 0; i < n; ++i) w[i] = 1.0;                    /* :286 */
    for (int j = 0; j < t->nlevels; ++j) {                          /* :293 */
        if (j >= t->nlists) break;
        const int64_t *list = t->list[j]; const uint8_t *fl = t->flags[j]; const int64_t *wt = t->weights[j];
        int64_t L = t->len[j];
#pragma omp parallel for schedule(static)
        for (int64_t k = 0; k < L - 1; ++k) {
            if (!fl[k]) continue;                                   /* left_mask = Flags[j] (:296) */
            int64_t i0 = list[k], i1 = list[k + 1];                 /* :301-302 */
            double w0 = (double)wt[k], w1 = (double)wt[k + 1];      /* :315-316 */
            double denom = w0 + w1;
            double a = sqrt(w0 / denom), b = sqrt(w1 / denom);      /* :321-322 */
            double nw = w[i0] + w[i1];                              /* :325-328 */
            w[i0] = nw; w[i1] = nw;
            double nb = -b;
            double *r0 = T + i0 * D, *r1 = T + i1 * D;
            for (int c = 0; c < D; ++c) {                           /* :331-334 */
                double x0 = r0[c], x1 = r1[c];
                double p0 = a * x0, p1 = b * x1, p2 = nb * x0, p3 = a * x1;
                r0[c] = p0 + p1;
                r1[c] = p2 + p3;
            }
        }
    }
}

/* (a8) inverse_RAHT_optimized.  python/iRAHT.py:40-114.  x0 = a*T0 - b*T1 ; x1 = b*T0 + a*T1. */
ORC_API void orc_raht_inverse(const orc_tree *t, const double *T, int D, double *C)
{
    int64_t n = t->n;
    memcpy(C, T, sizeof(double) * (size_t)n * (size_t)D);          /* :69 */
    for (int j = t->nlevels - 1; j >= 0; --j) {                      /* :76 */
        if (j >= t->nlists) continue;
        const int64_t *list = t->list[j]; const uint8_t *fl = t->flags[j]; const int64_t *wt = t->weights[j];
        int64_t L = t->len[j];
#pragma omp parallel for schedule(static)
        for (int64_t k = 0; k < L - 1; ++k) {
            if (!fl[k]) continue;
            int64_t i0 = list[k], i1 = list[k + 1];
            double w0 = (double)wt[k], w1 = (double)wt[k + 1];
            double denom = w0 + w1;
            double a = sqrt(w0 / denom), b = sqrt(w1 / denom);
            double *r0 = C + i0 * D, *r1 = C + i1 * D;
            for (int c = 0; c < D; ++c) {                            /* :107-111 */
                double t0 = r0[c], t1 = r1[c];
                double p0 = a * t0, p1 = b * t1, p2 = b * t0, p3 = a * t1;
                r0[c] = p0 - p1;
                r1[c] = p2 + p3;
            }
        }
    }
}

/* (a9) quantise / dequantise.  python/encode_3dgs.py:204, :261.  q = floor(T/step + 0.5). */
ORC_API void orc_quantize(const double *T, int64_t count, double step, double *q)
{
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < count; ++i) { double d = T[i] / step; q[i] = floor(d + 0.5); }
}
ORC_API void orc_dequantize(const double *q, int64_t count, double step, double *T)
{
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < count; ++i) T[i] = q[i] * step;
}

/* ------------------------------------------------------------------------------------------ */
/* (a12) RLGR.  python/PyRLGR/src/libs/rlgr/membuf.cpp: bit buffer :74-86,:166-189, close       */
/* :46-56, GR :228-256, rlgrRead :258-338, rlgrWrite :340-423.  Constants membuf.h:18-22.      */
/* ------------------------------------------------------------------------------------------ */
#define RL_L 4
#define RL_U0 3
#define RL_D0 1
#define RL_U1 2
#define RL_D1 1
#define RL_MASK(k) ((((uint64_t)1) << (k)) - 1)

typedef struct { uint64_t data; unsigned bits; uint8_t *buf; size_t pos, cap; int overflow; } bitw;

static void bw_flush(bitw *w)
{
    while (w->bits >= 8) {
        w->bits -= 8;
        uint8_t byte = (uint8_t)((w->data >> w->bits) & 0xff);
        if (w->pos < w->cap) w->buf[w->pos] = byte; else w->overflow = 1;
        w->pos++;
    }
}
static void bw_write1(bitw *w, unsigned bit)
{
    w->data <<= 1; if (bit) w->data++; w->bits++;
    if (w->bits >= 8) bw_flush(w);
}
static void bw_write(bitw *w, uint64_t data, unsigned bits)
{
    if (bits > 56) { bw_write(w, data >> 32, bits - 32); bw_write(w, data & RL_MASK(32), 32); return; }
    w->data = (w->data << bits) + data; w->bits += bits; bw_flush(w);
}
static void bw_gr(bitw *w, uint64_t data, unsigned bits)
{
    uint64_t p = data >> bits;
    if (p < 32) { bw_write(w, RL_MASK(p + 1) - 1, (unsigned)(p + 1)); bw_write(w, data & RL_MASK(bits), bits); }
    else { bw_write(w, RL_MASK(32), 32); bw_write(w, data, 32); }
}
static inline uint64_t s2u(int64_t v) { return v < 0 ? (((uint64_t)(-v)) << 1) - 1 : ((uint64_t)v) << 1; }
static inline int64_t u2s(uint64_t v) { int64_t d = (int64_t)(v >> 1); return (v & 1) ? -d - 1 : d; }

/* Returns the number of bytes of the closed stream (may exceed cap => caller must retry). */
ORC_API int64_t orc_rlgr_encode(const int64_t *seq, int64_t N, int flagSigned, uint8_t *out, int64_t cap)
{
    bitw w; memset(&w, 0, sizeof(w)); w.buf = out; w.cap = (size_t)cap;
    uint64_t u = 0, k_P = 0, k_RP = 2 * RL_L, m = 0, k = 0, k_R, p;
    if (N == 0) return 0;     /* reference reads an uninitialised k here; emits nothing in practice */
    for (int64_t n = 0; n < N; ++n) {
        u = flagSigned ? s2u(seq[n]) : (uint64_t)seq[n];
        k = k_P / RL_L; k_R = k_RP / RL_L;
        if (k) {
            if (u) {
                u--;
                bw_write1(&w, 0); bw_write(&w, m, (unsigned)k); bw_gr(&w, u, (unsigned)k_R);
                p = u >> k_R;
                if (p) { k_RP += p - 1; if (k_RP > 32 * RL_L) k_RP = 32 * RL_L; }
                else { if (k_RP < 2) k_RP = 0; else k_RP -= 2; }
                if (k_P < RL_D1) k_P = 0; else k_P -= RL_D1;
                m = 0;
            } else {
                m++;
                if (m == ((uint64_t)1 << k)) { bw_write1(&w, 1); k_P += RL_U1; m = 0; }
            }
        } else {
            bw_gr(&w, u, (unsigned)k_R);
            p = u >> k_R;
            if (p) { k_RP = k_RP + p - 1; if (k_RP > 32 * RL_L) k_RP = 32 * RL_L; }
            else { if (k_RP < 2) k_RP = 0; else k_RP -= 2; }
            if (u) { if (k_P < RL_D0) k_P = 0; else k_P -= RL_D0; } else k_P += RL_U0;
            m = 0;
        }
    }
    if (k && !u) { bw_write1(&w, 0); bw_write(&w, m, (unsigned)(k_P / RL_L)); }   /* :416-419 */
    unsigned r = w.bits % 8;                                                       /* close() :46-56 */
    if (r) bw_write(&w, 0, 8 - r); else bw_flush(&w);
    return (int64_t)w.pos;
}

typedef struct { uint64_t data; unsigned bits; const uint8_t *buf; size_t pos, size; } bitr;
static void br_fill(bitr *r)
{
    while (r->bits <= 56) {
        if (r->pos < r->size) { r->data = (r->data << 8) + r->buf[r->pos++]; r->bits += 8; }
        else return;
    }
}
static unsigned br_read1(bitr *r)
{
    if (!r->bits) br_fill(r);
    if (!r->bits) return 0;           /* past the end: the reference would underflow; treat as 0 */
    r->bits--; return (unsigned)((r->data >> r->bits) & 1);
}
static uint64_t br_read(bitr *r, unsigned bits)
{
    if (bits > 56) { uint64_t d = br_read(r, bits - 32) << 32; return d + br_read(r, 32); }
    br_fill(r);
    if (r->bits < bits) { uint64_t v = (r->data & RL_MASK(r->bits)) << (bits - r->bits); r->bits = 0; return v; }
    r->bits -= bits; return (r->data >> r->bits) & RL_MASK(bits);
}
static uint64_t br_gr(bitr *r, unsigned bits)
{
    uint64_t p = 0;
    while (br_read1(r)) { p++; if (p >= 32) return br_read(r, 32); }
    return (p << bits) + br_read(r, bits);
}

ORC_API void orc_rlgr_decode(const uint8_t *buf, int64_t nbytes, int64_t N, int flagSigned, int64_t *seq)
{
    bitr r; memset(&r, 0, sizeof(r)); r.buf = buf; r.size = (size_t)nbytes;
    uint64_t u, k_P = 0, k_RP = 2 * RL_L, m = 0, k, k_R, p;
    int64_t n = 0;
    while (n < N) {
        k = k_P / RL_L; k_R = k_RP / RL_L;
        if (k) {
            m = 0;
            while (br_read1(&r)) { m += (uint64_t)1 << k; k_P += RL_U1; k = k_P / RL_L; }
            m += br_read(&r, (unsigned)k);
            while (m--) { if (n < N) seq[n] = 0; n++; }    /* write guard only; a valid stream never overruns */
            if (n >= N) break;
            u = br_gr(&r, (unsigned)k_R);
            seq[n++] = flagSigned ? u2s(u + 1) : (int64_t)(u + 1);
            p = u >> k_R;
            if (p) { k_RP += p - 1; if (k_RP > 32 * RL_L) k_RP = 32 * RL_L; }
            else { if (k_RP < 2) k_RP = 0; else k_RP -= 2; }
            if (k_P < RL_D1) k_P = 0; else k_P -= RL_D1;
        } else {
            u = br_gr(&r, (unsigned)k_R);
            seq[n++] = flagSigned ? u2s(u) : (int64_t)u;
            p = u >> k_R;
            if (p) { k_RP = k_RP + p - 1; if (k_RP > 32 * RL_L) k_RP = 32 * RL_L; }
            else { if (k_RP < 2) k_RP = 0; else k_RP -= 2; }
            if (u) { if (k_P < RL_D0) k_P = 0; else k_P -= RL_D0; } else k_P += RL_U0;
        }
    }
}

/* Channel-parallel helpers for the CPU baseline legs of bench.py (one OpenMP thread per channel,
 * the parallelism the reference's per-channel loop offers: encode_3dgs.py:229-245).
 * q is channel-major int32 (nch, ld). out holds nch slots of cap bytes; nbytes[ch] returned. */
ORC_API void orc_rlgr_encode_channels(const int32_t *q, int64_t N, int64_t ld, int nch, int flagSigned,
                                      uint8_t *out, int64_t cap, int64_t *nbytes)
{
#pragma omp parallel for schedule(dynamic, 1)
    for (int ch = 0; ch < nch; ++ch) {
        int64_t *seq = (int64_t *)malloc(sizeof(int64_t) * (size_t)(N ? N : 1));
        for (int64_t i = 0; i < N; ++i) seq[i] = q[ch * ld + i];
        nbytes[ch] = orc_rlgr_encode(seq, N, flagSigned, out + (size_t)ch * (size_t)cap, cap);
        free(seq);
    }
}
ORC_API void orc_rlgr_decode_channels(const uint8_t *in, int64_t cap, const int64_t *nbytes, int64_t N,
                                      int64_t ld, int nch, int flagSigned, int32_t *q)
{
#pragma omp parallel for schedule(dynamic, 1)
    for (int ch = 0; ch < nch; ++ch) {
        int64_t *seq = (int64_t *)malloc(sizeof(int64_t) * (size_t)(N ? N : 1));
        orc_rlgr_decode(in + (size_t)ch * (size_t)cap, nbytes[ch], N, flagSigned, seq);
        for (int64_t i = 0; i < N; ++i) q[ch * ld + i] = (int32_t)seq[i];
        free(seq);
    }
}

/* python/utils.py:4-33 rgb_to_yuv (BT.709 matrix, clamp, *255), fp64. */
ORC_API void orc_rgb_to_yuv(const double *rgb, int64_t n, double *yuv)
{
    static const double Q[4][3] = {
        {0.21260000, -0.114572, 0.5}, {0.71520000, -0.385428, -0.454153},
        {0.07220000, 0.5, -0.045847}, {0.0, 0.50196078, 0.50196078}};
    for (int64_t i = 0; i < n; ++i) {
        double r[4] = {rgb[3 * i] / 255.0, rgb[3 * i + 1] / 255.0, rgb[3 * i + 2] / 255.0, 1.0};
        for (int c = 0; c < 3; ++c) {
            double acc = 0.0;
            for (int k = 0; k < 4; ++k) acc += r[k] * Q[k][c];
            if (acc < 0.0) acc = 0.0;
            if (acc > 1.0) acc = 1.0;
            yuv[3 * i + c] = acc * 255.0;
        }
    }
}

ORC_API int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
ORC_API void orc_set_num_threads(int t)
{
#ifdef _OPENMP
    omp_set_num_threads(t);
#else
    (void)t;
#endif
}
