// transform.cu -- forward / inverse RAHT butterflies.
// Replaces python/RAHT.py:253-336 (RAHT2_optimized) and python/iRAHT.py:40-114
// (inverse_RAHT_optimized): instead of a gather / butterfly / scatter chain per bit level (which
// moves every surviving row through HBM at every level), each CTA stages a tile of rows in shared
// memory -- one bulk async copy (TMA, UBLKCP) for the Morton-contiguous stage-0 tiles -- runs every
// butterfly of the plan's schedule for that tile there, round by round, and writes the tile back
// once.  Stage 0 therefore reads and writes each row exactly once; stages >= 1 touch only the few
// percent of rows whose butterflies span tile boundaries.
//
// Arithmetic (fp64 mode): a*x0 + b*x1 and (-b)*x0 + a*x1 with separately rounded products and sum
// (__dmul_rn / __dadd_rn never contract to FMA) -- bit-exact with torch on CPU (RAHT.py:331-332).
#include "plan.cuh"
#include <cstdlib>

namespace {

constexpr int TR_THREADS = 256;
constexpr int TR_WARPS = TR_THREADS / 32;
constexpr unsigned TMA_CHUNK = 32768;     // bytes per bulk copy

template <typename T> struct Arith;
template <> struct Arith<double> {
    static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
    static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
    static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
};
template <> struct Arith<float> {
    static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
    static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
    static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
};

// Quantisation fused into stage 0 (raht_forward_quant / raht_dequant_inverse): the rows a stage-0 tile finishes -- all
// but the ~15 % it defers to the upper stages -- never exist as fp64 coefficients in HBM.
//   QM_FWD: instead of the write-back, finished rows are quantised and stored at their order_RAGFT position in the
//           channel-major int32 matrix; deferred rows go to T as before (the upper stages work on T, and
//           quant_deferred_kernel quantises them afterwards);
//   QM_INV: instead of the load from T, finished rows are gathered from the int32 matrix and dequantised; deferred
//           rows come from C, where the upper stages left them (dequant_deferred_kernel fills T for those stages).
enum { QM_NONE = 0, QM_FWD = 1, QM_INV = 2 };
struct QuantArgs {
    const int32_t *pos;        // [n] position of row r in the coded order (inverse permutation of order_RAGFT)
    const double *steps;       // [D]
    int32_t *q;                // (D, ld) channel-major
    int64_t ld;
    int32_t *status;
};
__device__ __forceinline__ int32_t quantise_one(double v, double step, bool &ovf)
{
    const double qq = floor(__dadd_rn(__ddiv_rn(v, step), 0.5));            // encode_3dgs.py:204
    ovf |= !(qq >= -2147483648.0 && qq <= 2147483647.0);
    return (int32_t)qq;                                                     // :215 .to(torch.int32)
}

template <typename Tel> struct SmemLayout {
    size_t off_ab, off_sched, off_hdr, off_flag, off_bar, off_pos, off_step, off_def, total;
    __host__ __device__ SmemLayout(int R, int D)
    {
        size_t o = (size_t)R * D * sizeof(Tel);
        o = (o + 15) & ~(size_t)15;
        off_ab = o; o += (size_t)R * 2 * sizeof(Tel);
        off_sched = o; o += (size_t)R * 4;
        off_hdr = o; o += sizeof(TileHdr);
        off_bar = o; o += 8;
        off_flag = o; o += (size_t)R;
        o = (o + 7) & ~(size_t)7;
        off_step = o; o += (size_t)D * 8;
        off_pos = o; o += (size_t)R * 4;
        off_def = o; o += (size_t)PLAN_DEFER_MAX * 2 + 2;
        total = (o + 15) & ~(size_t)15;
    }
};

// ---- hand-over between the tiles of one launch (chained kernels below) ------------------------------------------------
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(unsigned *p, unsigned v)
{
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
template <typename T> __device__ __forceinline__ T ld_cg(const T *p) { return __ldcg(p); }

// CHAIN (stage0_tile / upper_tile): 0 = a launch per stage (the tile's inputs were written by earlier launches);
// 1 = forward chain, 2 = inverse chain: rows another CTA of the SAME launch wrote are read from L2 (ld.cg) behind an
// acquire of that CTA's hand-over, and this tile's own output is fenced before its hand-over.
enum { CHAIN_NONE = 0, CHAIN_FWD = 1, CHAIN_INV = 2 };

// ---- stage 0: tiles of consecutive rows ------------------------------------------------------------------
// A: source rows (row-major, Tin).  B: second source for the rows this tile deferred (inverse only: the
// upper stages have already rewritten them in the output), may be nullptr.  O: destination.
// FWD: forward butterflies in schedule order, else inverse butterflies in reverse order.
template <typename Tin, typename Tel, int DT, bool FWD, int QM, int CHAIN>
__device__ __forceinline__ void stage0_tile(const PlanDev &P, int64_t tile, const Tin *__restrict__ A, const Tel *B, Tel *O, int D_rt,
                                            int use_tma, const QuantArgs &Q, unsigned char *tr_smem, unsigned epoch)
{
    const int D = DT ? DT : D_rt;
    const int R = P.tile_rows;
    const SmemLayout<Tel> L(R, D);
    Tel *s_rows = (Tel *)tr_smem;
    Tel *s_ab = (Tel *)(tr_smem + L.off_ab);
    uint32_t *s_sched = (uint32_t *)(tr_smem + L.off_sched);
    TileHdr *s_hdr = (TileHdr *)(tr_smem + L.off_hdr);
    uint64_t *s_bar = (uint64_t *)(tr_smem + L.off_bar);
    uint8_t *s_flag = tr_smem + L.off_flag;
    double *s_step = (double *)(tr_smem + L.off_step);
    int32_t *s_pos = (int32_t *)(tr_smem + L.off_pos);
    uint16_t *s_def = (uint16_t *)(tr_smem + L.off_def);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t base = tile * R;

    if (tid < (int)(sizeof(TileHdr) / 4)) ((uint32_t *)s_hdr)[tid] = ((const uint32_t *)(P.hdr + tile))[tid];
    const bool patch = (B != nullptr) || QM == QM_FWD;       // (QM_FWD: the flags pick the rows that go to T)
    // The tile's deferred list sits at a fixed place of the plan: it is requested together with the header (at most
    // PLAN_DEFER_MAX entries, the count arrives with the header), so that the rows it names can be requested right behind
    // the header barrier -- with the schedule, not a round trip after it.
    static_assert(PLAN_DEFER_MAX <= TR_THREADS, "one deferred entry per thread");
    uint16_t dl_pre = 0;
    if (patch) {
        if (tid < PLAN_DEFER_MAX) dl_pre = P.def_local[(size_t)tile * PLAN_DEFER_MAX + tid];
        for (int j = tid; j < R; j += TR_THREADS) s_flag[j] = 0;
    }
    __syncthreads();
    const int n_rows = s_hdr->n_rows;
    const int n_bf = s_hdr->n_bf;
    const int n_rounds = s_hdr->n_rounds;
    const size_t bf_off = s_hdr->bf_off;
    const size_t tile_bytes = (size_t)n_rows * D * sizeof(Tel);

    // bulk copies need 16-byte granularity; tiles that do not qualify (ragged last tile) use plain loads
    bool tma_in = false, tma_out = false;
    if (use_tma && (tile_bytes % 16) == 0) {
        if constexpr (sizeof(Tin) == sizeof(Tel)) tma_in = QM != QM_INV && (((uintptr_t)(A + base * D)) % 16) == 0;
        tma_out = QM != QM_FWD && (((uintptr_t)(O + base * D)) % 16) == 0;
    }
    // a programmatic dependent may read T ahead of its wait only through the bulk copy (L2); plain loads wait first
    if constexpr (CHAIN == CHAIN_NONE && !FWD) { if (!tma_in) pdl_wait(); }
    if constexpr (QM != QM_NONE) {
        for (int c = tid; c < D; c += TR_THREADS) s_step[c] = Q.steps[c];
        for (int j = tid; j < n_rows; j += TR_THREADS) s_pos[j] = Q.pos[base + j];
    }
    if (tma_in) {
        if (tid == 0) {
            mbar_init(s_bar, 1);
            fence_proxy_async();
            mbar_expect_tx(s_bar, (unsigned)tile_bytes);
            const char *src = (const char *)(A + base * D);
            char *dst = (char *)s_rows;
            for (size_t off = 0; off < tile_bytes; off += TMA_CHUNK) {
                unsigned sz = (unsigned)((tile_bytes - off) < TMA_CHUNK ? (tile_bytes - off) : TMA_CHUNK);
                bulk_g2s(dst + off, src + off, sz, s_bar);
            }
        }
    }

    if (patch && tid < min((int)s_hdr->n_def, PLAN_DEFER_MAX)) {
        s_flag[dl_pre] = 1;
        s_def[tid] = dl_pre;
    }
    // the deferred list for the row requests below, and thread 0's mbarrier initialisation for everybody's wait on it
    // (nothing is in flight but the bulk copy: a short barrier)
    if (patch || tma_in) __syncthreads();

    // schedule and coefficients: a thread's first entry is requested here, into registers (plan data: part of the prologue
    // of a programmatic dependent launch), and stored behind the row requests below, so that neither waits for the other
    uint32_t sched_r = 0;
    double2 ab_r = make_double2(0.0, 0.0);
    if (tid < n_bf) {
        sched_r = P.sched[bf_off + tid];
        ab_r = P.ab[bf_off + tid];
    }

    // rows this tile deferred come from B: the first PF_ITEMS elements per thread are requested BEFORE the schedule is
    // stored and before the wait for the bulk copy (dependent rounds of loads behind either cost ~1 us per tile each), the
    // rest (tiles that defer more than ~34 rows) afterwards
    constexpr int PF_ITEMS = 8;
    Tel pf[PF_ITEMS];
    const int n_patch = (B != nullptr) ? (int)s_hdr->n_def * D : 0;
    if constexpr (CHAIN == CHAIN_INV) {
        // the rows this tile deferred are written by its parent tile in THIS launch: wait for its hand-over (the bulk
        // load of the tile and the schedule loads above are already in flight)
        if (B != nullptr && P.n_stages > 1) {
            if (tid == 0) {
                const unsigned *flag = P.done + P.tile_base[1] + tile / PLAN_FANIN;
                while (ld_acquire_u32(flag) != epoch) __nanosleep(64);
            }
            __syncthreads();
        }
    }
    // (launched as a programmatic dependent of the last upper stage: header, schedule and the bulk load of T above ran under
    // that kernel's tail; B and O are what it writes)
    if constexpr (CHAIN == CHAIN_NONE) pdl_wait();
    if (n_patch) {
#pragma unroll
        for (int u = 0; u < PF_ITEMS; ++u) {
            const int idx = tid + u * TR_THREADS;
            if (idx < n_patch) {
                const int j = idx / D, c = idx - j * D;
                const Tel *src = B + (size_t)(base + s_def[j]) * D + c;
                pf[u] = CHAIN == CHAIN_INV ? ld_cg(src) : *src;
            }
        }
    }

    if (tma_in) {
        mbar_wait(s_bar, 0);
    } else if constexpr (QM == QM_INV) {
        // a warp per channel: the positions of consecutive rows of one level bucket are consecutive
        for (int c = warp; c < D; c += TR_WARPS) {
            const int32_t *qc = Q.q + (size_t)c * Q.ld;
            const double st = s_step[c];
            for (int j = lane; j < n_rows; j += 32)
                if (!(B != nullptr && s_flag[j])) s_rows[j * D + c] = (Tel)__dmul_rn((double)qc[s_pos[j]], st);   // encode_3dgs.py:261
        }
        __syncthreads();
    } else {
        const Tin *src = A + base * D;
        const int n_el = n_rows * D;
        bool done = false;
        if constexpr (sizeof(Tin) == 4 && sizeof(Tel) == 8) {
            // fp32 rows widened on the way in: 16-byte requests, LD_BATCH of them in flight per thread before the first
            // use (a scalar loop kept ~4 KB per CTA in flight and was bound by the round trip, not by the bandwidth)
            if ((n_el & 3) == 0 && (((uintptr_t)src) & 15) == 0) {
                constexpr int LD_BATCH = 8;
                const float4 *src4 = (const float4 *)src;
                const int n4 = n_el >> 2;
                for (int i0 = tid; i0 < n4; i0 += LD_BATCH * TR_THREADS) {
                    float4 v[LD_BATCH];
#pragma unroll
                    for (int u = 0; u < LD_BATCH; ++u)
                        if (i0 + u * TR_THREADS < n4) v[u] = __ldg(src4 + i0 + u * TR_THREADS);
#pragma unroll
                    for (int u = 0; u < LD_BATCH; ++u)
                        if (i0 + u * TR_THREADS < n4) {
                            double2 *dst = (double2 *)(s_rows + 4 * (size_t)(i0 + u * TR_THREADS));
                            dst[0] = make_double2((double)v[u].x, (double)v[u].y);
                            dst[1] = make_double2((double)v[u].z, (double)v[u].w);
                        }
                }
                done = true;
            }
        }
        if (!done)
            for (int i = tid; i < n_el; i += TR_THREADS) s_rows[i] = (Tel)src[i];
        __syncthreads();
    }
    if (n_patch) {
#pragma unroll
        for (int u = 0; u < PF_ITEMS; ++u) {
            const int idx = tid + u * TR_THREADS;
            if (idx < n_patch) {
                const int j = idx / D, c = idx - j * D;
                s_rows[(int)s_def[j] * D + c] = pf[u];
            }
        }
        for (int idx = tid + PF_ITEMS * TR_THREADS; idx < n_patch; idx += TR_THREADS) {
            const int j = idx / D, c = idx - j * D;
            const Tel *src = B + (size_t)(base + s_def[j]) * D + c;
            s_rows[(int)s_def[j] * D + c] = CHAIN == CHAIN_INV ? ld_cg(src) : *src;
        }
    }
    if (tid < n_bf) {                                        // (requested behind the header barrier, with the rows)
        s_sched[tid] = sched_r;
        s_ab[2 * tid] = (Tel)ab_r.x;
        s_ab[2 * tid + 1] = (Tel)ab_r.y;
    }
    for (int j = tid + TR_THREADS; j < n_bf; j += TR_THREADS) {        // (tiles of more than TR_THREADS rows)
        s_sched[j] = P.sched[bf_off + j];
        double2 c = P.ab[bf_off + j];
        s_ab[2 * j] = (Tel)c.x;
        s_ab[2 * j + 1] = (Tel)c.y;
    }
    __syncthreads();

    // butterflies, one dependency round at a time
    using AR = Arith<Tel>;
    const int BF_CW = D <= 4 ? 4 : D <= 8 ? 8 : D <= 16 ? 16 : D <= 32 ? 32 : 64;      // channel slots per butterfly (power of two)
    for (int kk = 0; kk < n_rounds; ++kk) {
        const int k = FWD ? kk : n_rounds - 1 - kk;
        const int lo = k ? s_hdr->round_end[k - 1] : 0;
        const int hi = s_hdr->round_end[k];
        // A thread keeps its channel and walks the round's butterflies: the lanes of a warp are consecutive channels of ONE
        // butterfly.  (A linear index over (butterfly, channel) made almost every warp straddle two butterflies -- rows of
        // 59 doubles -- whose rows collide in the banks, and cost a division per element: rounds 11.3 -> 10.0 k clocks per
        // tile forward, 10.8 -> 9.1 k inverse.  The rounds are bound by shared-memory wavefronts)
        for (int cb0 = 0; cb0 < D; cb0 += BF_CW) {
            const int c = cb0 + (tid & (BF_CW - 1));
            if (c < D)
                for (int e = lo + tid / BF_CW; e < hi; e += TR_THREADS / BF_CW) {
                    const uint32_t ent = s_sched[e];
                    const int r = ent & 4095, s = (ent >> 12) & 4095;
                    const Tel ca = s_ab[2 * e], cb = s_ab[2 * e + 1];
                    Tel *p0 = s_rows + s * D + c, *p1 = s_rows + r * D + c;
                    const Tel x0 = *p0, x1 = *p1;
                    if (FWD) {
                        *p0 = AR::add(AR::mul(ca, x0), AR::mul(cb, x1));        // RAHT.py:331  low-pass, survives
                        *p1 = AR::add(AR::mul(-cb, x0), AR::mul(ca, x1));       // RAHT.py:332  high-pass, final
                    } else {
                        *p0 = AR::sub(AR::mul(ca, x0), AR::mul(cb, x1));        // iRAHT.py:108
                        *p1 = AR::add(AR::mul(cb, x0), AR::mul(ca, x1));        // iRAHT.py:109
                    }
                }
        }
        __syncthreads();
    }

    // write back
    if constexpr (QM == QM_FWD) {
        bool ovf = false;
        for (int c = warp; c < D; c += TR_WARPS) {
            int32_t *qc = Q.q + (size_t)c * Q.ld;
            const double st = s_step[c];
            for (int j = lane; j < n_rows; j += 32)
                if (!s_flag[j]) qc[s_pos[j]] = quantise_one((double)s_rows[j * D + c], st, ovf);
        }
        if (__any_sync(~0u, ovf) && lane == 0 && Q.status) atomicOr(Q.status, RAHT_ST_INT32_OVERFLOW);
        for (int j = warp; j < n_rows; j += TR_WARPS)
            if (s_flag[j]) {
                Tel *dst = O + (size_t)(base + j) * D;
                for (int c = lane; c < D; c += 32) dst[c] = s_rows[j * D + c];
            }
    } else if (tma_out) {
        fence_proxy_async();
        __syncthreads();
        if (tid == 0) {
            char *dst = (char *)(O + base * D);
            const char *src = (const char *)s_rows;
            for (size_t off = 0; off < tile_bytes; off += TMA_CHUNK) {
                unsigned sz = (unsigned)((tile_bytes - off) < TMA_CHUNK ? (tile_bytes - off) : TMA_CHUNK);
                bulk_s2g(dst + off, src + off, sz);
            }
            bulk_commit();
            if constexpr (CHAIN == CHAIN_NONE) bulk_wait_read0();
            else { bulk_wait0(); fence_proxy_async_all(); }   // the stores must have landed before the hand-over
        }
    } else {
        Tel *dst = O + base * D;
        const int n_el = n_rows * D;
        for (int i = tid; i < n_el; i += TR_THREADS) dst[i] = s_rows[i];
    }
}

template <typename Tin, typename Tel, int DT, bool FWD, int QM = QM_NONE>
__global__ void __launch_bounds__(TR_THREADS)
transform_stage0_kernel(PlanDev P, const Tin *__restrict__ A, const Tel *B, Tel *O, int D_rt, int use_tma, QuantArgs Q)
{
    extern __shared__ __align__(128) unsigned char tr_smem[];
    pdl_trigger();                                           // the next stage's tiles may run their prologue under this kernel's tail
    stage0_tile<Tin, Tel, DT, FWD, QM, CHAIN_NONE>(P, (int64_t)blockIdx.x, A, B, O, D_rt, use_tma, Q, tr_smem, 0u);
}

// ---- stages >= 1: tiles of gathered rows -------------------------------------------------------------------
// The entries of tile j are the rows its (up to) 8 child tiles deferred, at most 1032.  The channels are
// processed in passes of DS = min(D, UP_ELEMS / n_rows) so that any tile fits the same shared-memory
// budget; a typical tile (100-300 rows) takes one pass.  A: source of a row unless this tile deferred
// it and B is given (inverse: then B, which the upper stage already wrote).  O: destination.
// The butterflies of different channels are independent, so a stage with few tiles (the top of the hierarchy: 16, 2,
// 1 tiles, each a serial chain of up to 64 rounds) is spread over gridDim.y groups of channels: a lone CTA per tile
// took 40 us there whatever the tile held.
constexpr int UP_ELEMS = 8192;     // row elements staged per pass
constexpr int UP_GROUPS_MAX = 16;
constexpr int UP_GROUPS_MANY = 2, UP_ELEMS_MANY = 2560;      // stages with >= 2 tiles per SM (sweep: scripts/gpu_upper_sweep.sh)

template <typename Tel, bool FWD, int CHAIN>
__device__ __forceinline__ void upper_tile(const PlanDev &P, int stage, int64_t tile, const Tel *A, const Tel *B, Tel *O, int D,
                                           int up_elems, int cb, int ce, unsigned char *tr_smem)
{
    Tel *s_rows = (Tel *)tr_smem;                                          // up_elems
    Tel *s_ab = s_rows + up_elems;                                         // 2 * PLAN_UP_CAP
    uint32_t *s_sched = (uint32_t *)(s_ab + 2 * PLAN_UP_CAP);              // PLAN_UP_CAP
    uint32_t *s_rowid = s_sched + PLAN_UP_CAP;                             // PLAN_UP_CAP
    uint8_t *s_flag = (uint8_t *)(s_rowid + PLAN_UP_CAP);                  // PLAN_UP_CAP
    __shared__ TileHdr s_hdr;
    __shared__ int s_off[PLAN_FANIN + 1];

    const int tid = threadIdx.x;
    const int64_t gt = P.tile_base[stage] + tile;
    if (tid < (int)(sizeof(TileHdr) / 4)) ((uint32_t *)&s_hdr)[tid] = ((const uint32_t *)(P.hdr + gt))[tid];
    for (int j = tid; j < PLAN_UP_CAP; j += TR_THREADS) s_flag[j] = 0;   // (ahead of the header barrier: the deferred list sets them)
    const int64_t c0 = tile * PLAN_FANIN;
    const int64_t cbase = P.tile_base[stage - 1];
    if (tid < 32) {                                          // entries per child tile: eight independent loads, one scan
        int cnt = 0;
        if (tid < PLAN_FANIN && c0 + tid < P.tiles[stage - 1]) cnt = min((int)P.hdr[cbase + c0 + tid].n_def, PLAN_DEFER_MAX);
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < PLAN_FANIN; o <<= 1) {
            const int v = __shfl_up_sync(~0u, incl, o);
            if (tid >= o) incl += v;
        }
        if (tid < PLAN_FANIN) s_off[tid] = incl - cnt;
        if (tid == PLAN_FANIN - 1) s_off[PLAN_FANIN] = incl;
    }
    __syncthreads();
    const int n_rows = s_hdr.n_rows;
    const int n_bf = s_hdr.n_bf;
    const int n_rounds = s_hdr.n_rounds;
    const size_t bf_off = s_hdr.bf_off;
    // nothing scheduled: forward, every row stays as it is; the inverse still has to carry the tile's rows from T (or from
    // the parent's output) to C -- a forest root that became final in this tile is read from C by the child that deferred it
    // (a CTA that leaves early still waits: a dependent grid whose CTAs all returned at once would complete before the
    // stage ahead of it, and the stage behind it -- which waits for THIS grid only -- would read rows that are not written yet)
    if (n_bf == 0 && FWD) {
        if constexpr (CHAIN == CHAIN_NONE) pdl_wait();
        return;
    }
    // the deferred lists of the children, one flat pass with every load of a thread in flight at once (a loop per child
    // made eight dependent round trips of it: most of the ~10 us an upper tile takes whatever it holds)
    {
        constexpr int RB = (PLAN_UP_CAP + TR_THREADS - 1) / TR_THREADS;
        uint32_t rid[RB];
#pragma unroll
        for (int u = 0; u < RB; ++u) {
            const int j = tid + u * TR_THREADS;
            if (j < n_rows) {
                int c = 0;
#pragma unroll
                for (int k = 1; k < PLAN_FANIN; ++k) c += j >= s_off[k] ? 1 : 0;
                rid[u] = P.def_rows[(size_t)(cbase + c0 + c) * PLAN_DEFER_MAX + (j - s_off[c])];
            }
        }
#pragma unroll
        for (int u = 0; u < RB; ++u)
            if (tid + u * TR_THREADS < n_rows) s_rowid[tid + u * TR_THREADS] = rid[u];
    }
    for (int j = tid; j < n_bf; j += TR_THREADS) {
        s_sched[j] = P.sched[bf_off + j];
        double2 c = P.ab[bf_off + j];
        s_ab[2 * j] = (Tel)c.x;
        s_ab[2 * j + 1] = (Tel)c.y;
    }
    if (B) {                                                 // same round trip as the schedule (the flags were zeroed above)
        const int n_def = s_hdr.n_def;
        for (int j = tid; j < n_def; j += TR_THREADS) s_flag[P.def_local[(size_t)gt * PLAN_DEFER_MAX + j]] = 1;
    }
    __syncthreads();

    using AR = Arith<Tel>;
    // everything above read the plan only; the rows are what the stage before wrote (programmatic dependent launch)
    if constexpr (CHAIN == CHAIN_NONE) pdl_wait();
    if (cb >= ce) return;
    int DS = up_elems / n_rows;
    if (DS > ce - cb) DS = ce - cb;
    for (int d0 = cb; d0 < ce; d0 += DS) {
        const int dn = min(DS, ce - d0);
        const int n_el = n_rows * dn;
        // gather, UP_BATCH requests in flight per thread (one at a time, the stage was bound by the round trip of each)
        constexpr int UP_BATCH = 8;
        for (int i0 = tid; i0 < n_el; i0 += UP_BATCH * TR_THREADS) {
            Tel v[UP_BATCH];
#pragma unroll
            for (int u = 0; u < UP_BATCH; ++u) {
                const int i = i0 + u * TR_THREADS;
                if (i < n_el) {
                    const int j = i / dn, c = i - j * dn;
                    const Tel *src = ((B && s_flag[j]) ? B : A) + (size_t)s_rowid[j] * D + d0 + c;
                    // forward chain: every gathered row was written in this launch; inverse chain: those from B (the parent's)
                    v[u] = (CHAIN == CHAIN_FWD || (CHAIN == CHAIN_INV && B && s_flag[j])) ? ld_cg(src) : *src;
                }
            }
#pragma unroll
            for (int u = 0; u < UP_BATCH; ++u)
                if (i0 + u * TR_THREADS < n_el) s_rows[i0 + u * TR_THREADS] = v[u];      // = s_rows[j * dn + c]
        }
        __syncthreads();
        for (int kk = 0; kk < n_rounds; ++kk) {
            const int k = FWD ? kk : n_rounds - 1 - kk;
            const int lo = k ? s_hdr.round_end[k - 1] : 0;
            const int hi = s_hdr.round_end[k];
            const int work = (hi - lo) * dn;
            for (int idx = tid; idx < work; idx += TR_THREADS) {
                const int e = lo + idx / dn;
                const int c = idx - (idx / dn) * dn;
                const uint32_t ent = s_sched[e];
                const int r = ent & 4095, s = (ent >> 12) & 4095;
                const Tel ca = s_ab[2 * e], cbv = s_ab[2 * e + 1];
                Tel *p0 = s_rows + s * dn + c, *p1 = s_rows + r * dn + c;
                const Tel x0 = *p0, x1 = *p1;
                if (FWD) {
                    *p0 = AR::add(AR::mul(ca, x0), AR::mul(cbv, x1));
                    *p1 = AR::add(AR::mul(-cbv, x0), AR::mul(ca, x1));
                } else {
                    *p0 = AR::sub(AR::mul(ca, x0), AR::mul(cbv, x1));
                    *p1 = AR::add(AR::mul(cbv, x0), AR::mul(ca, x1));
                }
            }
            __syncthreads();
        }
        for (int i = tid; i < n_el; i += TR_THREADS) {
            const int j = i / dn, c = i - j * dn;
            O[(size_t)s_rowid[j] * D + d0 + c] = s_rows[j * dn + c];
        }
        __syncthreads();
    }
}

template <typename Tel, bool FWD>
__global__ void __launch_bounds__(TR_THREADS)
transform_upper_kernel(PlanDev P, int stage, const Tel *A, const Tel *B, Tel *O, int D, int up_elems)
{
    extern __shared__ __align__(128) unsigned char tr_smem[];
    pdl_trigger();
    const int per = (D + (int)gridDim.y - 1) / (int)gridDim.y;            // channels of this CTA's group
    const int cb = (int)blockIdx.y * per, ce = min(D, cb + per);
    upper_tile<Tel, FWD, CHAIN_NONE>(P, stage, (int64_t)blockIdx.x, A, B, O, D, up_elems, cb, ce, tr_smem);
}

// ---- the whole transform in ONE launch -------------------------------------------------------------------------------
// Round 1 ran a launch per stage: at 1 M rows the five upper stages (977 / 123 / 16 / 2 / 1 tiles, 7 % of the rows) cost
// 5 x 27 us of dependent launches next to 0.2 ms of stage 0.
// Forward (bottom-up): a CTA runs its stage-0 tile, then signs in at the parent tile's counter; the CTA that signs in
// last runs the parent (its children's rows are fenced to L2), and so on to the top -- nobody waits, the upper stages
// run in the shadow of the stage-0 stream, and only the last few tiles of the chain are exposed.
// Inverse (top-down): the tiles are handed out by TICKET in topological order (top tile, ..., stage-1 tiles, stage-0
// tiles); a tile spins on its parent's done-flag, which a smaller ticket -- a CTA that is already running -- sets:
// forward progress needs no co-residency.  A stage-0 tile has its bulk load in flight while it waits.
// arrive[] / ticket are left at zero for the next launch; done[] is compared with a per-launch epoch.
constexpr int UP_FIXED_BYTES = PLAN_UP_CAP * 9 + 16;                       // s_sched, s_rowid, s_flag
template <typename Tel> __host__ __device__ inline size_t upper_fixed_bytes() { return (size_t)2 * PLAN_UP_CAP * sizeof(Tel) + UP_FIXED_BYTES; }

template <typename Tin, typename Tel, int DT>
__global__ void __launch_bounds__(TR_THREADS)
transform_forward_chain_kernel(PlanDev P, const Tin *__restrict__ A, Tel *O, int D_rt, int use_tma, int up_elems)
{
    extern __shared__ __align__(128) unsigned char tr_smem[];
    __shared__ int s_last;
    const int D = DT ? DT : D_rt;
    stage0_tile<Tin, Tel, DT, true, QM_NONE, CHAIN_FWD>(P, (int64_t)blockIdx.x, A, (const Tel *)nullptr, O, D_rt, use_tma, QuantArgs(), tr_smem, 0u);
    int stage = 0;
    int64_t tile = blockIdx.x;
    while (stage + 1 < P.n_stages) {
        const int64_t parent = tile / PLAN_FANIN;
        const int64_t children = min((int64_t)PLAN_FANIN, P.tiles[stage] - parent * PLAN_FANIN);
        __threadfence();                                   // this tile's rows, before the sign-in
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned *ctr = P.arrive + P.tile_base[stage + 1] + parent;
            const unsigned prev = atomicAdd(ctr, 1u);
            s_last = prev + 1 == (unsigned)children;
            if (s_last) *ctr = 0u;                         // ready for the next launch
            __threadfence();
        }
        __syncthreads();
        if (!s_last) break;
        ++stage;
        tile = parent;
        upper_tile<Tel, true, CHAIN_FWD>(P, stage, tile, (const Tel *)O, (const Tel *)nullptr, O, D, up_elems, 0, D, tr_smem);
    }
}

template <typename Tel, int DT>
__global__ void __launch_bounds__(TR_THREADS)
transform_inverse_chain_kernel(PlanDev P, const Tel *__restrict__ T, Tel *C, int D_rt, int use_tma, int up_elems, unsigned epoch)
{
    extern __shared__ __align__(128) unsigned char tr_smem[];
    __shared__ unsigned s_ticket;
    const int D = DT ? DT : D_rt;
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(P.ticket, 1u);
        if (t + 1 == (unsigned)P.tiles_total) *P.ticket = 0u;             // every ticket is out: ready for the next launch
        s_ticket = t;
    }
    __syncthreads();
    // ticket -> (stage, tile): the top stage first
    int64_t rest = s_ticket;
    int stage = P.n_stages - 1;
    while (stage > 0 && rest >= P.tiles[stage]) { rest -= P.tiles[stage]; --stage; }
    const int64_t tile = rest;
    if (stage == 0) {
        stage0_tile<Tel, Tel, DT, false, QM_NONE, CHAIN_INV>(P, tile, T, P.n_stages > 1 ? (const Tel *)C : (const Tel *)nullptr, C, D_rt, use_tma,
                                                             QuantArgs(), tr_smem, epoch);
        return;
    }
    const bool top = stage == P.n_stages - 1;
    if (!top) {
        if (threadIdx.x == 0) {
            const unsigned *flag = P.done + P.tile_base[stage + 1] + tile / PLAN_FANIN;
            while (ld_acquire_u32(flag) != epoch) __nanosleep(64);
        }
        __syncthreads();
    }
    upper_tile<Tel, false, CHAIN_INV>(P, stage, tile, T, top ? (const Tel *)nullptr : (const Tel *)C, C, D, up_elems, 0, D, tr_smem);
    __threadfence();                                       // this tile's rows, before the hand-over to its children
    __syncthreads();
    if (threadIdx.x == 0) st_release_u32(P.done + P.tile_base[stage] + tile, epoch);
}

// The upper stages alone as one launch (stage 0 keeps its own launch and exits without waiting for its bulk store to land):
// forward: a CTA per stage-1 tile, last-arriver chaining from there up; inverse: tickets over the upper tiles, top first.
template <typename Tel>
__global__ void __launch_bounds__(TR_THREADS) transform_upper_forward_chain_kernel(PlanDev P, Tel *O, int D, int up_elems)
{
    extern __shared__ __align__(128) unsigned char tr_smem[];
    __shared__ int s_last;
    int stage = 1;
    int64_t tile = blockIdx.x;
    upper_tile<Tel, true, CHAIN_NONE>(P, stage, tile, (const Tel *)O, (const Tel *)nullptr, O, D, up_elems, 0, D, tr_smem);
    while (stage + 1 < P.n_stages) {
        const int64_t parent = tile / PLAN_FANIN;
        const int64_t children = min((int64_t)PLAN_FANIN, P.tiles[stage] - parent * PLAN_FANIN);
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned *ctr = P.arrive + P.tile_base[stage + 1] + parent;
            const unsigned prev = atomicAdd(ctr, 1u);
            s_last = prev + 1 == (unsigned)children;
            if (s_last) *ctr = 0u;
            __threadfence();
        }
        __syncthreads();
        if (!s_last) break;
        ++stage;
        tile = parent;
        upper_tile<Tel, true, CHAIN_FWD>(P, stage, tile, (const Tel *)O, (const Tel *)nullptr, O, D, up_elems, 0, D, tr_smem);
    }
}

template <typename Tel>
__global__ void __launch_bounds__(TR_THREADS) transform_upper_inverse_chain_kernel(PlanDev P, const Tel *__restrict__ T, Tel *C, int D,
                                                                                   int up_elems, unsigned epoch)
{
    extern __shared__ __align__(128) unsigned char tr_smem[];
    __shared__ unsigned s_ticket;
    const unsigned total = (unsigned)(P.tiles_total - P.tiles[0]);
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(P.ticket, 1u);
        if (t + 1 == total) *P.ticket = 0u;
        s_ticket = t;
    }
    __syncthreads();
    int64_t rest = s_ticket;
    int stage = P.n_stages - 1;
    while (stage > 1 && rest >= P.tiles[stage]) { rest -= P.tiles[stage]; --stage; }
    const int64_t tile = rest;
    const bool top = stage == P.n_stages - 1;
    if (!top) {
        if (threadIdx.x == 0) {
            const unsigned *flag = P.done + P.tile_base[stage + 1] + tile / PLAN_FANIN;
            while (ld_acquire_u32(flag) != epoch) __nanosleep(64);
        }
        __syncthreads();
    }
    upper_tile<Tel, false, CHAIN_INV>(P, stage, tile, T, top ? (const Tel *)nullptr : (const Tel *)C, C, D, up_elems, 0, D, tr_smem);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) st_release_u32(P.done + P.tile_base[stage] + tile, epoch);
}

// rows the stage-0 tiles deferred, a warp per row: after the upper stages of the forward transform (T -> q), or before
// those of the inverse (q -> T)
template <typename Tel, bool TO_Q>
__global__ void __launch_bounds__(TR_THREADS) deferred_rows_kernel(PlanDev P, Tel *T, int D, QuantArgs Q)
{
    const int64_t tile = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n_def = min((int)P.hdr[tile].n_def, PLAN_DEFER_MAX);
    bool ovf = false;
    for (int j = warp; j < n_def; j += TR_WARPS) {
        const size_t row = P.def_rows[(size_t)tile * PLAN_DEFER_MAX + j];
        const size_t at = (size_t)Q.pos[row];
        Tel *t = T + row * D;
        for (int c = lane; c < D; c += 32) {
            if (TO_Q) Q.q[(size_t)c * Q.ld + at] = quantise_one((double)t[c], Q.steps[c], ovf);
            else t[c] = (Tel)__dmul_rn((double)Q.q[(size_t)c * Q.ld + at], Q.steps[c]);
        }
    }
    if (TO_Q && __any_sync(~0u, ovf) && lane == 0 && Q.status) atomicOr(Q.status, RAHT_ST_INT32_OVERFLOW);
}

template <typename Tin, typename Tel, int DT, bool FWD, int QM>
int launch_stage0_q(const raht_plan *P, const Tin *A, const Tel *B, Tel *O, int D, int use_tma, const QuantArgs &Q, cudaStream_t st)
{
    const int R = P->d.tile_rows;
    SmemLayout<Tel> L(R, D);
    auto kern = transform_stage0_kernel<Tin, Tel, DT, FWD, QM>;
    if (L.total > 227 * 1024) {
        raht_set_error("transform: tile of %d rows x %d channels needs %zu bytes of shared memory (> 227 KB)", R, D, L.total);
        return RAHT_E_INVALID;
    }
    RAHT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total));
    (raht_count_launches(1), kern)<<<(unsigned)P->d.tiles[0], TR_THREADS, L.total, st>>>(P->d, A, B, O, D, use_tma, Q);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}
template <typename Tin, typename Tel, bool FWD, int QM>
int launch_stage0_qd(const raht_plan *P, const Tin *A, const Tel *B, Tel *O, int D, int use_tma, const QuantArgs &Q, cudaStream_t st)
{
    switch (D) {
    case 59: return launch_stage0_q<Tin, Tel, 59, FWD, QM>(P, A, B, O, D, use_tma, Q, st);
    default: return launch_stage0_q<Tin, Tel, 0, FWD, QM>(P, A, B, O, D, use_tma, Q, st);
    }
}

template <typename Tin, typename Tel, int DT, bool FWD>
int launch_stage0(const raht_plan *P, const Tin *A, const Tel *B, Tel *O, int D, int use_tma, cudaStream_t st)
{
    const int R = P->d.tile_rows;
    SmemLayout<Tel> L(R, D);
    auto kern = transform_stage0_kernel<Tin, Tel, DT, FWD>;
    if (L.total > 227 * 1024) {
        raht_set_error("transform: tile of %d rows x %d channels needs %zu bytes of shared memory (> 227 KB)", R, D, L.total);
        return RAHT_E_INVALID;
    }
    RAHT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total));
    raht_count_launches(1);
    const bool use_pdl = raht_use_pdl(PDL_TRANSFORM);
    // the inverse's stage 0 follows the upper stages: as their programmatic dependent its tiles load header, schedule and T
    // rows before the last upper stage has finished (B != nullptr marks that launch)
    if (use_pdl && !FWD && B != nullptr)
        RAHT_CUDA(pdl_launch(kern, dim3((unsigned)P->d.tiles[0]), dim3(TR_THREADS), L.total, st, P->d, A, B, O, D, use_tma, QuantArgs()));
    else kern<<<(unsigned)P->d.tiles[0], TR_THREADS, L.total, st>>>(P->d, A, B, O, D, use_tma, QuantArgs());
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

template <typename Tin, typename Tel, bool FWD>
int launch_stage0_d(const raht_plan *P, const Tin *A, const Tel *B, Tel *O, int D, int use_tma, cudaStream_t st)
{
    switch (D) {
    case 3: return launch_stage0<Tin, Tel, 3, FWD>(P, A, B, O, D, use_tma, st);
    case 56: return launch_stage0<Tin, Tel, 56, FWD>(P, A, B, O, D, use_tma, st);
    case 59: return launch_stage0<Tin, Tel, 59, FWD>(P, A, B, O, D, use_tma, st);
    default: return launch_stage0<Tin, Tel, 0, FWD>(P, A, B, O, D, use_tma, st);
    }
}

// head: the first kernel of a raht_inverse call.  It is a PLAIN launch: the later kernels of the call read T ahead of their
// waits (stage 0's bulk load), which is only safe if whatever produced T -- possibly a kernel that triggers its dependents
// early, like this library's own forward stages -- had completed before the call's first kernel started.
template <typename Tel, bool FWD>
int launch_upper(const raht_plan *P, int k, const Tel *A, const Tel *B, Tel *O, int D, cudaStream_t st, bool head = false)
{
    auto kern = transform_upper_kernel<Tel, FWD>;
    const int64_t tiles = P->d.tiles[k];
    // Stages with many tiles (typically ~70 entries each, up to 64 dependent rounds) are bound by how many of those
    // chains an SM holds, i.e. by shared memory per CTA: channel groups over gridDim.y shrink the staging area a tile
    // needs.  Stages with few tiles are spread over up to 16 channel groups to fill the GPU.
    static const int env_groups = getenv("RAHT_UP_GROUPS") ? atoi(getenv("RAHT_UP_GROUPS")) : UP_GROUPS_MANY;
    static const int env_elems = getenv("RAHT_UP_ELEMS") ? atoi(getenv("RAHT_UP_ELEMS")) : UP_ELEMS_MANY;
    int groups, up_elems;
    if (tiles >= 2 * sm_count_cached()) {
        groups = env_groups;
        up_elems = env_elems;
    } else {
        groups = (int)((4 * (int64_t)sm_count_cached() + tiles - 1) / tiles);
        if (groups > UP_GROUPS_MAX) groups = UP_GROUPS_MAX;
        up_elems = UP_ELEMS;
    }
    if (groups > D) groups = D;
    if (up_elems > UP_ELEMS) up_elems = UP_ELEMS;
    if (up_elems < PLAN_UP_CAP) up_elems = PLAN_UP_CAP;                // one channel of the largest tile must fit
    const size_t smem = (size_t)(up_elems + 2 * PLAN_UP_CAP) * sizeof(Tel) + (size_t)PLAN_UP_CAP * 9 + 16;
    RAHT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(UP_ELEMS + 2 * PLAN_UP_CAP) * (int)sizeof(Tel) + PLAN_UP_CAP * 9 + 16));
    raht_count_launches(1);
    const bool use_pdl = raht_use_pdl(PDL_TRANSFORM) && !head;
    if (use_pdl) RAHT_CUDA(pdl_launch(kern, dim3((unsigned)tiles, (unsigned)groups), dim3(TR_THREADS), smem, st, P->d, k, A, B, O, D, up_elems));
    else kern<<<dim3((unsigned)tiles, (unsigned)groups), TR_THREADS, smem, st>>>(P->d, k, A, B, O, D, up_elems);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

// RAHT_CHAIN=0: a launch per stage (round 1), for A/B timing and as the cross-check of the chained kernels in the tests
unsigned next_epoch()
{
    static unsigned epoch = 0;
    unsigned e = __atomic_add_fetch(&epoch, 1u, __ATOMIC_RELAXED);
    if (e == 0) e = __atomic_add_fetch(&epoch, 1u, __ATOMIC_RELAXED);     // 0 = "never done" (plan_init)
    return e;
}

// RAHT_CHAIN: 0 = a launch per stage (default); 1 = stage 0 + ONE launch for all upper stages; 2 = everything in one launch.
// The chained forms are bit-identical (tests) and were built to remove the 5 dependent upper-stage launches of a 1 M-row
// transform; measured on B200 (scripts/gpu_chain_ab.sh, 1 M / 6 M rows) they LOSE: mode 1 raht 0.337 -> 0.367 ms, iraht
// 0.372 -> 0.420 ms; mode 2 raht 0.370, iraht 0.375 ms.  The upper stages are not launch-bound but bound by the latency of a
// tile's dependent rounds, and a chained tile is run by ONE CTA over all channels where the per-stage launches spread a stage
// over channel groups (up to 16 CTAs per tile); in mode 2 every stage-0 CTA additionally has to wait for its bulk store to
// LAND before it may sign in at its parent, which holds its 63 KB of shared memory ~1-2 us longer.  Kept as options.
inline int chain_mode()
{
    static const int m = getenv("RAHT_CHAIN") ? atoi(getenv("RAHT_CHAIN")) : 0;
    return m;
}

template <typename Tel> int upper_chain_smem(int *up_elems, size_t *smem)
{
    const size_t fixed = upper_fixed_bytes<Tel>();
    *up_elems = 4096;                                                     // 32 KB of rows: 3-4 CTAs per SM
    *smem = fixed + (size_t)*up_elems * sizeof(Tel);
    return RAHT_OK;
}

template <typename Tel> int launch_upper_forward_chain(const raht_plan *P, Tel *T, int D, cudaStream_t st)
{
    int up_elems;
    size_t smem;
    upper_chain_smem<Tel>(&up_elems, &smem);
    auto kern = transform_upper_forward_chain_kernel<Tel>;
    RAHT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    (raht_count_launches(1), kern)<<<(unsigned)P->d.tiles[1], TR_THREADS, smem, st>>>(P->d, T, D, up_elems);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

template <typename Tel> int launch_upper_inverse_chain(const raht_plan *P, const Tel *T, Tel *C, int D, cudaStream_t st)
{
    int up_elems;
    size_t smem;
    upper_chain_smem<Tel>(&up_elems, &smem);
    auto kern = transform_upper_inverse_chain_kernel<Tel>;
    RAHT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    (raht_count_launches(1), kern)<<<(unsigned)(P->d.tiles_total - P->d.tiles[0]), TR_THREADS, smem, st>>>(P->d, T, C, D, up_elems, next_epoch());
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

// shared memory of a chained kernel: the stage-0 layout, or what the largest upper tile needs for one channel
template <typename Tel> size_t chain_smem_bytes(int R, int D, int *up_elems)
{
    SmemLayout<Tel> L(R, D);
    size_t smem = L.total;
    const size_t fixed = upper_fixed_bytes<Tel>();
    const size_t min_upper = fixed + (size_t)2048 * sizeof(Tel);
    if (smem < min_upper) smem = min_upper;
    int ue = (int)((smem - fixed) / sizeof(Tel));
    if (ue > UP_ELEMS) ue = UP_ELEMS;
    if (ue < PLAN_UP_CAP) {                                              // one channel of the largest tile must fit
        ue = PLAN_UP_CAP;
        smem = fixed + (size_t)ue * sizeof(Tel);
    }
    *up_elems = ue;
    return smem;
}

template <typename Tin, typename Tel, int DT>
int launch_forward_chain(const raht_plan *P, const Tin *C, Tel *T, int D, int use_tma, cudaStream_t st)
{
    int up_elems = 0;
    const size_t smem = chain_smem_bytes<Tel>(P->d.tile_rows, D, &up_elems);
    if (smem > 227 * 1024) {
        raht_set_error("transform: tile of %d rows x %d channels needs %zu bytes of shared memory (> 227 KB)", P->d.tile_rows, D, smem);
        return RAHT_E_INVALID;
    }
    auto kern = transform_forward_chain_kernel<Tin, Tel, DT>;
    RAHT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    (raht_count_launches(1), kern)<<<(unsigned)P->d.tiles[0], TR_THREADS, smem, st>>>(P->d, C, T, D, use_tma, up_elems);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

template <typename Tel, int DT>
int launch_inverse_chain(const raht_plan *P, const Tel *T, Tel *C, int D, int use_tma, cudaStream_t st)
{
    int up_elems = 0;
    const size_t smem = chain_smem_bytes<Tel>(P->d.tile_rows, D, &up_elems);
    if (smem > 227 * 1024) {
        raht_set_error("transform: tile of %d rows x %d channels needs %zu bytes of shared memory (> 227 KB)", P->d.tile_rows, D, smem);
        return RAHT_E_INVALID;
    }
    auto kern = transform_inverse_chain_kernel<Tel, DT>;
    RAHT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    (raht_count_launches(1), kern)<<<(unsigned)P->d.tiles_total, TR_THREADS, smem, st>>>(P->d, T, C, D, use_tma, up_elems, next_epoch());
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

template <typename Tin, typename Tel>
int forward_impl(const raht_plan *P, const Tin *C, Tel *T, int D, int use_tma, cudaStream_t st)
{
    if (chain_mode() == 1) {
        int rc = launch_stage0_d<Tin, Tel, true>(P, C, (const Tel *)nullptr, T, D, use_tma, st);
        if (rc == RAHT_OK && P->d.n_stages > 1) rc = launch_upper_forward_chain<Tel>(P, T, D, st);
        return rc;
    }
    if (chain_mode() == 2) {
        switch (D) {
        case 3: return launch_forward_chain<Tin, Tel, 3>(P, C, T, D, use_tma, st);
        case 56: return launch_forward_chain<Tin, Tel, 56>(P, C, T, D, use_tma, st);
        case 59: return launch_forward_chain<Tin, Tel, 59>(P, C, T, D, use_tma, st);
        default: return launch_forward_chain<Tin, Tel, 0>(P, C, T, D, use_tma, st);
        }
    }
    // stage 0: C -> T (every row); stages >= 1: in place on T
    int rc = launch_stage0_d<Tin, Tel, true>(P, C, (const Tel *)nullptr, T, D, use_tma, st);
    for (int k = 1; k < P->d.n_stages && rc == RAHT_OK; ++k)
        rc = launch_upper<Tel, true>(P, k, (const Tel *)T, (const Tel *)nullptr, T, D, st);
    return rc;
}

template <typename Tel>
int inverse_impl(const raht_plan *P, const Tel *T, Tel *C, int D, int use_tma, cudaStream_t st)
{
    if (chain_mode() == 1) {
        int rc = RAHT_OK;
        if (P->d.n_stages > 1) rc = launch_upper_inverse_chain<Tel>(P, T, C, D, st);
        if (rc == RAHT_OK) rc = launch_stage0_d<Tel, Tel, false>(P, T, P->d.n_stages > 1 ? (const Tel *)C : (const Tel *)nullptr, C, D, use_tma, st);
        return rc;
    }
    if (chain_mode() == 2) {
        switch (D) {
        case 3: return launch_inverse_chain<Tel, 3>(P, T, C, D, use_tma, st);
        case 56: return launch_inverse_chain<Tel, 56>(P, T, C, D, use_tma, st);
        case 59: return launch_inverse_chain<Tel, 59>(P, T, C, D, use_tma, st);
        default: return launch_inverse_chain<Tel, 0>(P, T, C, D, use_tma, st);
        }
    }
    // top stage first.  A row is read from C when an upper stage has already rewritten it (the rows a
    // tile defers to the next stage), else from T; every row of a tile that ran is written to C.
    int rc = RAHT_OK;
    const int S = P->d.n_stages;
    for (int k = S - 1; k >= 1 && rc == RAHT_OK; --k)
        rc = launch_upper<Tel, false>(P, k, T, k == S - 1 ? (const Tel *)nullptr : (const Tel *)C, C, D, st, k == S - 1);
    if (rc == RAHT_OK) rc = launch_stage0_d<Tel, Tel, false>(P, T, S > 1 ? (const Tel *)C : (const Tel *)nullptr, C, D, use_tma, st);
    return rc;
}

template <typename Tin, typename Tel>
int forward_quant_impl(const raht_plan *P, const Tin *C, Tel *T, int D, int use_tma, const QuantArgs &Q, cudaStream_t st)
{
    int rc = launch_stage0_qd<Tin, Tel, true, QM_FWD>(P, C, (const Tel *)nullptr, T, D, use_tma, Q, st);
    for (int k = 1; k < P->d.n_stages && rc == RAHT_OK; ++k)
        rc = launch_upper<Tel, true>(P, k, (const Tel *)T, (const Tel *)nullptr, T, D, st);
    if (rc == RAHT_OK && P->d.n_stages > 1) {
        raht_count_launches(1);
        deferred_rows_kernel<Tel, true><<<(unsigned)P->d.tiles[0], TR_THREADS, 0, st>>>(P->d, T, D, Q);
        RAHT_LAUNCH_CHECK();
    }
    return rc;
}

template <typename Tel>
int dequant_inverse_impl(const raht_plan *P, Tel *T, Tel *C, int D, int use_tma, const QuantArgs &Q, cudaStream_t st)
{
    int rc = RAHT_OK;
    const int S = P->d.n_stages;
    if (S > 1) {
        raht_count_launches(1);
        deferred_rows_kernel<Tel, false><<<(unsigned)P->d.tiles[0], TR_THREADS, 0, st>>>(P->d, T, D, Q);
        RAHT_LAUNCH_CHECK();
    }
    for (int k = S - 1; k >= 1 && rc == RAHT_OK; --k)                     // (the deferred-rows kernel above is the call's head)
        rc = launch_upper<Tel, false>(P, k, (const Tel *)T, k == S - 1 ? (const Tel *)nullptr : (const Tel *)C, C, D, st);
    if (rc == RAHT_OK)
        rc = launch_stage0_qd<Tel, Tel, false, QM_INV>(P, (const Tel *)T, S > 1 ? (const Tel *)C : (const Tel *)nullptr, C, D, use_tma, Q, st);
    return rc;
}

}  // namespace

RAHT_API int raht_forward_quant(const raht_plan *plan, const void *C, int c_dtype, int D, void *T, int t_dtype, const int32_t *pos,
                                const double *steps, int32_t *q, int64_t ld, int32_t *status, int flags, void *stream)
{
    RAHT_REQUIRE(plan && D >= 1 && ld >= plan->d.n, "raht_forward_quant: bad arguments");
    if (plan->d.n == 0) return RAHT_OK;
    RAHT_REQUIRE(C && T && C != T && pos && steps && q, "raht_forward_quant: null or aliased pointer");
    cudaStream_t st = (cudaStream_t)stream;
    const int use_tma = (flags & 1) ? 0 : 1;
    QuantArgs Q;
    Q.pos = pos; Q.steps = steps; Q.q = q; Q.ld = ld; Q.status = status;
    if (c_dtype == RAHT_F32 && t_dtype == RAHT_F64) return forward_quant_impl<float, double>(plan, (const float *)C, (double *)T, D, use_tma, Q, st);
    if (c_dtype == RAHT_F64 && t_dtype == RAHT_F64) return forward_quant_impl<double, double>(plan, (const double *)C, (double *)T, D, use_tma, Q, st);
    if (c_dtype == RAHT_F32 && t_dtype == RAHT_F32) return forward_quant_impl<float, float>(plan, (const float *)C, (float *)T, D, use_tma, Q, st);
    raht_set_error("raht_forward_quant: unsupported dtype pair");
    return RAHT_E_INVALID;
}

RAHT_API int raht_dequant_inverse(const raht_plan *plan, const int32_t *q, int64_t ld, const int32_t *pos, const double *steps, int D,
                                  void *T, int t_dtype, void *C, int flags, void *stream)
{
    RAHT_REQUIRE(plan && D >= 1 && ld >= plan->d.n, "raht_dequant_inverse: bad arguments");
    if (plan->d.n == 0) return RAHT_OK;
    RAHT_REQUIRE(C && T && C != T && pos && steps && q, "raht_dequant_inverse: null or aliased pointer");
    cudaStream_t st = (cudaStream_t)stream;
    const int use_tma = (flags & 1) ? 0 : 1;
    QuantArgs Q;
    Q.pos = pos; Q.steps = steps; Q.q = const_cast<int32_t *>(q); Q.ld = ld; Q.status = nullptr;
    if (t_dtype == RAHT_F64) return dequant_inverse_impl<double>(plan, (double *)T, (double *)C, D, use_tma, Q, st);
    if (t_dtype == RAHT_F32) return dequant_inverse_impl<float>(plan, (float *)T, (float *)C, D, use_tma, Q, st);
    raht_set_error("raht_dequant_inverse: bad dtype");
    return RAHT_E_INVALID;
}

RAHT_API int raht_forward(const raht_plan *plan, const void *C, int c_dtype, int D, void *T, int t_dtype, int flags,
                          void *stream)
{
    RAHT_REQUIRE(plan && D >= 1, "raht_forward: bad arguments");
    if (plan->d.n == 0) return RAHT_OK;
    RAHT_REQUIRE(C && T && C != T, "raht_forward: C and T must be distinct non-null device pointers");
    cudaStream_t st = (cudaStream_t)stream;
    const int use_tma = (flags & 1) ? 0 : 1;
    if (c_dtype == RAHT_F64 && t_dtype == RAHT_F64) return forward_impl<double, double>(plan, (const double *)C, (double *)T, D, use_tma, st);
    if (c_dtype == RAHT_F32 && t_dtype == RAHT_F64) return forward_impl<float, double>(plan, (const float *)C, (double *)T, D, use_tma, st);
    if (c_dtype == RAHT_F32 && t_dtype == RAHT_F32) return forward_impl<float, float>(plan, (const float *)C, (float *)T, D, use_tma, st);
    if (c_dtype == RAHT_F64 && t_dtype == RAHT_F32) return forward_impl<double, float>(plan, (const double *)C, (float *)T, D, use_tma, st);
    raht_set_error("raht_forward: bad dtype");
    return RAHT_E_INVALID;
}

RAHT_API int raht_inverse(const raht_plan *plan, const void *T, int t_dtype, int D, void *C, int flags, void *stream)
{
    RAHT_REQUIRE(plan && D >= 1, "raht_inverse: bad arguments");
    if (plan->d.n == 0) return RAHT_OK;
    RAHT_REQUIRE(C && T && C != T, "raht_inverse: T and C must be distinct non-null device pointers");
    cudaStream_t st = (cudaStream_t)stream;
    const int use_tma = (flags & 1) ? 0 : 1;
    if (t_dtype == RAHT_F64) return inverse_impl<double>(plan, (const double *)T, (double *)C, D, use_tma, st);
    if (t_dtype == RAHT_F32) return inverse_impl<float>(plan, (const float *)T, (float *)C, D, use_tma, st);
    raht_set_error("raht_inverse: bad dtype");
    return RAHT_E_INVALID;
}
