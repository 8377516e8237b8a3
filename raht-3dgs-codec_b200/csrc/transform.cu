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

struct StageArgs {
    const uint32_t *rows;      // nullptr: entries are the rows themselves (stage 0, contiguous tiles)
    const uint8_t *to_next;
    const uint32_t *sched;
    const double2 *ab;
    const TileHdr *hdr;
};

template <typename Tel> struct SmemLayout {
    size_t off_ab, off_sched, off_rowid, off_hdr, off_flag, off_bar, total;
    __host__ __device__ SmemLayout(int R, int D)
    {
        size_t o = (size_t)R * D * sizeof(Tel);
        o = (o + 15) & ~(size_t)15;
        off_ab = o; o += (size_t)R * 2 * sizeof(Tel);
        off_sched = o; o += (size_t)R * 4;
        off_rowid = o; o += (size_t)R * 4;
        off_hdr = o; o += sizeof(TileHdr);
        off_bar = o; o += 8;
        off_flag = o; o += (size_t)R;
        total = (o + 15) & ~(size_t)15;
    }
};

// A: primary source (row-major, Tin).  B: secondary source for rows flagged in sel (inverse: rows the
// upper stages already wrote), may be nullptr.  O: destination.  FWD: forward butterflies in schedule
// order, else inverse butterflies in reverse order.
template <typename Tin, typename Tel, int DT, bool FWD>
__global__ void __launch_bounds__(TR_THREADS)
transform_stage_kernel(StageArgs a, const Tin *__restrict__ A, const Tel *B, const uint8_t *__restrict__ sel, Tel *O,
                       int D_rt, int R, int use_tma)
{
    extern __shared__ __align__(128) unsigned char tr_smem[];
    const int D = DT ? DT : D_rt;
    const SmemLayout<Tel> L(R, D);
    Tel *s_rows = (Tel *)tr_smem;
    Tel *s_ab = (Tel *)(tr_smem + L.off_ab);
    uint32_t *s_sched = (uint32_t *)(tr_smem + L.off_sched);
    uint32_t *s_rowid = (uint32_t *)(tr_smem + L.off_rowid);
    TileHdr *s_hdr = (TileHdr *)(tr_smem + L.off_hdr);
    uint64_t *s_bar = (uint64_t *)(tr_smem + L.off_bar);
    uint8_t *s_flag = tr_smem + L.off_flag;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t tile = blockIdx.x;
    const int64_t base = tile * R;
    const bool contiguous = (a.rows == nullptr);

    if (tid < (int)(sizeof(TileHdr) / 4)) ((uint32_t *)s_hdr)[tid] = ((const uint32_t *)(a.hdr + tile))[tid];
    __syncthreads();
    const int n_rows = s_hdr->n_rows;
    const int n_bf = s_hdr->n_bf;
    const int n_rounds = s_hdr->n_rounds;
    const size_t tile_bytes = (size_t)n_rows * D * sizeof(Tel);

    // bulk copies need 16-byte granularity; tiles that do not qualify (ragged last tile) use plain loads
    bool tma_in = false, tma_out = false;
    if constexpr (sizeof(Tin) == sizeof(Tel)) {
        if (contiguous && use_tma && (tile_bytes % 16) == 0) {
            tma_in = (((uintptr_t)(A + base * D)) % 16) == 0;
            tma_out = (((uintptr_t)(O + base * D)) % 16) == 0;
        }
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

    // schedule, coefficients, row ids and source flags (overlaps the bulk copy)
    for (int j = tid; j < n_bf; j += TR_THREADS) {
        s_sched[j] = a.sched[base + j];
        double2 c = a.ab[base + j];
        s_ab[2 * j] = (Tel)c.x;
        s_ab[2 * j + 1] = (Tel)c.y;
    }
    const bool patch = (B != nullptr) && (sel != nullptr);
    for (int j = tid; j < n_rows; j += TR_THREADS) {
        s_rowid[j] = contiguous ? (uint32_t)(base + j) : a.rows[base + j];
        s_flag[j] = patch ? sel[base + j] : (uint8_t)0;
    }
    __syncthreads();

    if (contiguous) {
        if (tma_in) {
            mbar_wait(s_bar, 0);
        } else {
            const Tin *src = A + base * D;
            const int n_el = n_rows * D;
            for (int i = tid; i < n_el; i += TR_THREADS) s_rows[i] = (Tel)src[i];
        }
        if (patch) {
            if (!tma_in) __syncthreads();
            for (int j = warp; j < n_rows; j += TR_WARPS)
                if (s_flag[j]) {
                    const Tel *src = B + (size_t)(base + j) * D;
                    for (int c = lane; c < D; c += 32) s_rows[j * D + c] = src[c];
                }
        }
    } else {
        for (int j = warp; j < n_rows; j += TR_WARPS) {
            const size_t row = s_rowid[j];
            if (s_flag[j]) {
                const Tel *src = B + row * D;
                for (int c = lane; c < D; c += 32) s_rows[j * D + c] = src[c];
            } else {
                const Tin *src = A + row * D;
                for (int c = lane; c < D; c += 32) s_rows[j * D + c] = (Tel)src[c];
            }
        }
    }
    __syncthreads();

    // butterflies, one dependency round at a time
    using AR = Arith<Tel>;
    for (int kk = 0; kk < n_rounds; ++kk) {
        const int k = FWD ? kk : n_rounds - 1 - kk;
        const int lo = k ? s_hdr->round_end[k - 1] : 0;
        const int hi = s_hdr->round_end[k];
        const int work = (hi - lo) * D;
        for (int idx = tid; idx < work; idx += TR_THREADS) {
            const int e = lo + idx / D;
            const int c = idx - (idx / D) * D;
            const uint32_t ent = s_sched[e];
            const int r = ent & 1023, s = (ent >> 10) & 1023;
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
        __syncthreads();
    }

    // write back
    if (contiguous) {
        if (tma_out) {
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
                bulk_wait_read0();
            }
        } else {
            Tel *dst = O + base * D;
            const int n_el = n_rows * D;
            for (int i = tid; i < n_el; i += TR_THREADS) dst[i] = s_rows[i];
        }
    } else {
        for (int j = warp; j < n_rows; j += TR_WARPS) {
            Tel *dst = O + (size_t)s_rowid[j] * D;
            for (int c = lane; c < D; c += 32) dst[c] = s_rows[j * D + c];
        }
    }
}

template <typename Tin, typename Tel, int DT, bool FWD>
int launch_stage(const raht_plan *P, int k, const Tin *A, const Tel *B, const uint8_t *sel, Tel *O, int D, int use_tma,
                 cudaStream_t st)
{
    const PlanStage &S = P->st[k];
    const int R = P->tile_rows;
    SmemLayout<Tel> L(R, D);
    auto kern = transform_stage_kernel<Tin, Tel, DT, FWD>;
    if (L.total > 227 * 1024) {
        raht_set_error("transform: tile of %d rows x %d channels needs %zu bytes of shared memory (> 227 KB)", R, D, L.total);
        return RAHT_E_INVALID;
    }
    RAHT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.total));
    StageArgs a{S.rows, S.to_next, S.sched, S.ab, S.hdr};
    (raht_count_launches(1), kern)<<<(unsigned)S.n_tiles, TR_THREADS, L.total, st>>>(a, A, B, sel, O, D, R, use_tma);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

template <typename Tin, typename Tel, bool FWD>
int launch_stage_d(const raht_plan *P, int k, const Tin *A, const Tel *B, const uint8_t *sel, Tel *O, int D, int use_tma,
                   cudaStream_t st)
{
    switch (D) {
    case 3: return launch_stage<Tin, Tel, 3, FWD>(P, k, A, B, sel, O, D, use_tma, st);
    case 56: return launch_stage<Tin, Tel, 56, FWD>(P, k, A, B, sel, O, D, use_tma, st);
    case 59: return launch_stage<Tin, Tel, 59, FWD>(P, k, A, B, sel, O, D, use_tma, st);
    default: return launch_stage<Tin, Tel, 0, FWD>(P, k, A, B, sel, O, D, use_tma, st);
    }
}

template <typename Tin, typename Tel>
int forward_impl(const raht_plan *P, const Tin *C, Tel *T, int D, int use_tma, cudaStream_t st)
{
    // stage 0: C -> T (every row); stages >= 1: in place on T
    int rc = launch_stage_d<Tin, Tel, true>(P, 0, C, (const Tel *)nullptr, nullptr, T, D, use_tma, st);
    for (int k = 1; k < P->n_stages && rc == RAHT_OK; ++k)
        rc = launch_stage_d<Tel, Tel, true>(P, k, (const Tel *)T, (const Tel *)nullptr, nullptr, T, D, use_tma, st);
    return rc;
}

template <typename Tel>
int inverse_impl(const raht_plan *P, const Tel *T, Tel *C, int D, int use_tma, cudaStream_t st)
{
    // top stage first.  A row is read from C when an upper stage has already rewritten it (the rows a
    // stage carries into the next one), else from T; every row of a stage is written to C.
    int rc = RAHT_OK;
    for (int k = P->n_stages - 1; k >= 0 && rc == RAHT_OK; --k) {
        const bool top = (k == P->n_stages - 1);
        rc = launch_stage_d<Tel, Tel, false>(P, k, T, top ? (const Tel *)nullptr : (const Tel *)C,
                                             top ? nullptr : P->st[k].to_next, C, D, use_tma, st);
    }
    return rc;
}

}  // namespace

RAHT_API int raht_forward(const raht_plan *plan, const void *C, int c_dtype, int D, void *T, int t_dtype, int flags,
                          void *stream)
{
    RAHT_REQUIRE(plan && D >= 1, "raht_forward: bad arguments");
    if (plan->n == 0) return RAHT_OK;
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
    if (plan->n == 0) return RAHT_OK;
    RAHT_REQUIRE(C && T && C != T, "raht_inverse: T and C must be distinct non-null device pointers");
    cudaStream_t st = (cudaStream_t)stream;
    const int use_tma = (flags & 1) ? 0 : 1;
    if (t_dtype == RAHT_F64) return inverse_impl<double>(plan, (const double *)T, (double *)C, D, use_tma, st);
    if (t_dtype == RAHT_F32) return inverse_impl<float>(plan, (const float *)T, (float *)C, D, use_tma, st);
    raht_set_error("raht_inverse: bad dtype");
    return RAHT_E_INVALID;
}
