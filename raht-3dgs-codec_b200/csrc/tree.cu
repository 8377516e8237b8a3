// tree.cu -- RAHT tree construction.
// Replaces python/RAHT_param.py:6-83 (RAHT_param) and :191-279 (RAHT_param_reorder_fast): instead
// of up to 3*depth rounds of gather / xor / mask / boolean-compaction with a host sync each, one
// XOR+CLZ pass gives level[] (App. D), from which List/Flags/weights (any level, on demand),
// order_RAGFT (one stable 22-way partition), w, and the shared-memory transform plan are derived.
#include "plan.cuh"
#include <math.h>
#include <new>

namespace {

// ---- level[] ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) levels_kernel(const int64_t *__restrict__ MC, int64_t n, uint8_t *__restrict__ level,
                                                     int32_t *status)
{
    pdl_enter();
    bool bad = false;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        if (i == 0) { level[0] = 255; continue; }
        int64_t a = MC[i - 1], b = MC[i];
        bad |= !(a < b);
        uint64_t x = (uint64_t)a ^ (uint64_t)b;
        level[i] = x ? (uint8_t)(63 - __clzll((long long)x)) : (uint8_t)0;
    }
    if (__any_sync(~0u, bad) && lane_id() == 0 && status) atomicOr(status, RAHT_ST_NOT_SORTED_UNIQUE);
}

// ---- one level of List / Flags / weights (RAHT_param.py:226-248) -----------------------------------
struct LevelPred {
    const uint8_t *lev;
    int l;
    __device__ bool operator()(int64_t i) const { return i == 0 || lev[i] >= l; }
};
struct LevelEmit {
    int64_t *list;
    __device__ void operator()(unsigned long long pos, int64_t i) const { list[pos] = i; }
};
__global__ void __launch_bounds__(256) level_fill_kernel(const uint8_t *__restrict__ lev, int64_t n, int l,
                                                         const int64_t *__restrict__ list, const int64_t *count,
                                                         uint8_t *__restrict__ flags, int64_t *__restrict__ weights)
{
    const int64_t L = *count;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < L; k += (int64_t)gridDim.x * blockDim.x) {
        int64_t cur = list[k];
        int64_t nxt = (k + 1 < L) ? list[k + 1] : n;                    // :228-230
        if (weights) weights[k] = nxt - cur;
        if (flags) flags[k] = (k + 1 < L) && (lev[nxt] == l);           // :238-243
    }
}

__global__ void max_level_init(int32_t *m) { *m = -1; }
__global__ void __launch_bounds__(256) max_level_kernel(const uint8_t *__restrict__ lev, int64_t n, int32_t *out)
{
    int m = -1;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x + 1; i < n; i += (int64_t)gridDim.x * blockDim.x)
        m = max(m, (int)lev[i]);
    for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(~0u, m, o));
    if (lane_id() == 0 && m >= 0) atomicMax(out, m);
}

// ---- order_RAGFT: stable partition into 22 buckets (RAHT_param.py:219-274) -------------------------
constexpr int OR_BUCKETS = 22;       // bucket 0 = root, 1 + (20 - level/3) otherwise (coarse -> fine)
constexpr int OR_THREADS = 256;
constexpr int OR_IPT = 8;
constexpr int OR_TILE = OR_THREADS * OR_IPT;
constexpr int OR_WARPS = OR_THREADS / 32;

// the root is the row with level 255 (row 0 of a whole cloud; the first row of a Morton-prefix shard other than the
// first carries its level in the whole scene instead, shard.cu)
__device__ __forceinline__ int or_bucket(int64_t, uint8_t lev) { return lev == 255 ? 0 : 1 + (20 - min((int)lev, 62) / 3); }

__global__ void __launch_bounds__(OR_THREADS) order_count_kernel(const uint8_t *__restrict__ lev, int64_t n, int64_t nblocks,
                                                                unsigned long long *counts /* [bucket][block] */)
{
    pdl_enter();
    __shared__ int hist[OR_BUCKETS];
    if (threadIdx.x < OR_BUCKETS) hist[threadIdx.x] = 0;
    __syncthreads();
    int64_t base = (int64_t)blockIdx.x * OR_TILE;
    for (int k = 0; k < OR_IPT; ++k) {
        int64_t i = base + k * OR_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&hist[or_bucket(i, lev[i])], 1);
    }
    __syncthreads();
    if (threadIdx.x < OR_BUCKETS) counts[(size_t)threadIdx.x * nblocks + blockIdx.x] = (unsigned long long)hist[threadIdx.x];
}

__global__ void __launch_bounds__(OR_THREADS) order_scatter_kernel(const uint8_t *__restrict__ lev, int64_t n, int64_t nblocks,
                                                                  const unsigned long long *__restrict__ offsets,
                                                                  int64_t *__restrict__ order, int32_t *__restrict__ position)
{
    pdl_enter();
    __shared__ int whist[OR_WARPS][OR_BUCKETS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < OR_WARPS * OR_BUCKETS; i += OR_THREADS) (&whist[0][0])[i] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * OR_TILE;
    int bucket[OR_IPT], rank[OR_IPT];
#pragma unroll
    for (int j = 0; j < OR_IPT; ++j) {
        int64_t i = base + warp * (32 * OR_IPT) + j * 32 + lane;     // warp-striped: (warp, j, lane) == memory order
        int b = (i < n) ? or_bucket(i, lev[i]) : OR_BUCKETS;          // OR_BUCKETS = "none"
        unsigned peers = __match_any_sync(~0u, b);
        int leader = __ffs(peers) - 1;
        int before = __popc(peers & lanemask_lt());
        int old = 0;
        if (lane == leader && b < OR_BUCKETS) {
            old = whist[warp][b];
            whist[warp][b] = old + __popc(peers);
        }
        old = __shfl_sync(~0u, old, leader);
        bucket[j] = b;
        rank[j] = old + before;
        __syncwarp();
    }
    __syncthreads();
    if (tid < OR_BUCKETS) {
        int run = 0;
        for (int w = 0; w < OR_WARPS; ++w) {
            int c = whist[w][tid];
            whist[w][tid] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < OR_IPT; ++j) {
        int64_t i = base + warp * (32 * OR_IPT) + j * 32 + lane;
        if (i < n) {
            int b = bucket[j];
            unsigned long long dst = offsets[(size_t)b * nblocks + blockIdx.x] + (unsigned)(whist[warp][b] + rank[j]);
            if (order) order[dst] = i;
            if (position) position[i] = (int32_t)dst;
        }
    }
}

// ---- transform plan: one tile of one stage (plan.cuh) ------------------------------------------------
constexpr int PB_THREADS = 128;
constexpr int PB_WARPS = PB_THREADS / 32;

// arrive[]: one counter per upper tile = how many of its child tiles have been planned
__global__ void __launch_bounds__(256) plan_init_kernel(PlanDev P)
{
    pdl_enter();
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { *P.bump = (unsigned int)P.slots0; *P.ticket = 0u; }
    if (i < P.tiles_total) { P.arrive[i] = 0u; P.done[i] = 0u; }
}

// One tile of one stage.  Upper tiles read what OTHER CTAs of the same launch wrote (the deferred lists of their child
// tiles): those loads go to L2 (__ldcg), behind the fence / counter hand-over of plan_kernel.
__device__ __forceinline__ void plan_tile(const PlanDev &P, int stage, int64_t tile, const uint8_t *__restrict__ level,
                                          double *__restrict__ w)
{
    __shared__ uint32_t s_orig[PLAN_UP_CAP + 1];   // original row of each entry (+ virtual end of the cloud)
    __shared__ uint16_t s_s[PLAN_UP_CAP];          // local index of the left row, 0xffff = not schedulable here
    __shared__ uint16_t s_t[PLAN_UP_CAP];          // local index of the right boundary
    __shared__ uint8_t s_lev[PLAN_UP_CAP + 1];
    __shared__ int hist[PLAN_MAX_ROUNDS], offs[PLAN_MAX_ROUNDS], cur[PLAN_MAX_ROUNDS];
    __shared__ int s_off[PLAN_FANIN + 1];
    __shared__ int wsum[PB_WARPS];
    __shared__ int s_nbf, s_nr, s_bfoff, s_run, s_npend, s_lastroot, s_firstroot;
    __shared__ uint16_t s_pend[PLAN_UP_CAP];       // entries whose partners lie far away

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t gt = P.tile_base[stage] + tile;
    const bool last = tile == P.tiles[stage] - 1;
    const int64_t n_total = P.n;

    if (tid < PLAN_MAX_ROUNDS) { hist[tid] = 0; cur[tid] = 0; }
    if (tid == 0) { s_npend = 0; s_lastroot = -1; s_firstroot = 0x7fffffff; }
    int n_rows;
    if (stage == 0) {
        const int64_t base = tile * P.tile_rows;
        n_rows = (int)min((int64_t)P.tile_rows, n_total - base);
        for (int j = tid; j < n_rows; j += PB_THREADS) {
            s_lev[j] = level[base + j];
            s_orig[j] = (uint32_t)(base + j);
        }
    } else {
        const int64_t c0 = tile * PLAN_FANIN;
        const int64_t cbase = P.tile_base[stage - 1];
        if (tid < 32) {                                      // entries per child tile: eight independent loads, one scan
            int cnt = 0;
            if (tid < PLAN_FANIN && c0 + tid < P.tiles[stage - 1]) cnt = min((int)__ldcg(&P.hdr[cbase + c0 + tid].n_def), PLAN_DEFER_MAX);
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
        n_rows = s_off[PLAN_FANIN];
        // the children's deferred lists in one flat pass, the loads of a batch in flight at once (a loop per child was eight
        // dependent round trips); batches of 3 x 128 entries: typical upper tiles hold 70-150, and the registers of a
        // deeper batch cost the stage-0 tiles of the same kernel their occupancy
        constexpr int RB = 3;
        for (int j0 = 0; j0 < n_rows; j0 += RB * PB_THREADS) {
            uint32_t rid[RB];
            uint8_t rlv[RB];
#pragma unroll
            for (int u = 0; u < RB; ++u) {
                const int j = j0 + tid + u * PB_THREADS;
                if (j < n_rows) {
                    int c = 0;
#pragma unroll
                    for (int k = 1; k < PLAN_FANIN; ++k) c += j >= s_off[k] ? 1 : 0;
                    const size_t src = (size_t)(cbase + c0 + c) * PLAN_DEFER_MAX + (j - s_off[c]);
                    rid[u] = __ldcg(P.def_rows + src);
                    rlv[u] = __ldcg(P.def_lev + src);
                }
            }
#pragma unroll
            for (int u = 0; u < RB; ++u) {
                const int j = j0 + tid + u * PB_THREADS;
                if (j < n_rows) {
                    s_orig[j] = rid[u];
                    s_lev[j] = rlv[u];
                }
            }
        }
    }
    if (tid == 0 && last) { s_lev[n_rows] = 255; s_orig[n_rows] = (uint32_t)n_total; }   // virtual end of the cloud
    __syncthreads();
    for (int j = tid; j < n_rows; j += PB_THREADS)
        if (s_lev[j] == 255) { atomicMax(&s_lastroot, j); atomicMin(&s_firstroot, j); }
    __syncthreads();

    // partner search: s = nearest earlier entry with a higher level, t = nearest later one.  Nearly every entry finds
    // both within a few positions; the few that have to look far (the high levels of an upper tile: up to 1000 steps,
    // which used to set the duration of the whole stage) are handed to a warp each, 32 positions per step.
    const int lim = last ? n_rows + 1 : n_rows;
    constexpr int NEAR = 16;
    for (int j = tid; j < n_rows; j += PB_THREADS) {
        const int l = s_lev[j];
        if (l == 255) {                                   // a root: never a right sibling
            // Level 255 marks row 0 of a cloud, or the first row of every cell of a FOREST (the sharded encoder transforms
            // the aligned Morton-prefix cells of a rank side by side, shard.cu).  A root is the left partner of the
            // high-level entries of its own cell AND the right boundary of those of the cell before it, so it travels up
            // the hierarchy as long as either cell reaches out of this tile: the LAST root of a tile goes up unless the
            // cloud ends here, the FIRST one unless the cloud starts with it; a root between two others has both cells
            // entirely inside the tile and is final.  At most two roots per tile are handed up: with at most 63 + 63
            // ordinary entries (levels 0 .. 62) the 129-entry bound of the deferred list holds for any number of cells.
            const bool up = !((j != s_firstroot || s_orig[j] == 0) && (j != s_lastroot || last));
            s_s[j] = 0xffff;
            s_t[j] = up ? 1 : 0;                          // s_t doubles as the "deferred" flag for unschedulable entries
            if (!up && w && s_orig[j] == 0) w[0] = (double)n_total;   // RAHT.py:325-328: the root of a whole cloud carries every row
            continue;
        }
        int s = -1, t = -1;
        bool far = false;
        const int lo = max(0, j - NEAR);
        for (int i = j - 1; i >= lo; --i)
            if (s_lev[i] > l) { s = i; break; }
        if (s < 0 && lo > 0) far = true;
        if (!far && s >= 0) {
            const int hi = min(lim, j + 1 + NEAR);
            for (int i = j + 1; i < hi; ++i)
                if (s_lev[i] > l) { t = i; break; }
            if (t < 0 && hi < lim) far = true;
        }
        if (far) {
            s_pend[atomicAdd(&s_npend, 1)] = (uint16_t)j;
            continue;
        }
        const bool in_tile = (s >= 0) && (t >= 0);
        s_s[j] = in_tile ? (uint16_t)s : (uint16_t)0xffff;
        s_t[j] = in_tile ? (uint16_t)t : (uint16_t)1;
        if (in_tile) atomicAdd(&hist[min(l, PLAN_MAX_ROUNDS - 1)], 1);
    }
    __syncthreads();
    for (int p = warp; p < s_npend; p += PB_WARPS) {
        const int j = s_pend[p];
        const int l = s_lev[j];
        int s = -1, t = -1;
        for (int base = j - 1; base >= 0 && s < 0; base -= 32) {
            const int i = base - lane;
            const unsigned b = __ballot_sync(~0u, i >= 0 && s_lev[i] > l);
            if (b) s = base - (__ffs((int)b) - 1);
        }
        if (s >= 0)
            for (int base = j + 1; base < lim && t < 0; base += 32) {
                const int i = base + lane;
                const unsigned b = __ballot_sync(~0u, i < lim && s_lev[i] > l);
                if (b) t = base + (__ffs((int)b) - 1);
            }
        if (lane == 0) {
            const bool in_tile = (s >= 0) && (t >= 0);
            s_s[j] = in_tile ? (uint16_t)s : (uint16_t)0xffff;
            s_t[j] = in_tile ? (uint16_t)t : (uint16_t)1;
            if (in_tile) atomicAdd(&hist[min(l, PLAN_MAX_ROUNDS - 1)], 1);
        }
    }
    if (tid == 0) s_run = 0;
    __syncthreads();

    // deferred entries, in order, into this tile's slot of the next stage's input
    for (int base = 0; base < n_rows; base += PB_THREADS) {
        const int j = base + tid;
        const bool p = j < n_rows && s_s[j] == 0xffff && s_t[j] == 1;
        const unsigned b = __ballot_sync(~0u, p);
        if (lane == 0) wsum[warp] = __popc(b);
        __syncthreads();
        int woff = 0, tot = 0;
        for (int x = 0; x < PB_WARPS; ++x) {
            if (x < warp) woff += wsum[x];
            tot += wsum[x];
        }
        const int pos = s_run + woff + __popc(b & lanemask_lt());
        if (p && pos < PLAN_DEFER_MAX) {
            const size_t dst = (size_t)gt * PLAN_DEFER_MAX + pos;
            P.def_rows[dst] = s_orig[j];
            P.def_lev[dst] = s_lev[j];
            P.def_local[dst] = (uint16_t)j;
        }
        __syncthreads();
        if (tid == 0) s_run += tot;
        __syncthreads();
    }

    TileHdr *H = P.hdr + gt;
    if (tid < 32) {                                       // scan of the 64-bin level histogram by one warp
        int h0 = hist[tid], h1 = hist[tid + 32];
        int x0 = h0, x1 = h1;
        for (int o = 1; o < 32; o <<= 1) {
            int y0 = __shfl_up_sync(~0u, x0, o), y1 = __shfl_up_sync(~0u, x1, o);
            if (tid >= o) { x0 += y0; x1 += y1; }
        }
        int tot0 = __shfl_sync(~0u, x0, 31);
        offs[tid] = x0 - h0;
        offs[tid + 32] = tot0 + x1 - h1;
        unsigned m0 = __ballot_sync(~0u, h0 > 0), m1 = __ballot_sync(~0u, h1 > 0);
        if (h0 > 0) H->round_end[__popc(m0 & lanemask_lt())] = (uint16_t)x0;
        if (h1 > 0) H->round_end[__popc(m0) + __popc(m1 & lanemask_lt())] = (uint16_t)(tot0 + x1);
        if (tid == 31) {
            s_nbf = tot0 + x1;
            s_nr = __popc(m0) + __popc(m1);
            s_bfoff = stage == 0 ? (int)(tile * P.tile_rows) : (int)atomicAdd(P.bump, (unsigned int)(tot0 + x1));
        }
    }
    __syncthreads();
    const size_t bf_off = (size_t)(unsigned int)s_bfoff;
    for (int j = tid; j < n_rows; j += PB_THREADS) {
        const int s = s_s[j];
        if (s == 0xffff) continue;
        const int l = s_lev[j], t = s_t[j];
        const int p = offs[l] + atomicAdd(&cur[l], 1);
        const double w0 = (double)(s_orig[j] - s_orig[s]);
        const double w1 = (double)(s_orig[t] - s_orig[j]);
        const double den = __dadd_rn(w0, w1);                           // RAHT.py:317
        P.sched[bf_off + p] = sched_pack(j, s, l);
        P.ab[bf_off + p] = make_double2(__dsqrt_rn(__ddiv_rn(w0, den)), __dsqrt_rn(__ddiv_rn(w1, den)));   // :321-322
        if (w) w[s_orig[j]] = den;                                      // :325-328 (final value of w[i1])
    }
    if (tid == 0) {
        H->n_rows = (uint16_t)n_rows;
        H->n_bf = (uint16_t)s_nbf;
        H->n_rounds = (uint16_t)s_nr;
        H->n_def = (uint16_t)min(s_run, PLAN_DEFER_MAX);
        H->bf_off = (uint32_t)s_bfoff;
        H->pad = 0;
    }
}

// The whole hierarchy in ONE launch of tiles[0] CTAs: a CTA plans its stage-0 tile, then signs in at the parent tile's
// counter; the CTA that signs in last (all PLAN_FANIN children are done, their deferred lists fenced to L2) goes on to
// plan the parent, and so on up to the single top tile.  Nobody waits, so no co-residency is needed.  (Round 1 launched
// one kernel per stage: 6-7 dependent launches of ~20 us for a hierarchy whose upper stages have 977 / 123 / 16 / 2 / 1
// tiles at 1 M rows.)
__global__ void __launch_bounds__(PB_THREADS) plan_kernel(PlanDev P, const uint8_t *__restrict__ level, double *__restrict__ w)
{
    pdl_enter();
    __shared__ int s_last;
    int stage = 0;
    int64_t tile = blockIdx.x;
    for (;;) {
        plan_tile(P, stage, tile, level, w);
        if (stage + 1 >= P.n_stages) break;
        const int64_t parent = tile / PLAN_FANIN;
        const int64_t children = min((int64_t)PLAN_FANIN, P.tiles[stage] - parent * PLAN_FANIN);
        __threadfence();                                   // this tile's header and deferred list, before the sign-in
        __syncthreads();
        if (threadIdx.x == 0) {
            const unsigned prev = atomicAdd(P.arrive + P.tile_base[stage + 1] + parent, 1u);
            s_last = prev + 1 == (unsigned)children;
            if (s_last) P.arrive[P.tile_base[stage + 1] + parent] = 0u;       // zero again for the transforms that reuse it
            __threadfence();
        }
        __syncthreads();
        if (!s_last) break;
        ++stage;
        tile = parent;
    }
}

struct PlanLayout {
    PlanDev d;          // sizes and stage table filled in; pointers as byte offsets from the workspace start
    size_t off_hdr, off_def_rows, off_def_local, off_def_lev, off_sched, off_ab, off_bump, off_arrive, off_done, total;
};

PlanLayout plan_layout(int64_t n, int tile_rows)
{
    PlanLayout L;
    PlanDev &d = L.d;
    d = PlanDev();
    d.n = n;
    d.tile_rows = tile_rows;
    int k = 0;
    int64_t t = n > 0 ? (n + tile_rows - 1) / tile_rows : 0, base = 0;
    while (t > 0) {
        d.tiles[k] = t;
        d.tile_base[k] = base;
        base += t;
        ++k;
        if (t == 1 || k >= PLAN_MAX_STAGES) break;
        t = (t + PLAN_FANIN - 1) / PLAN_FANIN;
    }
    d.n_stages = k;
    d.tiles_total = base;
    d.slots0 = k ? d.tiles[0] * tile_rows : 0;
    size_t o = 0;
    L.off_hdr = o; o += align_up((size_t)base * sizeof(TileHdr), 256);
    L.off_def_rows = o; o += align_up((size_t)base * PLAN_DEFER_MAX * 4, 256);
    L.off_def_local = o; o += align_up((size_t)base * PLAN_DEFER_MAX * 2, 256);
    L.off_def_lev = o; o += align_up((size_t)base * PLAN_DEFER_MAX, 256);
    L.off_sched = o; o += align_up((size_t)(d.slots0 + n) * 4, 256);
    L.off_ab = o; o += align_up((size_t)(d.slots0 + n) * 16, 256);
    L.off_bump = o; o += 256;
    L.off_arrive = o; o += align_up((size_t)base * 4, 256);
    L.off_done = o; o += align_up((size_t)base * 4, 256);
    L.total = o;
    return L;
}

}  // namespace

RAHT_API int raht_tree_levels(const int64_t *MC, int64_t n, uint8_t *level, int32_t *status, void *stream)
{
    RAHT_REQUIRE(n >= 0 && (n == 0 || (MC && level)), "raht_tree_levels: bad arguments");
    if (n == 0) return RAHT_OK;
    RAHT_PDL_LAUNCH(PDL_TREE, levels_kernel, grid_for(n, 256, 8), 256, 0, (cudaStream_t)stream, MC, n, level, status);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API size_t raht_tree_level_workspace_bytes(int64_t n) { return sel_workspace_bytes(n < 0 ? 0 : n) + 256; }

RAHT_API int raht_tree_level(const uint8_t *level, int64_t n, int l, int64_t *list, uint8_t *flags, int64_t *weights,
                             int64_t *count, void *workspace, size_t workspace_bytes, void *stream)
{
    RAHT_REQUIRE(n >= 1 && level && list && count && workspace && l >= 0 && l < 64, "raht_tree_level: bad arguments");
    if (workspace_bytes < raht_tree_level_workspace_bytes(n)) {
        raht_set_error("raht_tree_level: workspace too small");
        return RAHT_E_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    LevelPred pred{level, l};
    LevelEmit emit{list};
    device_select(n, pred, emit, count, workspace, st);
    (raht_count_launches(1), level_fill_kernel)<<<grid_for(n, 256, 8), 256, 0, st>>>(level, n, l, list, count, flags, weights);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_tree_max_level(const uint8_t *level, int64_t n, int32_t *max_level, void *stream)
{
    RAHT_REQUIRE(n >= 0 && max_level && (n == 0 || level), "raht_tree_max_level: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    (raht_count_launches(1), max_level_init)<<<1, 1, 0, st>>>(max_level);
    if (n > 1) (raht_count_launches(1), max_level_kernel)<<<grid_for(n, 256, 8), 256, 0, st>>>(level, n, max_level);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API size_t raht_order_ragft_workspace_bytes(int64_t n)
{
    if (n < 0) n = 0;
    int64_t nb = (n + OR_TILE - 1) / OR_TILE;
    return align_up((size_t)(nb * OR_BUCKETS + 1) * 8, 256) + 256;
}

RAHT_API int raht_order_ragft(const uint8_t *level, int64_t n, int64_t *order, int32_t *position, void *workspace,
                              size_t workspace_bytes, void *stream)
{
    RAHT_REQUIRE(n >= 0 && (n == 0 || (level && workspace)), "raht_order_ragft: bad arguments");
    if (n == 0) return RAHT_OK;
    if (workspace_bytes < raht_order_ragft_workspace_bytes(n)) {
        raht_set_error("raht_order_ragft: workspace too small");
        return RAHT_E_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    int64_t nb = (n + OR_TILE - 1) / OR_TILE;
    unsigned long long *counts = (unsigned long long *)workspace;
    RAHT_PDL_LAUNCH(PDL_TREE, order_count_kernel, (unsigned)nb, OR_THREADS, 0, st, level, n, nb, counts);
    RAHT_PDL_LAUNCH(PDL_TREE, scan_counts_kernel, 1, 1024, 0, st, counts, nb * OR_BUCKETS, (int64_t *)nullptr);
    RAHT_PDL_LAUNCH(PDL_TREE, order_scatter_kernel, (unsigned)nb, OR_THREADS, 0, st, level, n, nb, (const unsigned long long *)counts, order, position);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

// ---- plan create / destroy -------------------------------------------------------------------------
RAHT_API void raht_plan_destroy(raht_plan *plan) { delete plan; }

RAHT_API size_t raht_plan_workspace_bytes(int64_t n, int tile_rows)
{
    if (n < 0 || tile_rows < 32 || tile_rows > 1024) return 0;
    return plan_layout(n, tile_rows).total + 256;
}

RAHT_API int raht_plan_create(const uint8_t *level, int64_t n, int tile_rows, double *w, void *workspace,
                              size_t workspace_bytes, raht_plan **plan, void *stream)
{
    RAHT_REQUIRE(plan, "raht_plan_create: plan is null");
    *plan = nullptr;
    RAHT_REQUIRE(n >= 0 && n < ((int64_t)1 << 32) - 1, "raht_plan_create: n out of range");
    RAHT_REQUIRE(tile_rows >= 32 && tile_rows <= 1024 && (tile_rows % 2) == 0,
                 "raht_plan_create: tile_rows must be even and in [32, 1024]");
    RAHT_REQUIRE(n == 0 || level, "raht_plan_create: level is null");
    PlanLayout L = plan_layout(n, tile_rows);
    RAHT_REQUIRE(L.d.n_stages < PLAN_MAX_STAGES, "raht_plan_create: too many stages");
    RAHT_REQUIRE((uint64_t)(L.d.slots0 + n) < ((uint64_t)1 << 32), "raht_plan_create: schedule does not fit 32-bit slots");
    if (n > 0 && (!workspace || workspace_bytes < L.total || ((uintptr_t)workspace & 255) != 0)) {
        raht_set_error("raht_plan_create: workspace too small or not 256-byte aligned (%zu < %zu)", workspace_bytes, L.total);
        return RAHT_E_WORKSPACE;
    }
    raht_plan *P = new (std::nothrow) raht_plan();
    RAHT_REQUIRE(P, "raht_plan_create: out of host memory");
    char *ws = (char *)workspace;
    P->d = L.d;
    P->bytes = L.total;
    P->d.hdr = (TileHdr *)(ws + L.off_hdr);
    P->d.def_rows = (uint32_t *)(ws + L.off_def_rows);
    P->d.def_local = (uint16_t *)(ws + L.off_def_local);
    P->d.def_lev = (uint8_t *)(ws + L.off_def_lev);
    P->d.sched = (uint32_t *)(ws + L.off_sched);
    P->d.ab = (double2 *)(ws + L.off_ab);
    P->d.bump = (unsigned int *)(ws + L.off_bump);
    P->d.arrive = (unsigned int *)(ws + L.off_arrive);
    P->d.done = (unsigned int *)(ws + L.off_done);
    P->d.ticket = (unsigned int *)(ws + L.off_bump) + 16;
    if (n > 0) {
        cudaStream_t st = (cudaStream_t)stream;
        RAHT_PDL_LAUNCH(PDL_TREE, plan_init_kernel, (unsigned)((P->d.tiles_total + 255) / 256), 256, 0, st, P->d);
        RAHT_PDL_LAUNCH(PDL_TREE, plan_kernel, (unsigned)P->d.tiles[0], PB_THREADS, 0, st, P->d, level, w);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) {
            raht_set_error("raht_plan_create: launch: %s", cudaGetErrorString(e));
            delete P;
            return RAHT_E_CUDA;
        }
    }
    *plan = P;
    return RAHT_OK;
}

RAHT_API int raht_plan_info(const raht_plan *plan, int64_t *info_host, int cap)
{
    RAHT_REQUIRE(plan && info_host && cap >= 4, "raht_plan_info: bad arguments");
    info_host[0] = plan->d.n;
    info_host[1] = plan->d.tile_rows;
    info_host[2] = plan->d.n_stages;
    info_host[3] = (int64_t)plan->bytes;
    for (int k = 0; k < plan->d.n_stages && 4 + k < cap; ++k) info_host[4 + k] = plan->d.tiles[k];
    return RAHT_OK;
}
