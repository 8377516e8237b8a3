// radix_sort.cu -- hand-written onesweep LSD radix sort (64-bit keys, 32-bit payloads), stable.
// Replaces torch.sort at python/voxelize_pc.py:101 (M_sort, idx = torch.sort(M)).  Ties keep the
// original order (the reference's tie order is unspecified).
//
// One histogram launch for all digits, then one launch per 8-bit digit: every CTA ranks a tile of
// keys in shared memory (warp-level match_any multisplit), publishes its per-digit counts and
// resolves its global offsets by decoupled look-back over the previous tiles of the same launch
// (tile ids come from an atomic ticket, so a tile only ever waits for tiles that already run).
// Traffic per pass: read 12 B + write 12 B per element -- the compulsory minimum for LSD.
#include "common.cuh"

namespace {

constexpr int RS_BITS = 8;
constexpr int RS_RADIX = 1 << RS_BITS;
constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_IPT = 12;
constexpr int RS_TILE = RS_THREADS * RS_IPT;   // 3072 keys per CTA
constexpr int RS_MAX_PASSES = 8;

constexpr uint32_t FLAG_AGG = 1u << 30;
constexpr uint32_t FLAG_INC = 2u << 30;
constexpr uint32_t FLAG_MASK = 3u << 30;
constexpr uint32_t VAL_MASK = ~FLAG_MASK;

// ---- all-digit histogram -------------------------------------------------------------------------
__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(const uint64_t *__restrict__ keys, int64_t n, int begin_bit,
                                                             int passes, int end_bit, unsigned long long *ghist)
{
    __shared__ unsigned int sh[RS_MAX_PASSES * RS_RADIX];
    pdl_trigger();
    for (int i = threadIdx.x; i < passes * RS_RADIX; i += RS_THREADS) sh[i] = 0;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * RS_THREADS + threadIdx.x; i < n; i += (int64_t)gridDim.x * RS_THREADS) {
        uint64_t k = keys[i];
        for (int p = 0; p < passes; ++p) {
            int sh_bit = begin_bit + p * RS_BITS;
            int nb = min(RS_BITS, end_bit - sh_bit);
            unsigned d = (unsigned)(k >> sh_bit) & ((1u << nb) - 1u);
            atomicAdd(&sh[p * RS_RADIX + d], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * RS_RADIX; i += RS_THREADS)
        if (sh[i]) atomicAdd(&ghist[i], (unsigned long long)sh[i]);
}

// exclusive scan of each pass's 256 bins, in place (one CTA per pass, 256 threads)
__global__ void __launch_bounds__(RS_RADIX) rs_scan_kernel(unsigned long long *ghist)
{
    __shared__ unsigned long long wt[RS_RADIX / 32];
    pdl_enter();
    unsigned long long *h = ghist + (size_t)blockIdx.x * RS_RADIX;
    unsigned long long v = h[threadIdx.x], x = v;
    for (int o = 1; o < 32; o <<= 1) {
        unsigned long long y = __shfl_up_sync(~0u, x, o);
        if ((int)lane_id() >= o) x += y;
    }
    if (lane_id() == 31) wt[threadIdx.x >> 5] = x;
    __syncthreads();
    unsigned long long base = 0;
    for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) base += wt[w];
    h[threadIdx.x] = base + x - v;
}

// ---- one onesweep pass ---------------------------------------------------------------------------
__global__ void __launch_bounds__(RS_THREADS)
rs_onesweep_kernel(const uint64_t *__restrict__ keys_in, uint64_t *__restrict__ keys_out,
                   const uint32_t *__restrict__ vals_in, uint32_t *__restrict__ vals_out, int64_t n, int shift,
                   unsigned digit_mask, const unsigned long long *__restrict__ ghist_excl,   // 256 bins of this pass
                   volatile uint32_t *tile_state,                                            // n_tiles x 256, zeroed
                   unsigned int *ticket, int32_t *status)
{
    pdl_enter();
    __shared__ uint64_t s_keys[RS_TILE];
    __shared__ uint32_t s_vals[RS_TILE];
    __shared__ uint32_t s_whist[RS_WARPS][RS_RADIX];
    __shared__ uint32_t s_digit_start[RS_RADIX];        // first local sorted position of each digit
    __shared__ long long s_global_base[RS_RADIX];       // global position of local sorted position 0, per digit
    __shared__ uint32_t s_wtot[RS_WARPS];
    __shared__ unsigned int s_tile;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    for (int i = tid; i < RS_WARPS * RS_RADIX; i += RS_THREADS) (&s_whist[0][0])[i] = 0;
    __syncthreads();
    const unsigned tile = s_tile;
    const int64_t tile_base = (int64_t)tile * RS_TILE;
    const int n_valid = (int)min((int64_t)RS_TILE, n - tile_base);

    // load: warp-striped so that (warp, j, lane) order == memory order (needed for stability)
    uint64_t key[RS_IPT];
    uint32_t val[RS_IPT];
    uint32_t rank[RS_IPT];
#pragma unroll
    for (int j = 0; j < RS_IPT; ++j) {
        int loc = warp * (32 * RS_IPT) + j * 32 + lane;
        int64_t g = tile_base + loc;
        if (loc < n_valid) {
            key[j] = keys_in[g];
            val[j] = vals_in ? vals_in[g] : (uint32_t)g;
        } else {
            key[j] = ~0ull;      // pads sort to the very end of the tile and are never written
            val[j] = 0;
        }
    }
    // rank inside the warp, in (j, lane) order
#pragma unroll
    for (int j = 0; j < RS_IPT; ++j) {
        unsigned d = (unsigned)(key[j] >> shift) & digit_mask;
        if (warp * (32 * RS_IPT) + j * 32 + lane >= n_valid) d = RS_RADIX - 1;
        unsigned peers = __match_any_sync(~0u, d);
        int leader = __ffs(peers) - 1;
        unsigned before = __popc(peers & lanemask_lt());
        uint32_t old = 0;
        if (lane == leader) {
            old = s_whist[warp][d];
            s_whist[warp][d] = old + __popc(peers);
        }
        old = __shfl_sync(~0u, old, leader);
        rank[j] = old + before;
        __syncwarp();
    }
    __syncthreads();

    // per digit (thread == digit): exclusive prefix over warps, tile total
    uint32_t total;
    {
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) {
            uint32_t c = s_whist[w][tid];
            s_whist[w][tid] = run;
            run += c;
        }
        total = run;
    }
    // publish the aggregate early so that later tiles can make progress
    volatile uint32_t *my_state = tile_state + (size_t)tile * RS_RADIX;
    if (tile == 0) my_state[tid] = FLAG_INC | total;
    else my_state[tid] = FLAG_AGG | total;

    // exclusive scan of the totals over digits -> s_digit_start
    {
        uint32_t x = total;
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(~0u, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) s_wtot[warp] = x;
        __syncthreads();
        uint32_t base = 0;
        for (int w = 0; w < warp; ++w) base += s_wtot[w];
        s_digit_start[tid] = base + x - total;
    }

    // decoupled look-back for digit `tid`
    {
        unsigned long long excl = 0;
        if (tile > 0) {
            int t = (int)tile - 1;
            unsigned spins = 0;
            while (true) {
                uint32_t s = tile_state[(size_t)t * RS_RADIX + tid];
                uint32_t f = s & FLAG_MASK;
                if (f == 0) {
                    if (++spins > (1u << 26)) {      // never expected; refuse to hang the GPU
                        if (status) atomicOr(status, RAHT_ST_SORT_STALL);
                        break;
                    }
                    __nanosleep(20);
                    continue;
                }
                spins = 0;
                excl += (s & VAL_MASK);
                if (f == FLAG_INC) break;
                --t;
            }
            my_state[tid] = FLAG_INC | (uint32_t)((excl + total) & VAL_MASK);
        }
        s_global_base[tid] = (long long)(ghist_excl[tid] + excl) - (long long)s_digit_start[tid];
    }
    __syncthreads();

    // local sorted position of every key; scatter into shared memory
    uint32_t pos[RS_IPT];
#pragma unroll
    for (int j = 0; j < RS_IPT; ++j) {
        unsigned d = (unsigned)(key[j] >> shift) & digit_mask;
        if (warp * (32 * RS_IPT) + j * 32 + lane >= n_valid) d = RS_RADIX - 1;
        pos[j] = s_digit_start[d] + s_whist[warp][d] + rank[j];
        s_keys[pos[j]] = key[j];
        s_vals[pos[j]] = val[j];
    }
    __syncthreads();
    // coalesced write-out: consecutive local positions of one digit are consecutive in global memory
    for (int i = tid; i < n_valid; i += RS_THREADS) {
        uint64_t k = s_keys[i];
        unsigned d = (unsigned)(k >> shift) & digit_mask;
        long long g = s_global_base[d] + i;
        keys_out[g] = k;
        vals_out[g] = s_vals[i];
    }
}

struct RsLayout {
    size_t off_ghist, off_ticket, off_state, off_keys, off_vals, total;
    int passes;
    int64_t tiles;
};

RsLayout rs_layout(int64_t n, int begin_bit, int end_bit)
{
    RsLayout L;
    L.passes = (end_bit - begin_bit + RS_BITS - 1) / RS_BITS;
    if (L.passes < 1) L.passes = 1;
    L.tiles = (n + RS_TILE - 1) / RS_TILE;
    size_t o = 0;
    L.off_ghist = o; o += align_up((size_t)RS_MAX_PASSES * RS_RADIX * 8, 256);
    L.off_ticket = o; o += 256;
    L.off_state = o; o += align_up((size_t)L.passes * (size_t)L.tiles * RS_RADIX * 4, 256);
    L.off_keys = o; o += align_up((size_t)n * 8, 256);
    L.off_vals = o; o += align_up((size_t)n * 4, 256);
    L.total = o + 256;
    return L;
}

}  // namespace

RAHT_API size_t raht_radix_sort_workspace_bytes(int64_t n, int begin_bit, int end_bit)
{
    if (n < 0) n = 0;
    return rs_layout(n, begin_bit, end_bit).total;
}

RAHT_API int raht_radix_sort_pairs(const uint64_t *keys_in, uint64_t *keys_out, const uint32_t *vals_in,
                                   uint32_t *vals_out, int64_t n, int begin_bit, int end_bit, int32_t *status,
                                   void *workspace, size_t workspace_bytes, void *stream)
{
    RAHT_REQUIRE(n >= 0 && n < ((int64_t)1 << 30), "raht_radix_sort_pairs: n must be in [0, 2^30)");
    RAHT_REQUIRE(begin_bit >= 0 && end_bit <= 64 && begin_bit < end_bit && end_bit - begin_bit <= RS_MAX_PASSES * RS_BITS,
                 "raht_radix_sort_pairs: bad bit range");
    if (n == 0) return RAHT_OK;
    RAHT_REQUIRE(keys_in && keys_out && vals_out && workspace, "raht_radix_sort_pairs: null pointer");
    RsLayout L = rs_layout(n, begin_bit, end_bit);
    if (workspace_bytes < L.total) {
        raht_set_error("raht_radix_sort_pairs: workspace too small (%zu < %zu)", workspace_bytes, L.total);
        return RAHT_E_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    char *ws = (char *)workspace;
    unsigned long long *ghist = (unsigned long long *)(ws + L.off_ghist);
    unsigned int *ticket = (unsigned int *)(ws + L.off_ticket);
    uint32_t *state = (uint32_t *)(ws + L.off_state);
    uint64_t *tmp_keys = (uint64_t *)(ws + L.off_keys);
    uint32_t *tmp_vals = (uint32_t *)(ws + L.off_vals);

    // zero histograms, tickets (one per pass) and tile states in one memset
    RAHT_CUDA(cudaMemsetAsync(ws, 0, L.off_keys, st));
    (raht_count_launches(1), rs_hist_kernel)<<<grid_for(n, RS_THREADS * 8, 4), RS_THREADS, 0, st>>>(keys_in, n, begin_bit, L.passes, end_bit, ghist);
    RAHT_PDL_LAUNCH(PDL_FRONT, rs_scan_kernel, L.passes, RS_RADIX, 0, st, ghist);
    RAHT_LAUNCH_CHECK();

    // ping-pong so that the last pass lands in keys_out / vals_out
    const uint64_t *src_k = keys_in;
    const uint32_t *src_v = vals_in;
    for (int p = 0; p < L.passes; ++p) {
        bool to_out = ((L.passes - 1 - p) % 2) == 0;
        uint64_t *dst_k = to_out ? keys_out : tmp_keys;
        uint32_t *dst_v = to_out ? vals_out : tmp_vals;
        int sh = begin_bit + p * RS_BITS;
        int nb = end_bit - sh < RS_BITS ? end_bit - sh : RS_BITS;
        RAHT_PDL_LAUNCH(PDL_FRONT, rs_onesweep_kernel, (unsigned)L.tiles, RS_THREADS, 0, st, src_k, dst_k, src_v, dst_v, n, sh, (1u << nb) - 1u,
                        (const unsigned long long *)(ghist + (size_t)p * RS_RADIX),
                        (volatile uint32_t *)(state + (size_t)p * (size_t)L.tiles * RS_RADIX), ticket + p, status);
        RAHT_LAUNCH_CHECK();
        src_k = dst_k;
        src_v = dst_v;
    }
    return RAHT_OK;
}
