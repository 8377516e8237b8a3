// container.cu -- device side of the compressed-scene container (the reference never
// serialises its bitstreams, python/encode_3dgs.py:247 only takes len(buf), and ships the voxel positions
// uncompressed in a PLY, python/test_voxelize_3dgs.py:284-288).
//
//  * geometry symbols: the Morton codes of the occupied voxels are sorted and unique, so they are coded as
//    d[0] = key[0], d[i] = key[i] - key[i-1] - 1 >= 0, every d split into three 21-bit digits (3 * 21 = 63 bits
//    covers every depth the Morton code allows, voxelize_pc.py:46-57) that go through the same RLGR coder as the
//    attributes, unsigned (flagSigned = 0): the high digits are zero runs, the low digit is a Golomb-Rice coded gap.
//  * keys back from the symbols: inclusive prefix sum of (d[i] + (i > 0)) (three-phase scan, 64-bit).
//  * stream packing: the nch RLGR streams of an encoder call (one slot per channel) laid end to end, every stream
//    padded with zero bytes to a multiple of 8, and the inverse.
#include "common.cuh"

namespace {

constexpr int GEO_DIGIT_BITS = 21;
constexpr int GEO_DIGITS = 3;
constexpr uint32_t GEO_MASK = (1u << GEO_DIGIT_BITS) - 1;
constexpr int SCAN_THREADS = 256, SCAN_ITEMS = 8, SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__global__ void __launch_bounds__(256) geometry_symbols_kernel(const int64_t *__restrict__ keys, int64_t n, int64_t ld,
                                                               int32_t *__restrict__ sym, int32_t *status)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const int64_t k = keys[i];
        uint64_t d;
        if (i == 0) {
            if (k < 0 && status) atomicOr(status, RAHT_ST_OUT_OF_BOUNDS);
            d = (uint64_t)k;
        } else {
            const int64_t p = keys[i - 1];
            if (!(p < k)) {
                if (status) atomicOr(status, RAHT_ST_NOT_SORTED_UNIQUE);
                d = 0;
            } else {
                d = (uint64_t)(k - p) - 1;
            }
        }
        sym[i] = (int32_t)(d & GEO_MASK);
        sym[ld + i] = (int32_t)((d >> GEO_DIGIT_BITS) & GEO_MASK);
        sym[2 * ld + i] = (int32_t)((d >> (2 * GEO_DIGIT_BITS)) & GEO_MASK);
    }
}

__device__ __forceinline__ uint64_t geo_step(const int32_t *__restrict__ sym, int64_t ld, int64_t i)
{
    const uint64_t d = (uint64_t)(uint32_t)sym[i] | ((uint64_t)(uint32_t)sym[ld + i] << GEO_DIGIT_BITS) |
                       ((uint64_t)(uint32_t)sym[2 * ld + i] << (2 * GEO_DIGIT_BITS));
    return d + (i > 0 ? 1 : 0);
}

__device__ __forceinline__ uint64_t block_sum_u64(uint64_t v, uint64_t *smem /*[32]*/)
{
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(~0u, v, o);
    if (lane_id() == 0) smem[threadIdx.x >> 5] = v;
    __syncthreads();
    uint64_t t = threadIdx.x < (blockDim.x >> 5) ? smem[threadIdx.x] : 0;
    if (threadIdx.x < 32)
        for (int o = 16; o; o >>= 1) t += __shfl_xor_sync(~0u, t, o);
    return t;                                                               // valid in warp 0
}

__global__ void __launch_bounds__(SCAN_THREADS) geometry_tile_sums_kernel(const int32_t *__restrict__ sym, int64_t n, int64_t ld,
                                                                         unsigned long long *__restrict__ sums)
{
    __shared__ uint64_t sh[32];
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
    uint64_t v = 0;
    for (int j = 0; j < SCAN_ITEMS; ++j) {
        const int64_t i = base + j * SCAN_THREADS + threadIdx.x;
        if (i < n) v += geo_step(sym, ld, i);
    }
    const uint64_t t = block_sum_u64(v, sh);
    if (threadIdx.x == 0) sums[blockIdx.x] = t;
}

// thread t of a tile owns items [t * SCAN_ITEMS, (t + 1) * SCAN_ITEMS): the loads of a warp cover one contiguous kilobyte
__global__ void __launch_bounds__(SCAN_THREADS) geometry_keys_kernel(const int32_t *__restrict__ sym, int64_t n, int64_t ld,
                                                                    const unsigned long long *__restrict__ tile_offsets,
                                                                    int64_t *__restrict__ keys)
{
    __shared__ uint64_t wsum[SCAN_THREADS / 32];
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    uint64_t v[SCAN_ITEMS];
    uint64_t run = 0;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; ++j) {
        const int64_t i = base + j;
        run += i < n ? geo_step(sym, ld, i) : 0;
        v[j] = run;
    }
    uint64_t x = run;                                                       // inclusive scan of the thread totals
    for (int o = 1; o < 32; o <<= 1) {
        const uint64_t y = __shfl_up_sync(~0u, x, o);
        if ((int)lane_id() >= o) x += y;
    }
    if (lane_id() == 31) wsum[threadIdx.x >> 5] = x;
    __syncthreads();
    uint64_t before = tile_offsets[blockIdx.x] + (x - run);
    for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) before += wsum[w];
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; ++j) {
        const int64_t i = base + j;
        if (i < n) keys[i] = (int64_t)(before + v[j]);
    }
}

// offsets[c] = sum over j < c of align8(nbytes[j]); offsets[nch] = total.  One block.
__global__ void __launch_bounds__(1024) pack_offsets_kernel(const int64_t *__restrict__ nbytes, int nch, int64_t slot_bytes,
                                                            int64_t *__restrict__ offsets, int64_t capacity, int32_t *status)
{
    __shared__ unsigned long long wtot[32];
    __shared__ unsigned long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nch; base += 1024) {
        const int i = base + threadIdx.x;
        unsigned long long v = 0;
        if (i < nch) {
            int64_t b = nbytes[i];
            if (b < 0) b = 0;
            if (b > slot_bytes) b = slot_bytes;
            v = (unsigned long long)((b + 7) & ~(int64_t)7);
        }
        unsigned long long x = v;
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long y = __shfl_up_sync(~0u, x, o);
            if ((int)lane_id() >= o) x += y;
        }
        if (lane_id() == 31) wtot[threadIdx.x >> 5] = x;
        __syncthreads();
        unsigned long long before = carry + (x - v);
        for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) before += wtot[w];
        if (i < nch) offsets[i] = (int64_t)before;
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        offsets[nch] = (int64_t)carry;
        if ((int64_t)carry > capacity && status) atomicOr(status, RAHT_ST_STREAM_OVERFLOW);
    }
}

// 8-byte words; the bytes of the last word past the stream's length are written as zeros
__global__ void __launch_bounds__(256) pack_copy_kernel(const uint8_t *__restrict__ streams, int64_t slot_bytes,
                                                        const int64_t *__restrict__ nbytes, const int64_t *__restrict__ offsets,
                                                        uint8_t *__restrict__ out, int64_t capacity)
{
    const int ch = blockIdx.y;
    int64_t nb = nbytes[ch];
    if (nb < 0) nb = 0;
    if (nb > slot_bytes) nb = slot_bytes;
    const int64_t off = offsets[ch];
    const int64_t words = (nb + 7) >> 3;
    if (off + words * 8 > capacity) return;
    const unsigned long long *src = (const unsigned long long *)(streams + (size_t)ch * slot_bytes);
    unsigned long long *dst = (unsigned long long *)(out + off);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += stride) {
        unsigned long long v = src[w];
        const int64_t left = nb - w * 8;
        if (left < 8) v &= (1ull << (8 * left)) - 1;                        // little-endian: the low bytes come first
        dst[w] = v;
    }
}

__global__ void __launch_bounds__(256) unpack_copy_kernel(const uint8_t *__restrict__ packed, const int64_t *__restrict__ offsets,
                                                          uint8_t *__restrict__ streams, int64_t slot_bytes)
{
    const int ch = blockIdx.y;
    const int64_t off = offsets[ch];
    int64_t bytes = offsets[ch + 1] - off;
    if (bytes > slot_bytes) bytes = slot_bytes & ~(int64_t)7;
    const int64_t words = bytes >> 3;
    const unsigned long long *src = (const unsigned long long *)(packed + off);
    unsigned long long *dst = (unsigned long long *)(streams + (size_t)ch * slot_bytes);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += stride) dst[w] = src[w];
}

}  // namespace

RAHT_API int raht_geometry_symbols(const int64_t *keys, int64_t n, int64_t ld, int32_t *symbols, int32_t *status, void *stream)
{
    RAHT_REQUIRE(n >= 0 && ld >= n, "raht_geometry_symbols: bad sizes");
    if (n == 0) return RAHT_OK;
    RAHT_REQUIRE(keys && symbols, "raht_geometry_symbols: null pointer");
    raht_count_launches(1);
    geometry_symbols_kernel<<<grid_for(n, 256, 8), 256, 0, (cudaStream_t)stream>>>(keys, n, ld, symbols, status);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API size_t raht_geometry_keys_workspace_bytes(int64_t n)
{
    const int64_t tiles = n <= 0 ? 0 : (n + SCAN_TILE - 1) / SCAN_TILE;
    return align_up((size_t)(tiles + 2) * 8, 256);
}

RAHT_API int raht_geometry_keys(const int32_t *symbols, int64_t n, int64_t ld, int64_t *keys, void *workspace,
                                size_t workspace_bytes, void *stream)
{
    RAHT_REQUIRE(n >= 0 && ld >= n, "raht_geometry_keys: bad sizes");
    if (n == 0) return RAHT_OK;
    RAHT_REQUIRE(symbols && keys, "raht_geometry_keys: null pointer");
    if (!workspace || workspace_bytes < raht_geometry_keys_workspace_bytes(n)) {
        raht_set_error("raht_geometry_keys: workspace too small");
        return RAHT_E_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    unsigned long long *sums = (unsigned long long *)workspace;
    raht_count_launches(3);
    geometry_tile_sums_kernel<<<(unsigned)tiles, SCAN_THREADS, 0, st>>>(symbols, n, ld, sums);
    scan_counts_kernel<<<1, 1024, 0, st>>>(sums, tiles, nullptr);
    geometry_keys_kernel<<<(unsigned)tiles, SCAN_THREADS, 0, st>>>(symbols, n, ld, sums, keys);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_streams_pack(const uint8_t *streams, int64_t slot_bytes, const int64_t *nbytes, int nch, uint8_t *out,
                               int64_t out_capacity, int64_t *offsets, int32_t *status, void *stream)
{
    RAHT_REQUIRE(nch >= 0 && nch < 65536 && slot_bytes >= 8 && (slot_bytes % 8) == 0 && out_capacity >= 0,
                 "raht_streams_pack: bad sizes (slot_bytes must be a positive multiple of 8)");
    RAHT_REQUIRE(offsets, "raht_streams_pack: null offsets");
    cudaStream_t st = (cudaStream_t)stream;
    if (nch == 0) {
        RAHT_CUDA(cudaMemsetAsync(offsets, 0, 8, st));
        return RAHT_OK;
    }
    RAHT_REQUIRE(streams && nbytes && out, "raht_streams_pack: null pointer");
    RAHT_REQUIRE((((uintptr_t)streams | (uintptr_t)out) & 7) == 0, "raht_streams_pack: buffers must be 8-byte aligned");
    raht_count_launches(2);
    pack_offsets_kernel<<<1, 1024, 0, st>>>(nbytes, nch, slot_bytes, offsets, out_capacity, status);
    int gx = (int)((slot_bytes / 8 + 255) / 256);
    const int cap = (sm_count_cached() * 8 + nch - 1) / nch;
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    pack_copy_kernel<<<dim3(gx, nch), 256, 0, st>>>(streams, slot_bytes, nbytes, offsets, out, out_capacity);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_streams_unpack(const uint8_t *packed, const int64_t *offsets, int nch, uint8_t *streams, int64_t slot_bytes,
                                 void *stream)
{
    RAHT_REQUIRE(nch >= 0 && nch < 65536 && slot_bytes >= 8 && (slot_bytes % 8) == 0,
                 "raht_streams_unpack: bad sizes (slot_bytes must be a positive multiple of 8)");
    if (nch == 0) return RAHT_OK;
    RAHT_REQUIRE(packed && offsets && streams, "raht_streams_unpack: null pointer");
    RAHT_REQUIRE((((uintptr_t)streams | (uintptr_t)packed) & 7) == 0, "raht_streams_unpack: buffers must be 8-byte aligned");
    int gx = (int)((slot_bytes / 8 + 255) / 256);
    const int cap = (sm_count_cached() * 8 + nch - 1) / nch;
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    raht_count_launches(1);
    unpack_copy_kernel<<<dim3(gx, nch), 256, 0, (cudaStream_t)stream>>>(packed, offsets, streams, slot_bytes);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}
