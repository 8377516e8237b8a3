// rlgr.cu -- adaptive Run-Length / Golomb-Rice entropy coder.
// Replaces python/PyRLGR/src/libs/rlgr/membuf.cpp:340-423 (rlgrWrite), :258-338 (rlgrRead),
// :228-256 (grWrite/grRead), :166-189 (bit writes), :46-56 (close) as bound by bindings.cpp:34-57
// and driven per channel by python/encode_3dgs.py:229-245.  The bitstream is the reference's, bit
// for bit: MSB-first packing, zero padding to a byte, constants L=4,U0=3,D0=1,U1=2,D1=1
// (membuf.h:18-22), including the tail rule (:416-419) and the 32-bit escape (:242-256).
//
// This file holds the channel-serial coder: one warp per channel, the adaptive state (k_P, k_RP, m)
// is carried symbol to symbol by one lane.  Channels are independent streams, so nch warps run
// concurrently on nch SMs.
#include "common.cuh"

namespace {

constexpr unsigned RL_L = 4, RL_U0 = 3, RL_D0 = 1, RL_U1 = 2, RL_D1 = 1;
constexpr unsigned RL_KRP_MAX = 32 * RL_L;

__device__ __forceinline__ uint32_t s2u32(int32_t v)      // membuf.cpp:4-14 (_s2u)
{
    return v < 0 ? (((uint32_t)(-(int64_t)v)) << 1) - 1u : ((uint32_t)v) << 1;
}
__device__ __forceinline__ int32_t u2s32(uint64_t u)
{
    int64_t d = (int64_t)(u >> 1);
    return (int32_t)((u & 1) ? -d - 1 : d);
}

struct BitWriter {
    uint32_t *out;        // big-endian 32-bit words
    int64_t cap_words;
    int64_t pos;
    uint64_t acc;
    unsigned nbits;       // < 32 between calls
    bool overflow;
    __device__ __forceinline__ void put(uint64_t value, unsigned len)      // len <= 32
    {
        acc = (acc << len) | value;
        nbits += len;
        if (nbits >= 32) {
            uint32_t w = (uint32_t)(acc >> (nbits - 32));
            if (pos < cap_words) out[pos] = __byte_perm(w, 0, 0x0123);
            else overflow = true;
            ++pos;
            nbits -= 32;
        }
    }
    __device__ __forceinline__ void gr(uint64_t u, unsigned kr)            // membuf.cpp:242-256 (grWrite)
    {
        uint64_t p = u >> kr;
        if (p < 32) {
            put((((uint64_t)1) << (p + 1)) - 2, (unsigned)p + 1);          // p ones and a zero
            if (kr) put(u & ((((uint64_t)1) << kr) - 1), kr);
        } else {
            put(0xffffffffull, 32);
            put(u & 0xffffffffull, 32);
        }
    }
};

__device__ __forceinline__ void adapt_krp(unsigned &k_RP, uint64_t u, unsigned k_R)
{
    uint64_t p = u >> k_R;
    if (p) { uint64_t t = (uint64_t)k_RP + p - 1; k_RP = t > RL_KRP_MAX ? RL_KRP_MAX : (unsigned)t; }
    else k_RP = k_RP < 2 ? 0 : k_RP - 2;
}

__global__ void __launch_bounds__(32) rlgr_encode_serial_kernel(const int32_t *__restrict__ q, int64_t n, int64_t ld,
                                                               int flag_signed, uint8_t *streams, int64_t slot_bytes,
                                                               int64_t *nbytes, int32_t *status)
{
    if (threadIdx.x != 0) return;
    const int ch = blockIdx.x;
    const int32_t *src = q + (size_t)ch * ld;
    BitWriter bw;
    bw.out = (uint32_t *)(streams + (size_t)ch * slot_bytes);
    bw.cap_words = slot_bytes / 4;
    bw.pos = 0; bw.acc = 0; bw.nbits = 0; bw.overflow = false;
    if (n == 0) { nbytes[ch] = 0; return; }

    unsigned k_P = 0, k_RP = 2 * RL_L, k = 0;
    uint64_t m = 0, u = 0;
    for (int64_t i = 0; i < n; ++i) {
        int32_t v = src[i];
        u = flag_signed ? (uint64_t)s2u32(v) : (uint64_t)(uint32_t)v;
        k = k_P / RL_L;
        const unsigned k_R = k_RP / RL_L;
        if (k) {
            if (u) {                                                       // membuf.cpp:353-381
                u--;
                bw.put(0, 1);
                bw.put(m, k);
                bw.gr(u, k_R);
                adapt_krp(k_RP, u, k_R);
                k_P = k_P < RL_D1 ? 0 : k_P - RL_D1;
                m = 0;
            } else {                                                       // :382-390
                m++;
                if (m == (((uint64_t)1) << k)) { bw.put(1, 1); k_P += RL_U1; m = 0; }
            }
        } else {                                                           // :391-413
            bw.gr(u, k_R);
            adapt_krp(k_RP, u, k_R);
            if (u) k_P = k_P < RL_D0 ? 0 : k_P - RL_D0;
            else k_P += RL_U0;
            m = 0;
        }
    }
    if (k && !u) { bw.put(0, 1); bw.put(m, k_P / RL_L); }                  // :416-419 (tail)
    // close(): pad with zero bits to a byte boundary (:46-56)
    int64_t total = bw.pos * 4 + (bw.nbits + 7) / 8;
    if (bw.nbits) {
        uint32_t w = (uint32_t)(bw.acc << (32 - bw.nbits));
        if (bw.pos < bw.cap_words) bw.out[bw.pos] = __byte_perm(w, 0, 0x0123);
        else bw.overflow = true;
    }
    nbytes[ch] = total;
    if (bw.overflow && status) atomicOr(status, RAHT_ST_STREAM_OVERFLOW);
}

struct BitReader {
    const uint32_t *in;
    int64_t n_words;      // readable words of the slot
    int64_t pos;
    uint64_t acc;
    unsigned nbits;
    __device__ __forceinline__ void refill()               // call when nbits <= 32
    {
        uint32_t w = pos < n_words ? __byte_perm(in[pos], 0, 0x0123) : 0u;
        ++pos;
        acc = (acc << 32) | w;
        nbits += 32;
    }
    __device__ __forceinline__ unsigned get1()
    {
        if (nbits == 0) refill();
        --nbits;
        return (unsigned)(acc >> nbits) & 1u;
    }
    __device__ __forceinline__ uint64_t get(unsigned len)  // len <= 32
    {
        if (len == 0) return 0;
        if (nbits < len) refill();
        nbits -= len;
        return (acc >> nbits) & ((((uint64_t)1) << len) - 1);
    }
    __device__ __forceinline__ uint64_t gr(unsigned kr)    // membuf.cpp:228-240 (grRead)
    {
        unsigned p = 0;
        while (get1()) {
            if (++p >= 32) return get(32);
        }
        return (((uint64_t)p) << kr) + get(kr);
    }
};

__global__ void __launch_bounds__(32) rlgr_decode_serial_kernel(const uint8_t *__restrict__ streams, int64_t slot_bytes,
                                                               const int64_t *__restrict__ nbytes, int64_t n, int64_t ld,
                                                               int flag_signed, int32_t *__restrict__ q)
{
    if (threadIdx.x != 0) return;
    const int ch = blockIdx.x;
    int32_t *dst = q + (size_t)ch * ld;
    BitReader br;
    br.in = (const uint32_t *)(streams + (size_t)ch * slot_bytes);
    int64_t nb = nbytes ? nbytes[ch] : slot_bytes;
    br.n_words = (min(nb, slot_bytes) + 3) / 4;
    br.pos = 0; br.acc = 0; br.nbits = 0;

    unsigned k_P = 0, k_RP = 2 * RL_L;
    int64_t i = 0;
    while (i < n) {                                                        // membuf.cpp:266-334
        unsigned k = k_P / RL_L;
        const unsigned k_R = k_RP / RL_L;
        if (k) {
            uint64_t m = 0;
            while (br.get1()) {                                            // complete runs
                m += ((uint64_t)1) << k;
                k_P += RL_U1;
                k = k_P / RL_L;
                if (m >= (uint64_t)(n - i)) break;                         // corrupt stream guard
            }
            m += br.get(k);
            int64_t stop = (m > (uint64_t)(n - i)) ? n : i + (int64_t)m;
            for (; i < stop; ++i) dst[i] = 0;
            if (i >= n) break;
            uint64_t u = br.gr(k_R);
            dst[i++] = flag_signed ? u2s32(u + 1) : (int32_t)(u + 1);
            adapt_krp(k_RP, u, k_R);
            k_P = k_P < RL_D1 ? 0 : k_P - RL_D1;
        } else {
            uint64_t u = br.gr(k_R);
            dst[i++] = flag_signed ? u2s32(u) : (int32_t)u;
            adapt_krp(k_RP, u, k_R);
            if (u) k_P = k_P < RL_D0 ? 0 : k_P - RL_D0;
            else k_P += RL_U0;
        }
    }
}

}  // namespace

RAHT_API size_t raht_rlgr_encode_workspace_bytes(int64_t n, int nch)
{
    (void)n; (void)nch;
    return 256;
}

RAHT_API int raht_rlgr_encode(const int32_t *q, int64_t n, int64_t ld, int nch, int flag_signed, uint8_t *streams,
                              int64_t slot_bytes, int64_t *nbytes, int32_t *status, void *workspace,
                              size_t workspace_bytes, void *stream)
{
    (void)workspace; (void)workspace_bytes;
    RAHT_REQUIRE(n >= 0 && nch >= 0 && ld >= n && slot_bytes >= 8 && (slot_bytes % 8) == 0,
                 "raht_rlgr_encode: bad sizes (slot_bytes must be a positive multiple of 8)");
    if (nch == 0) return RAHT_OK;
    RAHT_REQUIRE((q || n == 0) && streams && nbytes, "raht_rlgr_encode: null pointer");
    RAHT_REQUIRE(((uintptr_t)streams & 7) == 0, "raht_rlgr_encode: streams must be 8-byte aligned");
    (raht_count_launches(1), rlgr_encode_serial_kernel)<<<nch, 32, 0, (cudaStream_t)stream>>>(q, n, ld, flag_signed, streams, slot_bytes, nbytes, status);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

RAHT_API int raht_rlgr_decode(const uint8_t *streams, int64_t slot_bytes, const int64_t *nbytes, int64_t n, int64_t ld,
                              int nch, int flag_signed, int32_t *q, void *stream)
{
    RAHT_REQUIRE(n >= 0 && nch >= 0 && ld >= n && slot_bytes >= 8 && (slot_bytes % 8) == 0,
                 "raht_rlgr_decode: bad sizes (slot_bytes must be a positive multiple of 8)");
    if (nch == 0 || n == 0) return RAHT_OK;
    RAHT_REQUIRE(streams && q, "raht_rlgr_decode: null pointer");
    RAHT_REQUIRE(((uintptr_t)streams & 7) == 0, "raht_rlgr_decode: streams must be 8-byte aligned");
    (raht_count_launches(1), rlgr_decode_serial_kernel)<<<nch, 32, 0, (cudaStream_t)stream>>>(streams, slot_bytes, nbytes, n, ld, flag_signed, q);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}
