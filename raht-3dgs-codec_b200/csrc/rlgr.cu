// rlgr.cu -- adaptive Run-Length / Golomb-Rice entropy coder.
// Replaces python/PyRLGR/src/libs/rlgr/membuf.cpp:340-423 (rlgrWrite), :258-338 (rlgrRead),
// :228-256 (grWrite/grRead), :166-189 (bit writes), :46-56 (close) as bound by bindings.cpp:34-57
// and driven per channel by python/encode_3dgs.py:229-245.  The bitstream is the reference's, bit
// for bit: MSB-first packing, zero padding to a byte, constants L=4,U0=3,D0=1,U1=2,D1=1
// (membuf.h:18-22), including the tail rule (:416-419) and the 32-bit escape (:242-256).
//
// The reference coder is a sequential state machine per channel.  Here a channel is cut into lane
// blocks of 128 symbols at EVENT boundaries (rlgr_core.cuh: between events the state is (k_P,k_RP));
// 31 consecutive blocks form a warp chunk, solved by one warp:
//   K0  last_nonzero     per block: index of its last non-zero symbol; max-scanned per channel, so every
//                        block knows how many zeros precede it (its first event's run starts there);
//   A   speculate        every chunk in parallel.  Each lane codes its block (lengths only) from two
//                        candidate start states; lane j+1 then restarts from lane j's end states until
//                        nothing changes (exact inside the chunk after <= 32 rounds, 2 in practice).
//                        Lane 0 holds the block BEFORE the chunk as warm-up from canonical states;
//                        later sweeps start from the previous chunk's end states instead;
//   B   resolve          one warp per channel walks the chunk records: the true start state selects
//                        the matching candidate (or the chunk is re-solved exactly), giving each chunk
//                        its start state and bit offset, the channel its byte length and -- optionally
//                        -- the sync index that lets the decoder start at every chunk;
//   C   emit             every lane re-codes its block from its now-known state and bit offset.
// All of it is exact: speculation only decides how much is recomputed.  The decoder is chunk-parallel
// with the sync index and channel-serial without it (streams written by the reference coder).
#include "common.cuh"
#include "rlgr_core.cuh"
#include <stdlib.h>

namespace {

using namespace rlgr;

constexpr int SYNC_STRIDE = 4;                                                // lane blocks per sync entry
constexpr int SYNC_PER_CHUNK = (CHUNK_LANES + SYNC_STRIDE - 1) / SYNC_STRIDE;     // 8 entries per warp chunk
constexpr uint32_t SYNC_NONE = 1023;                                          // no non-zero symbol in the entry's range
constexpr int ROW_PITCH = BLOCK_SYMS + 1;                         // odd pitch: lane-strided reads are conflict-free
constexpr int CHUNK_SMEM_WORDS = 32 * ROW_PITCH + 32 * MASK_WORDS;

struct Geometry {
    int64_t n, ld;
    int64_t nblk;        // lane blocks per channel
    int64_t nchunks;     // warp chunks per channel
    int nch, flag_signed;
};

__host__ __device__ inline int64_t blocks_of(int64_t n) { return (n + BLOCK_SYMS - 1) / BLOCK_SYMS; }
__host__ __device__ inline int64_t chunks_of(int64_t n) { return (blocks_of(n) + CHUNK_LANES - 1) / CHUNK_LANES; }

// ---- K0: last non-zero symbol of every lane block ------------------------------------------------------
constexpr int K0_WARPS = 8;
__global__ void __launch_bounds__(32 * K0_WARPS) rlgr_lastnz_kernel(const int32_t *__restrict__ q, Geometry G,
                                                                   int32_t *__restrict__ lastnz)
{
    const int lane = threadIdx.x & 31;
    const int64_t group = (int64_t)blockIdx.x * K0_WARPS + (threadIdx.x >> 5);      // 32 lane blocks per warp
    const int ch = blockIdx.y;
    const int64_t blk0 = group * 32;
    if (blk0 >= G.nblk) return;
    const int32_t *src = q + (size_t)ch * G.ld;
    int32_t mine = -1;
    for (int b = 0; b < 32; ++b) {
        const int64_t s0 = (blk0 + b) * BLOCK_SYMS;
        if (s0 >= G.n) break;
        int32_t v[MASK_WORDS];
#pragma unroll
        for (int w = 0; w < MASK_WORDS; ++w) {
            const int64_t i = s0 + 32 * w + lane;
            v[w] = i < G.n ? src[i] : 0;
        }
        int32_t last = -1;
#pragma unroll
        for (int w = 0; w < MASK_WORDS; ++w) {
            const unsigned m = __ballot_sync(~0u, v[w] != 0);
            if (m) last = (int32_t)(s0 + 32 * w + 31 - __clz((int)m));
        }
        if (lane == b) mine = last;
    }
    if (blk0 + lane < G.nblk) lastnz[(size_t)ch * G.nblk + blk0 + lane] = mine;
}

// inclusive max-scan of lastnz per channel: afterwards lastnz[b] = last non-zero symbol at or before
// the end of block b (-1: none).  One CTA per channel.
__global__ void __launch_bounds__(1024) rlgr_lastnz_scan_kernel(Geometry G, int32_t *__restrict__ lastnz)
{
    __shared__ int32_t wmax[32];
    __shared__ int32_t carry_s;
    int32_t *a = lastnz + (size_t)blockIdx.x * G.nblk;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = -1;
    __syncthreads();
    for (int64_t base = 0; base < G.nblk; base += 1024) {
        const int64_t i = base + threadIdx.x;
        int32_t x = i < G.nblk ? a[i] : -1;
        for (int o = 1; o < 32; o <<= 1) {
            int32_t y = __shfl_up_sync(~0u, x, o);
            if (lane >= o) x = max(x, y);
        }
        if (lane == 31) wmax[warp] = x;
        __syncthreads();
        int32_t pre = carry_s;
        for (int w = 0; w < warp; ++w) pre = max(pre, wmax[w]);
        x = max(x, pre);
        if (i < G.nblk) a[i] = x;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = x;
        __syncthreads();
    }
}

// ---- one warp chunk in shared memory -------------------------------------------------------------------
struct LaneBlock {
    const uint32_t *row;       // mapped symbols of this lane's block (shared memory)
    const uint32_t *mask;      // MASK_WORDS non-zero masks
    int len;                   // symbols of the block inside the channel (0: none)
    uint32_t zeros_before;     // zeros between the previous non-zero of the channel and the block
    bool is_tail;              // block holds the channel's last symbol
    int last_nz;               // index of the block's last non-zero symbol, -1: none
    bool dense;                // warp-uniform: >= 25 % non-zero -> per-symbol branch-free counting
    bool single;               // warp-uniform: >= 50 % non-zero -> one canonical warm-up state is enough
};

// Lane l of chunk w holds block 31 w - 1 + l (lane 0: the block before the chunk).
__device__ __forceinline__ LaneBlock load_chunk(const int32_t *__restrict__ q, const Geometry &G, int ch, int64_t w,
                                                const int32_t *__restrict__ lastnz_ch, uint32_t *smem)
{
    const int lane = threadIdx.x & 31;
    uint32_t *s_rows = smem;
    uint32_t *s_mask = smem + 32 * ROW_PITCH;
    const int32_t *src = q + (size_t)ch * G.ld;
    const int64_t sym0 = (w * CHUNK_LANES - 1) * BLOCK_SYMS;
    __syncwarp();
    for (int b = 0; b < 32; ++b) {
        const int64_t s0 = sym0 + (int64_t)b * BLOCK_SYMS;
        uint32_t u[MASK_WORDS];
#pragma unroll
        for (int wd = 0; wd < MASK_WORDS; ++wd) {
            const int64_t i = s0 + 32 * wd + lane;
            const int32_t v = (i >= 0 && i < G.n) ? src[i] : 0;
            u[wd] = G.flag_signed ? s2u(v) : (uint32_t)v;
        }
#pragma unroll
        for (int wd = 0; wd < MASK_WORDS; ++wd) {
            s_rows[b * ROW_PITCH + 32 * wd + lane] = u[wd];
            const unsigned m = __ballot_sync(~0u, u[wd] != 0);
            if (lane == wd) s_mask[b * MASK_WORDS + wd] = m;
        }
    }
    __syncwarp();
    LaneBlock B;
    const int64_t blk = w * CHUNK_LANES - 1 + lane;
    const int64_t start = blk * BLOCK_SYMS;
    B.row = s_rows + lane * ROW_PITCH;
    B.mask = s_mask + lane * MASK_WORDS;
    int64_t len = (blk < 0) ? 0 : G.n - start;
    len = len < 0 ? 0 : (len > BLOCK_SYMS ? BLOCK_SYMS : len);
    B.len = (int)len;
    B.zeros_before = (blk <= 0 || len == 0) ? 0u : (uint32_t)(start - 1 - (int64_t)lastnz_ch[blk - 1]);
    B.is_tail = len > 0 && start + len == G.n;
    B.last_nz = -1;
    uint32_t nz = 0;
    for (int wd = 0; wd < MASK_WORDS; ++wd) {
        const uint32_t m = B.mask[wd];
        nz += __popc(m);
        if (m) B.last_nz = 32 * wd + 31 - __clz((int)m);
    }
    for (int o = 16; o; o >>= 1) nz += __shfl_xor_sync(~0u, nz, o);
    B.dense = nz * 4 >= (uint32_t)(CHUNK_LANES * BLOCK_SYMS);   // >= 25 % non-zero: the per-symbol form is cheaper
    B.single = nz * 2 >= (uint32_t)(CHUNK_LANES * BLOCK_SYMS);
    return B;
}

// Exact solution of a chunk given the states at the start of lane 1 (pin == true: lane 0 just forwards
// `first`) or a warm-up of lane 0's block from `first` (pin == false).  On return, for each candidate c:
// in[c] = state at the start of this lane's block, out[c] = state after it, bits[c] = code length of it.
// Blocks without events (and without the tail) leave the state alone, so every lane takes its start
// state from the nearest earlier lane that has events: a correction crosses empty blocks in one round.
// Returns true when no lane of the chunk (1..31) codes anything: the chunk is the identity on the state.
__device__ __forceinline__ bool solve_chunk(const LaneBlock &B, bool pin, const State first[NCAND], State in[NCAND],
                                            State out[NCAND], uint32_t bits[NCAND], int *rounds)
{
    const int lane = threadIdx.x & 31;
    bool has = B.is_tail;
    for (int w = 0; w < MASK_WORDS; ++w) has |= B.mask[w] != 0;
    if (lane == 0) has = !pin && has;                       // a pinned lane 0 only forwards `first`
    const unsigned active = __ballot_sync(~0u, has);
    const unsigned before = active & lanemask_lt();
    const int src = before ? 31 - __clz((int)before) : 0;  // nearest earlier lane that changes the state
    for (int c = 0; c < NCAND; ++c) { in[c] = first[c]; out[c] = first[c]; bits[c] = 0; }
    bool need = has;
    int r = 0;
    for (; r < 34; ++r) {
        if (need) {
            if (B.dense) {
                for (int c = 0; c < NCAND; ++c) out[c] = in[c];
                if (in[1] == in[0]) {
                    count_block_dense<1>(B.row, B.len, B.last_nz, B.zeros_before, B.is_tail, out, bits);
                    out[1] = out[0];
                    bits[1] = bits[0];
                } else {
                    count_block_dense<NCAND>(B.row, B.len, B.last_nz, B.zeros_before, B.is_tail, out, bits);
                }
            } else {
                for (int c = 0; c < NCAND; ++c) {
                    if (c > 0 && in[c] == in[0]) { out[c] = out[0]; bits[c] = bits[0]; continue; }
                    State s = in[c];
                    CountSink k{0};
                    code_block(B.row, B.len, B.mask, B.zeros_before, B.is_tail, s, k);
                    out[c] = s;
                    bits[c] = k.bits;
                }
            }
        }
        need = false;
        for (int c = 0; c < NCAND; ++c) {
            const uint32_t prev = __shfl_sync(~0u, pack_state(out[c]), src);
            if (lane > 0 && prev != pack_state(in[c])) {
                in[c] = unpack_state(prev);
                if (has) need = true;
                else out[c] = in[c];
            }
        }
        if (!__any_sync(~0u, need)) break;
    }
    // lanes without events behind the last correction: forward once more so that in == out == source's out
    for (int c = 0; c < NCAND; ++c) {
        const uint32_t prev = __shfl_sync(~0u, pack_state(out[c]), src);
        if (lane > 0 && !has) { in[c] = unpack_state(prev); out[c] = in[c]; }
    }
    if (lane == 0) { bits[0] = 0; bits[1] = 0; }           // the warm-up block belongs to the previous chunk
    if (rounds) *rounds = r + 1;
    return (active & ~1u) == 0;
}

// ---- records -----------------------------------------------------------------------------------------------
// chunk record (uint4): x = cand0 | cand1 << 16 (assumed states at the chunk start), y = end0 | end1 << 16,
//                       z = bits0, w = bits1 (code length of the chunk under each candidate)
// lane record  (uint2): x = start0 | start1 << 16, y = bits0 | bits1 << 16      (bits of a block < 2^16)
// chunk offset (u64)  : bit offset of the chunk in its stream << 1 | chosen candidate
// sync index   (uint4): x,y = bit offset (lo, hi), z = first symbol coded at that offset, w = packed state
__device__ __forceinline__ uint32_t warp_sum(uint32_t v)
{
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(~0u, v, o);
    return v;
}

constexpr uint32_t REC_IDENTITY = 0x80000000u;     // in rec.z: no lane of the chunk codes anything

__device__ __forceinline__ uint4 finish_chunk(const State in[NCAND], const State out[NCAND], const uint32_t bits[NCAND],
                                              bool identity, uint2 *lanerec)
{
    const int lane = threadIdx.x & 31;
    lanerec[lane] = make_uint2(pack_state(in[0]) | (pack_state(in[1]) << 16), bits[0] | (bits[1] << 16));
    uint4 rec;
    rec.x = __shfl_sync(~0u, pack_state(in[0]) | (pack_state(in[1]) << 16), 1);
    rec.y = __shfl_sync(~0u, pack_state(out[0]) | (pack_state(out[1]) << 16), 31);
    rec.z = warp_sum(bits[0]) | (identity ? REC_IDENTITY : 0u);
    rec.w = warp_sum(bits[1]);
    return rec;
}

// the two canonical start states of a warm-up block: the coder's initial state and a run-mode state of
// the other k_P parity (sparse channels never leave run mode, where differences in k_P keep their parity)
__device__ __forceinline__ void canonical_states(State c[NCAND])
{
    c[0] = initial_state();
    c[1].kP = 9;
    c[1].kRP = RL_KRP_INIT;
}

// ---- A: speculate ----------------------------------------------------------------------------------------
// sweep 0: warm-up from canonical states.  sweep > 0: start from the end states chunk w-1 reported in the
// previous sweep (prev records); a chunk whose assumption did not change just copies its record.
__global__ void __launch_bounds__(32) rlgr_speculate_kernel(const int32_t *__restrict__ q, Geometry G,
                                                            const int32_t *__restrict__ lastnz, int sweep,
                                                            const uint4 *__restrict__ rec_prev, uint4 *__restrict__ rec_cur,
                                                            uint2 *__restrict__ lanerec, unsigned long long *stats)
{
    extern __shared__ uint32_t rl_smem[];
    const int ch = blockIdx.y;
    const int64_t w = blockIdx.x;
    const size_t ci = (size_t)ch * G.nchunks + w;
    State first[NCAND];
    bool pin;
    if (w == 0) {
        first[0] = first[1] = initial_state();
        pin = true;
        if (sweep > 0) { if (threadIdx.x == 0) rec_cur[ci] = rec_prev[ci]; return; }
    } else if (sweep == 0) {
        canonical_states(first);
        pin = false;
    } else {
        const uint4 mine = rec_prev[ci];
        const uint32_t guess = rec_prev[ci - 1].y;
        if (guess == mine.x || (mine.z & REC_IDENTITY)) { if (threadIdx.x == 0) rec_cur[ci] = mine; return; }
        first[0] = unpack_state(guess & 0xffff);
        first[1] = unpack_state(guess >> 16);
        pin = true;
    }
    const LaneBlock B = load_chunk(q, G, ch, w, lastnz + (size_t)ch * G.nblk, rl_smem);
    // mostly non-zero chunks re-synchronise through k_P = 0 whatever the parity: one canonical state is enough
    if (sweep == 0 && B.single) first[1] = first[0];
    State in[NCAND], out[NCAND];
    uint32_t bits[NCAND];
    int rounds;
    const bool identity = solve_chunk(B, pin, first, in, out, bits, &rounds);
    const uint4 rec = finish_chunk(in, out, bits, identity, lanerec + ci * 32);
    if (threadIdx.x == 0) {
        rec_cur[ci] = rec;
        if (stats) { atomicAdd(stats + 0, 1ull); atomicAdd(stats + 1, (unsigned long long)rounds); }
    }
}

// ---- B: resolve ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) rlgr_resolve_kernel(const int32_t *__restrict__ q, Geometry G,
                                                          const int32_t *__restrict__ lastnz, uint4 *__restrict__ rec,
                                                          uint2 *__restrict__ lanerec, unsigned long long *__restrict__ chunkoff,
                                                          int64_t slot_bytes, int64_t *__restrict__ nbytes,
                                                          unsigned long long *__restrict__ totals, int32_t *status,
                                                          unsigned long long *stats)
{
    extern __shared__ uint32_t rl_smem[];
    const int ch = blockIdx.x;
    const int lane = threadIdx.x & 31;
    const int32_t *lastnz_ch = lastnz + (size_t)ch * G.nblk;
    State s = initial_state();
    unsigned long long off = 0;
    unsigned long long repairs = 0;
    for (int64_t base = 0; base < G.nchunks; base += 32) {
        const int64_t w = base + lane;
        const size_t ci = (size_t)ch * G.nchunks + w;
        uint4 r = make_uint4(0, 0, 0, 0);
        if (w < G.nchunks) r = rec[ci];
        unsigned long long my_off = 0;
        const int cnt = (int)min((int64_t)32, G.nchunks - base);
        for (int i = 0; i < cnt; ++i) {
            const uint32_t cand = __shfl_sync(~0u, r.x, i);
            uint32_t ends = __shfl_sync(~0u, r.y, i);
            uint32_t b0 = __shfl_sync(~0u, r.z, i), b1 = __shfl_sync(~0u, r.w, i);
            const uint32_t ps = pack_state(s);
            if (b0 & REC_IDENTITY) {                             // nothing coded in this chunk: state and offset pass through
                if (lane == i) my_off = off << 1;
                continue;
            }
            int b = (ps == (cand & 0xffff)) ? 0 : (ps == (cand >> 16)) ? 1 : -1;
            if (b < 0) {                                          // assumption failed: solve this chunk exactly
                const LaneBlock B = load_chunk(q, G, ch, base + i, lastnz_ch, rl_smem);
                State first[NCAND] = {s, s}, in[NCAND], out[NCAND];
                uint32_t bits[NCAND];
                solve_chunk(B, true, first, in, out, bits, nullptr);
                const uint4 fixed = finish_chunk(in, out, bits, false, lanerec + ((size_t)ch * G.nchunks + base + i) * 32);
                if (lane == i) r = fixed;
                ends = fixed.y; b0 = fixed.z; b1 = fixed.w;
                b = 0;
                ++repairs;
            }
            if (lane == i) my_off = (off << 1) | (unsigned)b;
            off += b ? b1 : b0;
            s = unpack_state(b ? (ends >> 16) : (ends & 0xffff));
        }
        if (w < G.nchunks) chunkoff[ci] = my_off;
    }
    if (lane == 0) {
        const int64_t nb = (int64_t)((off + 7) >> 3);
        nbytes[ch] = nb;
        totals[ch] = off;
        if (nb > slot_bytes && status) atomicOr(status, RAHT_ST_STREAM_OVERFLOW);
        if (stats) atomicAdd(stats + 2, repairs);
    }
}

// zero the words the emit pass will OR into: ceil(bits / 64) 8-byte words per channel, capped to the slot
__global__ void __launch_bounds__(256) rlgr_zero_kernel(uint8_t *__restrict__ streams, int64_t slot_bytes,
                                                        const unsigned long long *__restrict__ totals)
{
    const int ch = blockIdx.y;
    unsigned long long words = (totals[ch] + 63) >> 6;
    const unsigned long long cap = (unsigned long long)slot_bytes >> 3;
    if (words > cap) words = cap;
    unsigned long long *dst = (unsigned long long *)(streams + (size_t)ch * slot_bytes);
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < words;
         i += (unsigned long long)gridDim.x * blockDim.x)
        dst[i] = 0ull;
}

// ---- C: emit ---------------------------------------------------------------------------------------------
// MSB-first bit writer into pre-zeroed big-endian 32-bit words.  A lane owns bits [begin, end) of the
// stream; words it shares with its neighbours are OR-ed atomically, the others stored.
struct BitSink {
    uint32_t *words;
    int64_t cap_words;
    int64_t wi;           // index of the word being assembled
    uint64_t acc;
    unsigned nb;          // bits in acc (the first word starts with begin % 32 zero bits that are not ours)
    bool shared_word;     // the word being assembled has bits of the previous lane
    bool overflow;
    __device__ __forceinline__ void init(uint8_t *stream, int64_t slot_bytes, unsigned long long begin_bit)
    {
        words = (uint32_t *)stream;
        cap_words = slot_bytes >> 2;
        wi = (int64_t)(begin_bit >> 5);
        nb = (unsigned)(begin_bit & 31);
        acc = 0;
        shared_word = nb != 0;
        overflow = false;
    }
    __device__ __forceinline__ void flush_word(uint32_t w, bool shared)
    {
        if (wi < cap_words) {
            const uint32_t be = __byte_perm(w, 0, 0x0123);
            if (shared) { if (be) atomicOr(words + wi, be); }
            else words[wi] = be;
        } else overflow = true;
        ++wi;
    }
    __device__ __forceinline__ void put(uint64_t value, unsigned len)      // len <= 32
    {
        acc = (acc << len) | value;
        nb += len;
        if (nb >= 32) {
            flush_word((uint32_t)(acc >> (nb - 32)), shared_word);
            shared_word = false;
            nb -= 32;
        }
    }
    __device__ __forceinline__ void finish()
    {
        if (nb) flush_word((uint32_t)(acc << (32 - nb)), true);           // last word: shared with the next lane
    }
};

__global__ void __launch_bounds__(32) rlgr_emit_kernel(const int32_t *__restrict__ q, Geometry G,
                                                       const int32_t *__restrict__ lastnz, const uint2 *__restrict__ lanerec,
                                                       const unsigned long long *__restrict__ chunkoff,
                                                       uint8_t *__restrict__ streams, int64_t slot_bytes,
                                                       uint2 *__restrict__ sync_index, int32_t *status)
{
    extern __shared__ uint32_t rl_smem[];
    const int ch = blockIdx.y;
    const int64_t w = blockIdx.x;
    const int lane = threadIdx.x & 31;
    const size_t ci = (size_t)ch * G.nchunks + w;
    const LaneBlock B = load_chunk(q, G, ch, w, lastnz + (size_t)ch * G.nblk, rl_smem);
    const unsigned long long co = chunkoff[ci];
    const int b = (int)(co & 1);
    const uint2 lr = lanerec[ci * 32 + lane];
    State s = unpack_state(b ? (lr.x >> 16) : (lr.x & 0xffff));
    const uint32_t mybits = lane == 0 ? 0u : (b ? (lr.y >> 16) : (lr.y & 0xffff));
    uint32_t x = mybits;                                   // exclusive prefix over the lanes
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(~0u, x, o);
        if (lane >= o) x += y;
    }
    const unsigned long long begin = (co >> 1) + (x - mybits);
    if (sync_index) {
        // sync entry of lanes 1+4k..4+4k: where (bit offset, state) the coder stands before the range's first
        // event and where that event's non-zero symbol sits.  Lanes without events have no bits and pass the
        // state through, so the leader lane's (begin, s) are those of the first lane with an event.
        uint32_t fn = SYNC_NONE;
        for (int wd = MASK_WORDS - 1; wd >= 0; --wd)
            if (B.mask[wd]) fn = (uint32_t)(32 * wd + __ffs((int)B.mask[wd]) - 1);
        uint32_t first = fn;
        for (int o = 1; o < SYNC_STRIDE; ++o) {                   // nearest lane with an event wins
            const uint32_t f = __shfl_down_sync(~0u, fn, o);
            if (first == SYNC_NONE && lane + o < 32 && f != SYNC_NONE) first = (uint32_t)(o * BLOCK_SYMS) + f;
        }
        if (lane >= 1 && ((lane - 1) % SYNC_STRIDE) == 0) {
            const int k = (lane - 1) / SYNC_STRIDE;
            const uint32_t hi = (uint32_t)((begin >> 32) & 0x7f) | ((uint32_t)(s.kP & 0x7f) << 7) | ((uint32_t)(s.kRP & 0xff) << 14) |
                                (first << 22);
            sync_index[ci * SYNC_PER_CHUNK + k] = make_uint2((uint32_t)begin, hi);
        }
    }
    if (lane == 0 || mybits == 0) return;
    BitSink sink;
    sink.init(streams + (size_t)ch * slot_bytes, slot_bytes, begin);
    code_block(B.row, B.len, B.mask, B.zeros_before, B.is_tail, s, sink);
    sink.finish();
    if (sink.overflow && status) atomicOr(status, RAHT_ST_STREAM_OVERFLOW);
}

// ---- decoder (BitReader and decode_range live in rlgr_core.cuh) ---------------------------------------------
// without a sync index: one thread per channel
__global__ void __launch_bounds__(32) rlgr_decode_serial_kernel(const uint8_t *__restrict__ streams, int64_t slot_bytes,
                                                               const int64_t *__restrict__ nbytes, Geometry G,
                                                               int32_t *__restrict__ q)
{
    if (threadIdx.x != 0) return;
    const int ch = blockIdx.x;
    int64_t nb = nbytes ? nbytes[ch] : slot_bytes;
    if (nb > slot_bytes) nb = slot_bytes;
    BitReader br;
    br.init(streams + (size_t)ch * slot_bytes, nb, 0);
    decode_range(br, initial_state(), -1, G.n, G.flag_signed, q + (size_t)ch * G.ld);
}

// with the encoder's sync index: one thread per entry (SYNC_STRIDE lane blocks)
constexpr int DEC_THREADS = 128;
__global__ void __launch_bounds__(DEC_THREADS) rlgr_decode_indexed_kernel(const uint8_t *__restrict__ streams, int64_t slot_bytes,
                                                                         const int64_t *__restrict__ nbytes,
                                                                         const uint2 *__restrict__ sync_index, Geometry G,
                                                                         int32_t *__restrict__ q)
{
    const int ch = blockIdx.y;
    const int64_t e = (int64_t)blockIdx.x * DEC_THREADS + threadIdx.x;
    const int64_t entries = G.nchunks * SYNC_PER_CHUNK;
    if (e >= entries) return;
    const int64_t w = e / SYNC_PER_CHUNK;
    const int k = (int)(e - w * SYNC_PER_CHUNK);
    const int64_t start = (w * CHUNK_LANES + (int64_t)k * SYNC_STRIDE) * BLOCK_SYMS;
    if (start >= G.n) return;
    const int nblocks = min(SYNC_STRIDE, CHUNK_LANES - k * SYNC_STRIDE);
    int64_t end = start + (int64_t)nblocks * BLOCK_SYMS;
    if (end > G.n) end = G.n;
    const uint2 ent = sync_index[(size_t)ch * entries + e];
    const uint32_t first = ent.y >> 22;
    if (first == SYNC_NONE || start + first >= end) return;                // nothing but zeros in this range
    State s;
    s.kP = (int)((ent.y >> 7) & 0x7f);
    s.kRP = (int)((ent.y >> 14) & 0xff);
    int64_t nb = nbytes ? nbytes[ch] : slot_bytes;
    if (nb > slot_bytes) nb = slot_bytes;
    BitReader br;
    br.init(streams + (size_t)ch * slot_bytes, nb, (unsigned long long)ent.x | ((unsigned long long)(ent.y & 0x7f) << 32));
    decode_range(br, s, start + first, end, G.flag_signed, q + (size_t)ch * G.ld);
}

struct EncLayout {
    size_t off_lastnz, off_recA, off_recB, off_lanerec, off_chunkoff, off_totals, off_stats, total;
};
EncLayout enc_layout(int64_t n, int nch)
{
    EncLayout L;
    const size_t nblk = (size_t)blocks_of(n), nchunks = (size_t)chunks_of(n), c = (size_t)(nch > 0 ? nch : 0);
    size_t o = 0;
    L.off_lastnz = o; o += align_up(c * nblk * 4, 256);
    L.off_recA = o; o += align_up(c * nchunks * 16, 256);
    L.off_recB = o; o += align_up(c * nchunks * 16, 256);
    L.off_lanerec = o; o += align_up(c * nchunks * 32 * 8, 256);
    L.off_chunkoff = o; o += align_up(c * nchunks * 8, 256);
    L.off_totals = o; o += align_up(c * 8, 256);
    L.off_stats = o; o += 256;
    L.total = o + 256;
    return L;
}

}  // namespace

RAHT_API int64_t raht_rlgr_sync_entries(int64_t n) { return n <= 0 ? 0 : chunks_of(n) * SYNC_PER_CHUNK; }

RAHT_API size_t raht_rlgr_encode_workspace_bytes(int64_t n, int nch) { return enc_layout(n < 0 ? 0 : n, nch).total; }

RAHT_API int raht_rlgr_encode(const int32_t *q, int64_t n, int64_t ld, int nch, int flag_signed, uint8_t *streams,
                              int64_t slot_bytes, int64_t *nbytes, void *sync_index, int32_t *status, void *workspace,
                              size_t workspace_bytes, void *stream)
{
    RAHT_REQUIRE(n >= 0 && n < ((int64_t)1 << 31) - 256 && nch >= 0 && nch < 65536 && ld >= n && slot_bytes >= 8 &&
                     (slot_bytes % 8) == 0,
                 "raht_rlgr_encode: bad sizes (slot_bytes must be a positive multiple of 8)");
    if (nch == 0) return RAHT_OK;
    RAHT_REQUIRE((q || n == 0) && streams && nbytes, "raht_rlgr_encode: null pointer");
    RAHT_REQUIRE(((uintptr_t)streams & 7) == 0, "raht_rlgr_encode: streams must be 8-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) {                                            // nothing written: zero-length streams
        RAHT_CUDA(cudaMemsetAsync(nbytes, 0, (size_t)nch * 8, st));
        return RAHT_OK;
    }
    const EncLayout L = enc_layout(n, nch);
    if (!workspace || workspace_bytes < L.total) {
        raht_set_error("raht_rlgr_encode: workspace too small (%zu < %zu)", workspace_bytes, L.total);
        return RAHT_E_WORKSPACE;
    }
    Geometry G;
    G.n = n; G.ld = ld; G.nblk = blocks_of(n); G.nchunks = chunks_of(n); G.nch = nch; G.flag_signed = flag_signed;
    RAHT_REQUIRE(G.nchunks < ((int64_t)1 << 31), "raht_rlgr_encode: too many chunks");
    char *ws = (char *)workspace;
    int32_t *lastnz = (int32_t *)(ws + L.off_lastnz);
    uint4 *recA = (uint4 *)(ws + L.off_recA), *recB = (uint4 *)(ws + L.off_recB);
    uint2 *lanerec = (uint2 *)(ws + L.off_lanerec);
    unsigned long long *chunkoff = (unsigned long long *)(ws + L.off_chunkoff);
    unsigned long long *totals = (unsigned long long *)(ws + L.off_totals);
    unsigned long long *stats = (unsigned long long *)(ws + L.off_stats);
    const size_t smem = (size_t)CHUNK_SMEM_WORDS * 4;
    RAHT_CUDA(cudaMemsetAsync(stats, 0, 256, st));

    const unsigned groups = (unsigned)((G.nblk + 32 * K0_WARPS - 1) / (32 * K0_WARPS));
    (raht_count_launches(1), rlgr_lastnz_kernel)<<<dim3(groups, nch), 32 * K0_WARPS, 0, st>>>(q, G, lastnz);
    (raht_count_launches(1), rlgr_lastnz_scan_kernel)<<<nch, 1024, 0, st>>>(G, lastnz);
    const dim3 cgrid((unsigned)G.nchunks, nch);
    // speculation sweeps: the first from canonical states, later ones from the previous chunk's end states
    static int sweeps = -1;
    if (sweeps < 0) {
        const char *e = getenv("RAHT_RLGR_SWEEPS");
        sweeps = e ? atoi(e) : 2;
        if (sweeps < 1) sweeps = 1;
        if (sweeps > 8) sweeps = 8;
    }
    uint4 *rec_cur = recA, *rec_prev = recB;
    for (int sw = 0; sw < sweeps; ++sw) {
        (raht_count_launches(1), rlgr_speculate_kernel)<<<cgrid, 32, smem, st>>>(q, G, lastnz, sw, rec_prev, rec_cur, lanerec, stats);
        uint4 *t = rec_cur; rec_cur = rec_prev; rec_prev = t;
    }
    recB = rec_prev;                                          // records of the last sweep
    (raht_count_launches(1), rlgr_resolve_kernel)<<<nch, 32, smem, st>>>(q, G, lastnz, recB, lanerec, chunkoff, slot_bytes, nbytes,
                                                                       totals, status, stats);
    (raht_count_launches(1), rlgr_zero_kernel)<<<dim3(64, nch), 256, 0, st>>>(streams, slot_bytes, totals);
    (raht_count_launches(1), rlgr_emit_kernel)<<<cgrid, 32, smem, st>>>(q, G, lastnz, lanerec, chunkoff, streams, slot_bytes,
                                                                      (uint2 *)sync_index, status);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

// diagnostics of the last raht_rlgr_encode on this workspace: out[0] = chunk solves in the speculation
// sweeps, out[1] = their fixed-point rounds, out[2] = chunks re-solved by the resolve pass.  Synchronises.
RAHT_API int raht_rlgr_encode_stats(const void *workspace, int64_t n, int nch, unsigned long long *out3_host, void *stream)
{
    RAHT_REQUIRE(workspace && out3_host && n > 0, "raht_rlgr_encode_stats: bad arguments");
    const EncLayout L = enc_layout(n, nch);
    RAHT_CUDA(cudaMemcpyAsync(out3_host, (const char *)workspace + L.off_stats, 24, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    RAHT_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return RAHT_OK;
}

RAHT_API int raht_rlgr_decode(const uint8_t *streams, int64_t slot_bytes, const int64_t *nbytes, const void *sync_index,
                              int64_t n, int64_t ld, int nch, int flag_signed, int32_t *q, void *stream)
{
    RAHT_REQUIRE(n >= 0 && n < ((int64_t)1 << 31) - 256 && nch >= 0 && nch < 65536 && ld >= n && slot_bytes >= 8 &&
                     (slot_bytes % 8) == 0,
                 "raht_rlgr_decode: bad sizes (slot_bytes must be a positive multiple of 8)");
    if (nch == 0 || n == 0) return RAHT_OK;
    RAHT_REQUIRE(streams && q, "raht_rlgr_decode: null pointer");
    RAHT_REQUIRE(((uintptr_t)streams & 7) == 0, "raht_rlgr_decode: streams must be 8-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    Geometry G;
    G.n = n; G.ld = ld; G.nblk = blocks_of(n); G.nchunks = chunks_of(n); G.nch = nch; G.flag_signed = flag_signed;
    RAHT_CUDA(cudaMemsetAsync(q, 0, ((size_t)(nch - 1) * ld + n) * 4, st));      // the decoders store non-zeros only
    if (sync_index) {
        const dim3 grid((unsigned)((G.nchunks * SYNC_PER_CHUNK + DEC_THREADS - 1) / DEC_THREADS), nch);
        (raht_count_launches(1), rlgr_decode_indexed_kernel)<<<grid, DEC_THREADS, 0, st>>>(streams, slot_bytes, nbytes,
                                                                                          (const uint2 *)sync_index, G, q);
    } else {
        (raht_count_launches(1), rlgr_decode_serial_kernel)<<<nch, 32, 0, st>>>(streams, slot_bytes, nbytes, G, q);
    }
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}
