// rlgr.cu -- adaptive Run-Length / Golomb-Rice entropy coder.
// Replaces python/PyRLGR/src/libs/rlgr/membuf.cpp:340-423 (rlgrWrite), :258-338 (rlgrRead),
// :228-256 (grWrite/grRead), :166-189 (bit writes), :46-56 (close) as bound by bindings.cpp:34-57
// and driven per channel by python/encode_3dgs.py:229-245.  The bitstream is the reference's, bit
// for bit: MSB-first packing, zero padding to a byte, constants L=4,U0=3,D0=1,U1=2,D1=1
// (membuf.h:18-22), including the tail rule (:416-419) and the 32-bit escape (:242-256).
//
// The reference coder is a sequential state machine per channel.  Here every channel is first
// compacted to its EVENT list (rlgr_core.cuh: event = zero run + the non-zero symbol that ends it;
// between events the state is (k_P,k_RP)), stored lane-interleaved so that the 32 lanes of a warp
// stream 32 consecutive blocks of 128 events with fully coalesced loads and no shared memory:
//   K1  count / scan / scatter   non-zeros per 128-symbol block, exclusive scan per channel, then
//                        every non-zero symbol writes (gap, u) to its slot of the event store;
//   A   speculate        every chunk (31 blocks + the block before it as warm-up) in parallel: each
//                        lane codes its block (lengths only) from two canonical start states; lane
//                        j+1 then restarts from lane j's end states until nothing changes (exact
//                        inside the chunk after <= 32 rounds, ~2 in practice);
//   B   resolve          one warp per channel walks the chunk records: the true start state selects
//                        the matching candidate (or the chunk is re-solved exactly), giving each chunk
//                        its start state and bit offset and the channel its byte length;
//   C   emit             every lane re-codes its block from its now-known state at its now-known bit
//                        offset, and writes the sync index (one entry per 256 symbols) on the way.
// All of it is exact: speculation only decides how much is recomputed.  Work, warm-up length and the
// number of chunks are proportional to the number of EVENTS, so zero runs cost nothing and sparse
// channels (which never re-synchronise) have few chunks to repair.  The decoder is range-parallel
// with the sync index and channel-serial without it (streams written by the reference coder).
#include "common.cuh"
#include "rlgr_core.cuh"
#include <stdlib.h>

namespace {

using namespace rlgr;

struct Geometry {
    int64_t n, ld;
    int64_t nsb;         // 128-symbol blocks per channel (compaction passes)
    int64_t W;           // chunk slots per channel in the event store (enough for n events)
    int nch, flag_signed;
    const int32_t *q;    // the symbols (dense channels are coded straight from them)
    const int32_t *lastnz;   // after the scan pass: last non-zero symbol at or before the end of every 128-symbol block
    int vec;             // q rows are 16-byte aligned: whole units of 8 symbols are read as two 16-byte loads
};

__host__ __device__ inline int64_t sym_blocks_of(int64_t n) { return (n + SYM_BLOCK - 1) / SYM_BLOCK; }
__host__ __device__ inline int64_t ev_blocks_of(int64_t events) { return events > 0 ? (events + EV_BLOCK - 1) / EV_BLOCK : 1; }
__host__ __device__ inline int64_t chunks_of_blocks(int64_t nb) { return (nb + CHUNK_LANES - 1) / CHUNK_LANES; }
__host__ __device__ inline int64_t chunk_slots_of(int64_t n) { return chunks_of_blocks(ev_blocks_of(n)); }

// per-channel header written by the scan pass
struct ChanInfo {
    uint32_t events;     // non-zero symbols of the channel
    uint32_t nblocks;    // lane blocks: blocks of 128 events, or (dense) of 128 symbols (>= 1: the block of an empty channel codes the tail)
    uint32_t nchunks;
    uint32_t tail_zeros; // zero symbols after the last event
    uint32_t dense;      // at least half of the symbols are non-zero: coded in SYMBOL space, straight from q (no event store)
    uint32_t tail_block; // dense: the symbol block that holds the last non-zero symbol
    uint32_t pad0, pad1;
};

// ---- K1a: non-zero count and last non-zero symbol of every 128-symbol block -----------------------------
constexpr int K1_WARPS = 8;
__global__ void __launch_bounds__(32 * K1_WARPS) rlgr_count_kernel(const int32_t *__restrict__ q, Geometry G,
                                                                  uint32_t *__restrict__ cnt, int32_t *__restrict__ lastnz)
{
    pdl_trigger();
    const int lane = threadIdx.x & 31;
    const int64_t group = (int64_t)blockIdx.x * K1_WARPS + (threadIdx.x >> 5);      // 32 symbol blocks per warp
    const int ch = blockIdx.y;
    const int64_t blk0 = group * 32;
    if (blk0 >= G.nsb) return;
    const int32_t *src = q + (size_t)ch * G.ld;
    int32_t mine_last = -1;
    uint32_t mine_cnt = 0;
#pragma unroll 4
    for (int b = 0; b < 32; ++b) {
        const int64_t s0 = (blk0 + b) * SYM_BLOCK;
        if (s0 >= G.n) break;
        int32_t v[4];
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            const int64_t i = s0 + 32 * w + lane;
            v[w] = i < G.n ? src[i] : 0;
        }
        int32_t last = -1;
        uint32_t c = 0;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            const unsigned m = __ballot_sync(~0u, v[w] != 0);
            c += __popc(m);
            if (m) last = (int32_t)(s0 + 32 * w + 31 - __clz((int)m));
        }
        if (lane == b) { mine_last = last; mine_cnt = c; }
    }
    if (blk0 + lane < G.nsb) {
        lastnz[(size_t)ch * G.nsb + blk0 + lane] = mine_last;
        cnt[(size_t)ch * G.nsb + blk0 + lane] = mine_cnt;
    }
}

// ---- K1b: per channel, exclusive sum-scan of cnt (-> index of the block's first event) and inclusive
// max-scan of lastnz (-> last non-zero symbol at or before the end of the block).  One CTA per channel.
__global__ void __launch_bounds__(1024) rlgr_scan_kernel(Geometry G, uint32_t *__restrict__ cnt, int32_t *__restrict__ lastnz,
                                                         ChanInfo *__restrict__ info)
{
    pdl_enter();
    __shared__ int32_t wmax[32];
    __shared__ uint32_t wsum[32];
    __shared__ int32_t carry_max;
    __shared__ uint32_t carry_sum;
    int32_t *a = lastnz + (size_t)blockIdx.x * G.nsb;
    uint32_t *c = cnt + (size_t)blockIdx.x * G.nsb;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { carry_max = -1; carry_sum = 0; }
    __syncthreads();
    for (int64_t base = 0; base < G.nsb; base += 1024) {
        const int64_t i = base + threadIdx.x;
        int32_t x = i < G.nsb ? a[i] : -1;
        const uint32_t own = i < G.nsb ? c[i] : 0u;
        uint32_t s = own;
        for (int o = 1; o < 32; o <<= 1) {
            const int32_t y = __shfl_up_sync(~0u, x, o);
            const uint32_t t = __shfl_up_sync(~0u, s, o);
            if (lane >= o) { x = max(x, y); s += t; }
        }
        if (lane == 31) { wmax[warp] = x; wsum[warp] = s; }
        __syncthreads();
        int32_t pre = carry_max;
        uint32_t pres = carry_sum;
        for (int w = 0; w < warp; ++w) { pre = max(pre, wmax[w]); pres += wsum[w]; }
        x = max(x, pre);
        s += pres;
        if (i < G.nsb) { a[i] = x; c[i] = s - own; }
        __syncthreads();
        if (threadIdx.x == 1023) { carry_max = x; carry_sum = s; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        ChanInfo I;
        I.events = carry_sum;
        I.dense = (int64_t)carry_sum * 2 >= G.n ? 1u : 0u;
        I.nblocks = I.dense ? (uint32_t)G.nsb : (uint32_t)ev_blocks_of(carry_sum);
        I.nchunks = (uint32_t)chunks_of_blocks(I.nblocks);
        I.tail_zeros = (uint32_t)(G.n - 1 - (int64_t)carry_max);
        I.tail_block = carry_max < 0 ? 0u : (uint32_t)(carry_max / SYM_BLOCK);
        I.pad0 = I.pad1 = 0u;
        info[blockIdx.x] = I;
    }
}

// ---- K1c: scatter every non-zero symbol to its slot(s) of the event store ----------------------------------
// store: uint2 (gap, u) [channel][chunk slot w][unit j = i / 4][lane][i % 4]; block b = 31 w - 1 + lane, so the
// last block of a chunk is written twice (lane 31 of chunk w and lane 0, the warm-up, of chunk w + 1).
// blockpos[channel][b] = symbol index of the block's first event.
__device__ __forceinline__ void store_event(uint2 *__restrict__ evc, uint32_t *__restrict__ bpos, int64_t W, uint32_t e,
                                            uint32_t gap, uint32_t u, uint32_t pos)
{
    const uint32_t b = e / EV_BLOCK;
    const int i = (int)(e % EV_BLOCK);
    const uint32_t w = b / CHUNK_LANES;
    const int l = (int)(b - w * CHUNK_LANES) + 1;
    evc[(size_t)w * CHUNK_SLOTS + ev_slot(l, i)] = make_uint2(gap, u);
    if (l == CHUNK_LANES && (int64_t)w + 1 < W) evc[(size_t)(w + 1) * CHUNK_SLOTS + ev_slot(0, i)] = make_uint2(gap, u);
    if (i == 0) bpos[b] = pos;
}

__global__ void __launch_bounds__(32 * K1_WARPS) rlgr_scatter_kernel(const int32_t *__restrict__ q, Geometry G,
                                                                    const uint32_t *__restrict__ evbase,
                                                                    const int32_t *__restrict__ lastnz, uint2 *__restrict__ ev,
                                                                    uint32_t *__restrict__ blockpos, const ChanInfo *__restrict__ info)
{
    pdl_enter();
    const int lane = threadIdx.x & 31;
    const int64_t group = (int64_t)blockIdx.x * K1_WARPS + (threadIdx.x >> 5);
    const int ch = blockIdx.y;
    const int64_t blk0 = group * 32;
    if (blk0 >= G.nsb || info[ch].dense) return;            // dense channels are coded from q itself
    const int32_t *src = q + (size_t)ch * G.ld;
    uint2 *evc = ev + (size_t)ch * G.W * CHUNK_SLOTS;
    uint32_t *bpos = blockpos + (size_t)ch * G.nsb;
    uint32_t e = evbase[(size_t)ch * G.nsb + blk0];
    int64_t last = blk0 ? (int64_t)lastnz[(size_t)ch * G.nsb + blk0 - 1] : -1;          // warp-uniform
    // SC_BATCH blocks (4 KB of symbols per warp) are requested before the first is used: only the sparse channels come here,
    // a few hundred warps on the whole GPU, and with two blocks in flight the pass was a chain of 16 round trips (44 us for
    // 3 channels of 1 M symbols)
    constexpr int SC_BATCH = 8;
    for (int b0 = 0; b0 < 32; b0 += SC_BATCH) {
        if ((blk0 + b0) * SYM_BLOCK >= G.n) break;
        int32_t v[SC_BATCH][4];
#pragma unroll
        for (int t = 0; t < SC_BATCH; ++t) {
            const int64_t s0 = (blk0 + b0 + t) * SYM_BLOCK;
#pragma unroll
            for (int w = 0; w < 4; ++w) {
                const int64_t i = s0 + 32 * w + lane;
                v[t][w] = i < G.n ? __ldg(src + i) : 0;
            }
        }
#pragma unroll
        for (int t = 0; t < SC_BATCH; ++t) {
            const int64_t s0 = (blk0 + b0 + t) * SYM_BLOCK;
#pragma unroll
            for (int w = 0; w < 4; ++w) {
                const unsigned m = __ballot_sync(~0u, v[t][w] != 0);
                if (!m) continue;
                const int64_t wbase = s0 + 32 * w;
                if (v[t][w] != 0) {
                    const unsigned before = m & lanemask_lt();
                    const int64_t prev = before ? wbase + 31 - __clz((int)before) : last;
                    const int64_t pos = wbase + lane;
                    const uint32_t u = G.flag_signed ? s2u(v[t][w]) : (uint32_t)v[t][w];
                    store_event(evc, bpos, G.W, e + __popc(before), (uint32_t)(pos - prev - 1), u, (uint32_t)pos);
                }
                e += __popc(m);
                last = wbase + 31 - __clz((int)m);
            }
        }
    }
}

// ---- one lane block of the event store ---------------------------------------------------------------------
struct LaneBlock {
    const uint4 *base;         // the chunk in the store, as 16-byte units: unit j of store lane `slot` at base[(j * 32 + slot) * 2 + {0,1}]
    int slot;                  // store lane of the block (= the thread's lane, except in the table pass where a warp shares one block)
    int cnt;                   // events of this lane's block (0: none); dense channels: its SYMBOLS
    bool is_tail;              // the block holds the channel's last event (or the channel has none): it also codes the tail
    uint32_t tail_zeros;
    // dense channels (symbol space): the block is 128 consecutive symbols of q, its events are found on the fly
    const int32_t *sym;        // nullptr: event space
    uint32_t gap0;             // zero symbols pending in front of the block (since the last non-zero symbol before it)
    int64_t pos0;              // index of the block's first symbol
    int vec, flag_signed;
    int avail;                 // symbols of the channel from the block's start (<= 128): what may be read
};

// Lane l of chunk w holds block 31 w - 1 + l (lane 0: the block before the chunk).
__device__ __forceinline__ LaneBlock lane_block(const uint2 *__restrict__ ev, const Geometry &G, int ch, int64_t w,
                                                const ChanInfo &I)
{
    const int lane = threadIdx.x & 31;
    LaneBlock B;
    B.base = (const uint4 *)(ev + ((size_t)ch * G.W + w) * CHUNK_SLOTS);
    B.slot = lane;
    const int64_t b = w * CHUNK_LANES - 1 + lane;
    B.tail_zeros = I.tail_zeros;
    B.sym = nullptr;
    B.gap0 = 0;
    B.pos0 = 0;
    B.vec = G.vec; B.flag_signed = G.flag_signed; B.avail = 0;
    if (I.dense) {                                          // symbol space: blocks up to the one with the last non-zero symbol
        const bool live = b >= 0 && b <= (int64_t)I.tail_block;
        B.pos0 = b * SYM_BLOCK;
        B.is_tail = b == (int64_t)I.tail_block;
        B.sym = G.q + (size_t)ch * G.ld + (live ? B.pos0 : 0);
        B.cnt = 0;
        B.avail = 0;
        if (live) {                                         // the block ends with its last non-zero symbol (an event boundary)
            const int32_t *ln = G.lastnz + (size_t)ch * G.nsb + b;
            const int64_t last = ln[0];
            const int64_t prev = b > 0 ? (int64_t)ln[-1] : -1;
            B.cnt = last >= B.pos0 ? (int)(last - B.pos0 + 1) : 0;
            B.gap0 = (uint32_t)(B.pos0 - 1 - prev);
            const int64_t av = G.n - B.pos0;
            B.avail = (int)(av > SYM_BLOCK ? SYM_BLOCK : av);
        }
        return B;
    }
    int64_t cnt = (b < 0 || b >= (int64_t)I.nblocks) ? 0 : (int64_t)I.events - b * EV_BLOCK;
    cnt = cnt < 0 ? 0 : (cnt > EV_BLOCK ? EV_BLOCK : cnt);
    B.cnt = (int)cnt;
    B.is_tail = b == (int64_t)I.nblocks - 1;
    return B;
}

// ---- dense channels: a lane block in symbol space -------------------------------------------------------------
// Every lane walks its own block of q -- the symbols from the block's start to its last non-zero symbol -- in units of
// 8 symbols (two 16-byte loads = one 32-byte sector, so DRAM traffic is exactly the symbols), two units requested
// ahead of the one being coded, and feeds every symbol to the per-symbol machine of rlgr_core.cuh: all lanes of the
// warp do the same work per symbol.  f(index in block, mapped symbol); whole units run unguarded.
template <class F> __device__ __forceinline__ void walk_symbols(const LaneBlock &B, F &&f)
{
    const int n = B.cnt;
    const int32_t *p = B.sym;
    const int units = (n + 7) >> 3;
    const int vunits = B.vec ? (B.avail >> 3) : 0;           // units that may be read as two 16-byte loads
    const uint4 *v = (const uint4 *)p;
    // _s2u (membuf.cpp:4-14) as the zig-zag (x << 1) ^ (x >> 31); unsigned symbols pass through
    const int sh = B.flag_signed ? 1 : 0;
    const uint32_t sm = B.flag_signed ? 0xffffffffu : 0u;
    uint4 a1, b1, a2, b2;
    a1 = b1 = a2 = b2 = make_uint4(0u, 0u, 0u, 0u);
    if (0 < units && 0 < vunits) { a1 = __ldg(v); b1 = __ldg(v + 1); }
    if (1 < units && 1 < vunits) { a2 = __ldg(v + 2); b2 = __ldg(v + 3); }
    for (int j = 0; j < units; ++j) {
        uint4 a = a1, b = b1;
        a1 = a2; b1 = b2;
        if (j + 2 < units && j + 2 < vunits) { a2 = __ldg(v + 2 * (j + 2)); b2 = __ldg(v + 2 * (j + 2) + 1); }
        if (j >= vunits) {                                   // the channel's last, partial unit, or rows that are not 16-byte aligned
            int32_t y[8];
#pragma unroll
            for (int t = 0; t < 8; ++t) y[t] = j * 8 + t < B.avail ? __ldg(p + j * 8 + t) : 0;
            a = make_uint4((uint32_t)y[0], (uint32_t)y[1], (uint32_t)y[2], (uint32_t)y[3]);
            b = make_uint4((uint32_t)y[4], (uint32_t)y[5], (uint32_t)y[6], (uint32_t)y[7]);
        }
        const int32_t x[8] = {(int32_t)a.x, (int32_t)a.y, (int32_t)a.z, (int32_t)a.w, (int32_t)b.x, (int32_t)b.y, (int32_t)b.z, (int32_t)b.w};
        if (j * 8 + 8 <= n) {
#pragma unroll
            for (int t = 0; t < 8; ++t) f(j * 8 + t, ((uint32_t)x[t] << sh) ^ ((uint32_t)(x[t] >> 31) & sm));
        } else {
#pragma unroll
            for (int t = 0; t < 8; ++t)
                if (j * 8 + t < n) f(j * 8 + t, ((uint32_t)x[t] << sh) ^ ((uint32_t)(x[t] >> 31) & sm));
        }
    }
}

__device__ __forceinline__ uint32_t last_mapped(const LaneBlock &B)      // the block's last (non-zero) symbol, mapped
{
    const int32_t x = B.cnt > 0 ? __ldg(B.sym + B.cnt - 1) : 0;
    return B.flag_signed ? s2u(x) : (uint32_t)x;
}

// code length of the block from state st (in/out)
__device__ __forceinline__ void count_lane_sym(const LaneBlock &B, State &st, uint32_t &bits_out)
{
    SymState t;
    t.kP = st.kP; t.kRP = st.kRP; t.m = 0;
    uint32_t bits = 0;
    if (B.gap0 && B.cnt > 0) {
        CountSink z{0};
        sym_begin(t, B.gap0, z);
        bits = z.bits;
    }
    walk_symbols(B, [&](int, uint32_t u) { sym_count(t, u, bits); });
    State s;
    s.kP = t.kP; s.kRP = t.kRP;
    if (B.is_tail) {
        CountSink k{0};
        code_tail(s, B.tail_zeros, B.cnt > 0 && sym_tail_quirk(t, last_mapped(B)), k);
        bits += k.bits;
    }
    st = s;
    bits_out = bits;
}

// Code lengths of the lane's block for NC candidate start states side by side.  s[] in/out.  The candidates are
// independent dependency chains kept in registers; for NC > 1 their zero runs advance together in one loop of
// predicated steps (separate data-dependent loops would serialise the chains and multiply the trip count the warp
// pays by NC).
template <int NC> __device__ __forceinline__ void count_lane(const LaneBlock &B, State *s_io, uint32_t *bits_out)
{
    if (NC == 1 && B.sym) {                                  // (warp-uniform: a chunk belongs to one channel)
        count_lane_sym(B, s_io[0], bits_out[0]);
        return;
    }
    const int lane = B.slot;
    int kP[NC], kRP[NC];
    uint32_t bits[NC];
    bool quirk[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) { kP[c] = s_io[c].kP; kRP[c] = s_io[c].kRP; bits[c] = 0; quirk[c] = false; }
    const int units = (B.cnt + EV_GROUP - 1) / EV_GROUP;
    // the units of a block are at fixed addresses: two of them are requested ahead of the one being coded (the coder's
    // state chain is long, but shorter than a round trip to the event store)
    uint4 a1, b1, a2, b2;
    a1 = b1 = a2 = b2 = make_uint4(0u, 0u, 0u, 0u);
    if (units > 0) { a1 = __ldg(B.base + lane * 2); b1 = __ldg(B.base + lane * 2 + 1); }
    if (units > 1) { a2 = __ldg(B.base + (32 + lane) * 2); b2 = __ldg(B.base + (32 + lane) * 2 + 1); }
    for (int j = 0; j < units; ++j) {
        const uint4 a = a1, b = b1;
        a1 = a2; b1 = b2;
        if (j + 2 < units) { a2 = __ldg(B.base + ((j + 2) * 32 + lane) * 2); b2 = __ldg(B.base + ((j + 2) * 32 + lane) * 2 + 1); }
        const uint32_t g[4] = {a.x, a.z, b.x, b.z}, u[4] = {a.y, a.w, b.y, b.w};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            if (j * EV_GROUP + t >= B.cnt) break;
            if (NC == 1) {
                State st;
                st.kP = kP[0]; st.kRP = kRP[0];
                quirk[0] = count_event(st, g[t], u[t], bits[0]);
                kP[0] = st.kP; kRP[0] = st.kRP;
                continue;
            }
            if (g[t]) {                                     // the zero run, all candidates together (code_zeros, lengths only)
                uint32_t gc[NC];
#pragma unroll
                for (int c = 0; c < NC; ++c) gc[c] = g[t];
                bool more = true;
                while (more) {
                    more = false;
#pragma unroll
                    for (int c = 0; c < NC; ++c) {
                        const int k = kP[c] >> 2;
                        const uint32_t run = 1u << (k & 31);
                        const bool z0 = k == 0 && gc[c] != 0;               // one zero in no-run mode (membuf.cpp:391-413, u == 0)
                        const bool zr = k != 0 && gc[c] >= run;             // a complete run (:382-390)
                        bits[c] += z0 ? 1u + (uint32_t)(kRP[c] >> 2) : (zr ? 1u : 0u);
                        kRP[c] = z0 ? (kRP[c] < 2 ? 0 : kRP[c] - 2) : kRP[c];
                        kP[c] += z0 ? RL_U0 : (zr ? RL_U1 : 0);
                        gc[c] -= z0 ? 1u : (zr ? run : 0u);
                        more |= z0 || zr;
                    }
                }
            }
#pragma unroll
            for (int c = 0; c < NC; ++c) {                  // the non-zero symbol (count_event without its zero run)
                const int k = kP[c] >> 2;
                const unsigned kR = (unsigned)(kRP[c] >> 2);
                const uint32_t v = u[t] - (k ? 1u : 0u);
                const uint32_t p = kR > 31 ? 0u : (v >> kR);
                bits[c] += (p < 32 ? p + 1 + kR : 64u) + (k ? 1u + (uint32_t)k : 0u);
                const int up = kRP[c] + (int)(p > 256u ? 256u : p) - 1;
                kRP[c] = p ? (up > RL_KRP_MAX ? RL_KRP_MAX : up) : (kRP[c] < 2 ? 0 : kRP[c] - 2);
                kP[c] = kP[c] < 1 ? 0 : kP[c] - 1;
                quirk[c] = k != 0 && v == 0;
            }
        }
    }
#pragma unroll
    for (int c = 0; c < NC; ++c) {
        State st;
        st.kP = kP[c]; st.kRP = kRP[c];
        if (B.is_tail) {
            CountSink k{0};
            code_tail(st, B.tail_zeros, quirk[c], k);
            bits[c] += k.bits;
        }
        s_io[c] = st;
        bits_out[c] = bits[c];
    }
}

// ---- one chunk: speculative block states made consistent lane to lane -----------------------------------------
// Every lane carries NC candidate start states of its block with the end state and code length under each
// (NC = 1 for channels that are mostly non-zero: they re-synchronise through k_P = 0 within a block; NC = 8 = the
// state classes of rlgr_core.cuh::expand_classes otherwise).  pin == false (speculation): lane 0 codes the block
// before the chunk as warm-up; pin == true (chunk 0, repair): lane 0 forwards the known state at the chunk start
// and lane 1 keeps it.  Round 1: every other lane codes its block from the canonical states.  After round 1 a lane
// takes the classes around its predecessor's end states as candidates; later, a lane whose predecessor ended in a
// state that is not among its candidates takes the predecessor's end states (results of states it already knows
// are kept).  Candidates are LINKED, not index-matched: candidate c of a lane continues in whichever candidate of
// the next lane equals its end state, so a chain may change slots from lane to lane, and only one chain has to be
// complete: a candidate of the wrong class never re-synchronises (its chain would need one round per lane), the
// right one merges with the true trajectory within a few events.  The rounds stop when the pinned chain (pin) /
// any chain (speculation) reaches the last lane, or after max_rounds; chains that are incomplete then are marked
// invalid.  With pin the true state advances at least one lane per round: 33 rounds always complete the chain.
constexpr uint32_t STATE_INVALID = 0xffffu;
constexpr int SEL_NONE = 15;

template <int NC> struct ChunkSol {
    uint32_t in[NC], out[NC];       // packed states
    uint32_t bits[NC];
    uint32_t sel;                   // nibble c: this lane's candidate on the chain of chunk candidate c (15: broken)
    unsigned complete;              // warp-uniform mask: chain c reaches the last lane
    uint32_t end[NC], total[NC];    // warp-uniform: packed state after the chunk / code length of the chunk on chain c
    int rounds;
};

// The 8-candidate solves of the sparse channels are bound by the latency of a lone warp (a few chunks per channel,
// every lane coding its block from up to 8 states one after the other: measured 200 k cycles per round against 30 k
// for a dense chunk).  Their CTA therefore carries 7 more warps that each code the blocks from ONE of the distinct
// states: the solving warp posts the states in shared memory, all 8 warps run the single-state coder side by side.
struct Helpers {
    uint32_t state[8][32], out[8][32], bits[8][32];      // [distinct state k][lane]
    int m[32];                                           // distinct states posted by the lane
    long long w;                                         // chunk being solved
    int cmd;                                             // 1: states posted, 0: finished
};

__device__ __forceinline__ void helper_loop(Helpers *H, const uint2 *__restrict__ ev, const Geometry &G, int ch, const ChanInfo &I)
{
    const int lane = threadIdx.x & 31, k = threadIdx.x >> 5;
    for (;;) {
        __syncthreads();                                                        // (A) a command is posted
        if (H->cmd == 0) break;
        if (H->m[lane] > k) {
            const LaneBlock B = lane_block(ev, G, ch, (int64_t)H->w, I);
            State t[1] = {unpack_state(H->state[k][lane])};
            uint32_t b[1];
            count_lane<1>(B, t, b);
            H->out[k][lane] = pack_state(t[0]);
            H->bits[k][lane] = b[0];
        }
        __syncthreads();                                                        // (B) results are in
    }
}
__device__ __forceinline__ void helpers_release(Helpers *H)                      // by the solving warp, all lanes
{
    __syncwarp();
    if ((threadIdx.x & 31) == 0) H->cmd = 0;
    __syncthreads();                                                            // (A)
}

template <int NC> __device__ __forceinline__ void recode(const LaneBlock &B, ChunkSol<NC> &S, unsigned dirty, Helpers *H)
{
    if (NC == 1) {
        if (!dirty) return;
        State t[1] = {unpack_state(S.in[0])};
        uint32_t b[1];
        count_lane<1>(B, t, b);
        S.out[0] = pack_state(t[0]);
        S.bits[0] = b[0];
        return;
    }
    if (!H && !dirty) return;
    // the distinct states among the dirty candidates, coded side by side
    uint32_t ds[NC];
    int rep[NC], m = 0;
    for (int c = 0; c < NC; ++c) {
        rep[c] = 0;
        if (!(dirty >> c & 1)) continue;
        int k = 0;
        while (k < m && ds[k] != S.in[c]) ++k;
        if (k == m) ds[m++] = S.in[c];
        rep[c] = k;
    }
    State t[NC];
    uint32_t b[NC];
    for (int k = 0; k < NC; ++k) { t[k] = unpack_state(ds[k < m ? k : 0]); b[k] = 0; }
    if (H) {                                             // (the whole warp is here: the barriers are CTA-wide)
        const int lane = threadIdx.x & 31;
        if (!__any_sync(~0u, m > 1)) {                   // one state per lane: nothing to share out
            if (m) count_lane<1>(B, t, b);
        } else {
            H->m[lane] = m;
            for (int k = 1; k < NC; ++k) if (k < m) H->state[k][lane] = ds[k];
            if (lane == 0) H->cmd = 1;
            __syncthreads();                                                    // (A)
            if (m) count_lane<1>(B, t, b);
            __syncthreads();                                                    // (B)
            for (int k = 1; k < NC; ++k) if (k < m) { t[k] = unpack_state(H->out[k][lane]); b[k] = H->bits[k][lane]; }
        }
    } else if (m == 1) count_lane<1>(B, t, b);
    else if (m == 2) count_lane<(NC >= 2 ? 2 : 1)>(B, t, b);
    else if (m <= 4) count_lane<(NC >= 4 ? 4 : 1)>(B, t, b);
    else count_lane<NC>(B, t, b);
    for (int c = 0; c < NC; ++c)
        if (dirty >> c & 1) { S.out[c] = pack_state(t[rep[c]]); S.bits[c] = b[rep[c]]; }
}

// chains from lane 1's candidates through the lanes in order (src skips the lanes without a block)
template <int NC> __device__ __forceinline__ void walk_chains(ChunkSol<NC> &S, bool has, int src)
{
    const int lane = threadIdx.x & 31;
    uint32_t from = 0;               // nibble c: my candidate that continues candidate c of my predecessor
    for (int c = 0; c < NC; ++c) {
        const uint32_t o = __shfl_sync(~0u, S.out[c], src);
        int f = SEL_NONE;
        for (int k = NC - 1; k >= 0; --k) if (S.in[k] == o) f = k;
        if (!has) f = c;
        from |= (uint32_t)f << (4 * c);
    }
    uint32_t sel = 0;
    for (int c = 0; c < NC; ++c) sel |= (uint32_t)(lane == 1 ? c : SEL_NONE) << (4 * c);
    for (int step = 2; step < 32; ++step) {
        const uint32_t p = __shfl_sync(~0u, sel, src);
        if (lane == step) {
            sel = 0;
            for (int c = 0; c < NC; ++c) {
                const int pc = (int)(p >> (4 * c)) & 15;
                sel |= (uint32_t)(pc == SEL_NONE ? SEL_NONE : (int)(from >> (4 * pc)) & 15) << (4 * c);
            }
        }
    }
    S.sel = sel;
    S.complete = 0;
    for (int c = 0; c < NC; ++c) {
        const int sc = (int)(sel >> (4 * c)) & 15;
        const unsigned broken = __ballot_sync(~0u, sc == SEL_NONE) & ~1u;          // lanes 1..31
        if (!broken) S.complete |= 1u << c;
        uint32_t b = 0, e = 0;
        for (int k = 0; k < NC; ++k) if (k == sc) { b = S.bits[k]; e = S.out[k]; }
        if (lane == 0) b = 0;                                                     // the warm-up block belongs to the previous chunk
        for (int o = 16; o; o >>= 1) b += __shfl_xor_sync(~0u, b, o);
        S.total[c] = b;
        S.end[c] = __shfl_sync(~0u, e, 31);
    }
}

// pin_states[NC]: the candidates of lane 1 when pin (the known state at the chunk start NC times, or the states the
// previous chunk can end in); pin_need: mask of the pinned candidates whose chains must be complete before stopping.
template <int NC>
__device__ __forceinline__ void solve_chunk(const LaneBlock &B, bool pin, const uint32_t *pin_states, unsigned pin_need,
                                            int max_rounds, ChunkSol<NC> &S, Helpers *H = nullptr)
{
    const int lane = threadIdx.x & 31;
    bool has = B.cnt > 0 || B.is_tail;
    if (lane == 0) has = !pin && has;                       // a pinned lane 0 only forwards the known state
    const unsigned active = __ballot_sync(~0u, has);
    const unsigned before = active & lanemask_lt();
    const int src = before ? 31 - __clz((int)before) : 0;  // nearest earlier lane that changes the state
    State canon[2];
    canon[0] = initial_state();
    canon[1].kP = 9;                                        // run mode, the other k_P parity
    canon[1].kRP = RL_KRP_INIT;
    for (int c = 0; c < NC; ++c) {
        S.in[c] = (pin && lane < 2) ? pin_states[c] : pack_state(canon[c < NC / 2 ? 0 : 1]);
        S.out[c] = S.in[c];
        S.bits[c] = 0;
    }
    unsigned dirty = has ? (1u << NC) - 1 : 0u;
    int r = 0;
    for (;;) {
        recode<NC>(B, S, dirty, H);
        ++r;
        dirty = 0;
        uint32_t o[NC];
        for (int c = 0; c < NC; ++c) o[c] = __shfl_sync(~0u, S.out[c], src);
        if (lane > 0 && !has)
            for (int c = 0; c < NC; ++c) S.in[c] = S.out[c] = o[c];             // no block: pass the states through
        if (r >= 2) {
            walk_chains<NC>(S, has, src);
            if (pin ? ((S.complete & pin_need) == pin_need) : (S.complete != 0)) break;
        }
        if (lane > 0 && has && r < max_rounds && !(pin && lane == 1)) {
            uint32_t nw[NC];
            bool replace = false;
            bool expanded = false;
            if constexpr (NC == 8) {
                if (r == 1) {                              // the classes around the predecessor's end states
                    expand_classes(unpack_state(o[0]), unpack_state(o[NC / 2]), nw);
                    replace = expanded = true;
                }
            }
            if (!expanded && !(NC > 1 && lane == 1)) {           // (lane 1's candidates are the chunk's: fixed after the expansion)
                for (int c = 0; c < NC; ++c) {
                    bool cov = false;
                    for (int k = 0; k < NC; ++k) cov |= o[c] == S.in[k];
                    replace |= !cov;
                    nw[c] = o[c];
                }
            }
            if (replace) {
                uint32_t oin[NC], oout[NC], obits[NC];
                for (int c = 0; c < NC; ++c) { oin[c] = S.in[c]; oout[c] = S.out[c]; obits[c] = S.bits[c]; }
                for (int c = 0; c < NC; ++c) {
                    int m = -1;
                    for (int k = NC - 1; k >= 0; --k) if (nw[c] == oin[k]) m = k;
                    S.in[c] = nw[c];
                    if (m >= 0) { S.out[c] = oout[m]; S.bits[c] = obits[m]; }
                    else dirty |= 1u << c;
                }
            }
        }
        if (!__any_sync(~0u, dirty != 0)) {
            if (r < 2) walk_chains<NC>(S, has, src);
            break;
        }
    }
    S.rounds = r;
}

// ---- records -----------------------------------------------------------------------------------------------
// chunk record (4 x uint4): [0] candidates (8 x 16 bit: assumed states at the chunk start, 0xffff: chain incomplete),
//                           [1] end states (8 x 16 bit), [2], [3] code length of the chunk on each chain (8 x 32 bit)
// lane record  (3 x uint4): [0] the lane's 8 candidates (16 bit each), [1] code lengths of its block (8 x 16 bit),
//                           [2].x nibble c = the lane's candidate on the chain of chunk candidate c
//                           [3] end states of its block under each candidate (8 x 16 bit; speculation records only:
//                               what lets a repair join the recorded chains instead of coding every block again)
// chunk offset (u64)      : bit offset of the chunk in its stream << 3 | chosen chunk candidate
// sync index   (uint2)    : x = bit offset (low 32), y = offset bits 32..38 | k_P << 7 | k_RP << 14 | first << 22
constexpr int REC_WORDS = 4, LANEREC_WORDS = 4;

template <int NC> __device__ __forceinline__ void finish_chunk(const ChunkSol<NC> &S, uint4 *rec, uint4 *lanerec)
{
    const int lane = threadIdx.x & 31;
    uint32_t in16[4] = {0, 0, 0, 0}, b16[4] = {0, 0, 0, 0}, out16[4] = {0, 0, 0, 0};
    uint32_t cand16[4] = {~0u, ~0u, ~0u, ~0u}, end16[4] = {0, 0, 0, 0}, tot[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int c = 0; c < NC; ++c) {
        in16[c >> 1] |= (S.in[c] & 0xffffu) << (16 * (c & 1));
        b16[c >> 1] |= (S.bits[c] & 0xffffu) << (16 * (c & 1));
        out16[c >> 1] |= (S.out[c] & 0xffffu) << (16 * (c & 1));
        const uint32_t l1 = __shfl_sync(~0u, S.in[c], 1);
        const uint32_t cand = (S.complete >> c & 1) ? l1 : STATE_INVALID;
        cand16[c >> 1] = (cand16[c >> 1] & ~(0xffffu << (16 * (c & 1)))) | (cand << (16 * (c & 1)));
        end16[c >> 1] |= (S.end[c] & 0xffffu) << (16 * (c & 1));
        tot[c] = S.total[c];
    }
    uint4 *lr = lanerec + (size_t)lane * LANEREC_WORDS;
    lr[0] = make_uint4(in16[0], in16[1], in16[2], in16[3]);
    lr[1] = make_uint4(b16[0], b16[1], b16[2], b16[3]);
    lr[2] = make_uint4(S.sel, 0u, 0u, 0u);
    lr[3] = make_uint4(out16[0], out16[1], out16[2], out16[3]);
    if (NC < 8) lr[3] = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);      // (only the 8-candidate records are joined)
    if (lane == 0) {
        rec[0] = make_uint4(cand16[0], cand16[1], cand16[2], cand16[3]);
        rec[1] = make_uint4(end16[0], end16[1], end16[2], end16[3]);
        rec[2] = make_uint4(tot[0], tot[1], tot[2], tot[3]);
        rec[3] = make_uint4(tot[4], tot[5], tot[6], tot[7]);
    }
}

__device__ __forceinline__ uint32_t half_of(const uint4 &v, int c)       // 16-bit entry c of 8
{
    const uint32_t w = (c >> 1) == 0 ? v.x : (c >> 1) == 1 ? v.y : (c >> 1) == 2 ? v.z : v.w;
    return (w >> (16 * (c & 1))) & 0xffffu;
}
__device__ __forceinline__ uint32_t word_of(const uint4 &a, const uint4 &b, int c)   // 32-bit entry c of 8
{
    const uint4 &v = c < 4 ? a : b;
    return (c & 3) == 0 ? v.x : (c & 3) == 1 ? v.y : (c & 3) == 2 ? v.z : v.w;
}

// channels that are mostly non-zero run with one candidate (they re-synchronise through k_P = 0 within a block)
__device__ __forceinline__ bool channel_is_dense(const ChanInfo &I, const Geometry &) { return I.dense != 0; }

// ---- A: speculate ----------------------------------------------------------------------------------------
constexpr int SPEC_WARPS = 4;        // table / emit passes: chunks per CTA
constexpr int SOLVE_WARPS = 8;       // speculation and resolve: 8 dense chunks per CTA, or one sparse chunk and its 7 helper warps
constexpr int SPEC_ROUNDS = 4;       // speculation: a chain that is not complete by then is left invalid
constexpr int RESPEC_ROUNDS = 6;     // second sweep (candidates = the predecessor's end states)
constexpr int PIN_ROUNDS = 34;       // chunk 0 and repairs: the pinned chain always completes

// sweep 0: every chunk from the canonical warm-up.  sweep 1: a chunk whose predecessor can end (on one of its
// complete chains) in a state that is not among the chunk's own valid candidates is solved again with exactly
// those end states as its candidates -- after the predecessor's 31 blocks they are far better guesses than one
// warm-up block, and the chains of a run of such chunks are repaired together instead of one after the other in
// the resolve pass.  Records are double-buffered (rec_prev -> rec) so that the sweep reads a consistent state.
template <int NC>
__device__ __forceinline__ void speculate_chunk(const Geometry &G, const ChanInfo *__restrict__ info, const uint2 *__restrict__ ev,
                                                int sweep, const uint4 *__restrict__ rec_prev, uint4 *__restrict__ rec,
                                                uint4 *__restrict__ lanerec, unsigned long long *stats, Helpers *H)
{
    const int ch = blockIdx.y;
    const int warp = threadIdx.x >> 5;
    // dense: a chunk per warp; sparse: a chunk per CTA, warp 0 solves and the others code single states for it
    const int64_t w = NC == 1 ? (int64_t)blockIdx.x * SOLVE_WARPS + warp : (int64_t)blockIdx.x;
    const int lane = threadIdx.x & 31;
    const ChanInfo I = info[ch];
    if (w >= (int64_t)I.nchunks || channel_is_dense(I, G) != (NC == 1)) return;
    const size_t ci = (size_t)ch * G.W + w;
    uint32_t pins[NC];
    unsigned need = 1u;
    bool pin = w == 0;
    for (int c = 0; c < NC; ++c) pins[c] = pack_state(initial_state());
    if (sweep > 0) {
        const uint4 *mine = rec_prev + ci * REC_WORDS;
        bool again = false;
        if (w > 0 && NC == 1) {                                   // (the sparse channels go through the tables instead)
            const uint4 pc = rec_prev[(ci - 1) * REC_WORDS], pe = rec_prev[(ci - 1) * REC_WORDS + 1], mc = mine[0];
            uint32_t fill = STATE_INVALID;
            for (int c = 0; c < NC; ++c)
                if (half_of(pc, c) != STATE_INVALID) {            // a complete chain of the predecessor: its end state
                    const uint32_t e = half_of(pe, c);
                    if (fill == STATE_INVALID) fill = e;
                    bool have = false;
                    for (int k = 0; k < NC; ++k) have |= half_of(mc, k) == e;
                    again |= !have;
                }
            if (again) {
                need = 0;
                for (int c = 0; c < NC; ++c) {
                    const bool ok = half_of(pc, c) != STATE_INVALID;
                    pins[c] = ok ? half_of(pe, c) : fill;
                    bool dup = false;
                    for (int k = 0; k < c; ++k) dup |= pins[k] == pins[c];
                    if (ok && !dup) need |= 1u << c;
                }
                pin = true;
            }
        }
        if (!again) {                                             // keep the record of the first sweep
            if (warp == 0 || NC == 1)
                if (lane < REC_WORDS) rec[ci * REC_WORDS + lane] = mine[lane];
            return;
        }
    }
    if (NC > 1 && warp > 0) {                                     // helper warps of a sparse chunk
        helper_loop(H, ev, G, ch, I);
        return;
    }
    if (NC > 1 && lane == 0) H->w = (long long)w;                 // (read by the helpers after the first barrier)
    const LaneBlock B = lane_block(ev, G, ch, w, I);
    ChunkSol<NC> S;
    const long long t0 = clock64();
    // (chunk 0 runs until its pinned chain is complete: with the helper warps a sparse chunk's rounds are cheap, and
    // the re-solve it would otherwise need sits on the critical path of the resolve pass)
    solve_chunk<NC>(B, pin, pins, need, w == 0 ? PIN_ROUNDS : (sweep ? RESPEC_ROUNDS : SPEC_ROUNDS), S, NC > 1 ? H : nullptr);
    if (NC > 1) helpers_release(H);
    finish_chunk<NC>(S, rec + ci * REC_WORDS, lanerec + ci * 32 * LANEREC_WORDS);
    if ((threadIdx.x & 31) == 0 && stats) {
        atomicAdd(stats + 0, 1ull);
        atomicAdd(stats + 1, (unsigned long long)S.rounds);
        // slowest chunk of the pass: cycles << 24 | channel << 8 | rounds
        atomicMax(stats + 3, ((unsigned long long)(clock64() - t0) << 24) | ((unsigned long long)(ch & 0xffff) << 8) | (unsigned)S.rounds);
    }
}

// one launch for both kinds of channel (blockIdx.z): the few chunks of the sparse channels (z = 0, scheduled first:
// one chunk per CTA, bound by the latency of their chains) run under the many chunks of the dense channels
// (z = 1: 8 chunks per CTA, bound by issue slots)
__global__ void __launch_bounds__(32 * SOLVE_WARPS, 4) rlgr_speculate_kernel(Geometry G, const ChanInfo *__restrict__ info,
                                                                         const uint2 *__restrict__ ev, int sweep,
                                                                         const uint4 *__restrict__ rec_prev, uint4 *__restrict__ rec,
                                                                         uint4 *__restrict__ lanerec, unsigned long long *stats)
{
    pdl_enter();
    __shared__ Helpers helpers;
    if (blockIdx.z == 0) speculate_chunk<8>(G, info, ev, sweep, rec_prev, rec, lanerec, stats, &helpers);
    else speculate_chunk<1>(G, info, ev, sweep, rec_prev, rec, lanerec, stats, nullptr);
}

// ---- table mode for the sparse channels ---------------------------------------------------------------------------
// A sparse channel in run mode with k_R = 0 does not re-synchronise: its block map is one-to-one on the states it
// visits, and when the values are one-signed (position channels: no p = 0 event ever clamps k_RP) the classes are
// more than a handful of candidates can cover.  There the block's transition is tabulated instead of guessed:
//   T1  warp per block: the block coded from all 128 states of a window (16 k_P x 8 k_RP values around the state the
//       first sweep estimated at the block's start), 4 states per lane as independent chains, events broadcast;
//   T2  warp per chunk: the 128 states of its first block's window followed through the 31 tables;
//   resolve looks the true state up in the chunk's table (falls back to the candidate records / a pinned repair when
//       it lies outside the window or its chain leaves a later window);
//   T3  the chunks resolved that way get their lane records from one more walk through the tables.
// Exact by construction: a table entry is the coder's own arithmetic from that state.
constexpr int TAB_KP = 16, TAB_KRP = 8, TAB_STATES = TAB_KP * TAB_KRP;       // 128 = 32 lanes x 4
constexpr uint32_t TAB_DEAD = 0xffffu;

__device__ __forceinline__ uint32_t tab_window(uint32_t centre)      // packed (k_P0, k_RP0) of the window around a state
{
    const State c = unpack_state(centre);
    State w;
    w.kP = c.kP > 6 ? c.kP - 6 : 0;
    w.kRP = c.kRP > 3 ? c.kRP - 3 : 0;
    return pack_state(w);
}
__device__ __forceinline__ int tab_index(uint32_t window, uint32_t state)    // -1: outside
{
    if (window == 0xffffffffu) return -1;
    const State w = unpack_state(window), s = unpack_state(state);
    const int a = s.kP - w.kP, b = s.kRP - w.kRP;
    return (a >= 0 && a < TAB_KP && b >= 0 && b < TAB_KRP) ? a * TAB_KRP + b : -1;
}
__device__ __forceinline__ uint32_t tab_state(uint32_t window, int idx)
{
    State s = unpack_state(window);
    s.kP += idx / TAB_KRP;
    s.kRP += idx % TAB_KRP;
    return pack_state(s);
}

// Only the chunks the candidate records may fail on are tabulated: no complete chain of their own, or a complete
// chain of the predecessor ending in a state they do not have (in a stretch of such chunks every one qualifies).
constexpr uint32_t TAB_NO_WINDOW = 0xffffffffu;
__device__ __forceinline__ bool chunk_suspect1(const uint4 *__restrict__ rec, size_t ci, int64_t w)
{
    const uint4 mc = rec[ci * REC_WORDS];
    bool any = false;
    for (int c = 0; c < 8; ++c) any |= half_of(mc, c) != STATE_INVALID;
    if (!any) return true;
    if (w == 0) return false;
    const uint4 pc = rec[(ci - 1) * REC_WORDS], pe = rec[(ci - 1) * REC_WORDS + 1];
    for (int c = 0; c < 8; ++c)
        if (half_of(pc, c) != STATE_INVALID) {
            const uint32_t e = half_of(pe, c);
            bool have = false;
            for (int k = 0; k < 8; ++k) have |= half_of(mc, k) == e;
            if (!have) return true;
        }
    return false;
}

// ... and the two chunks behind such a chunk: the state a table hands on need not be one the next chunks guessed
__device__ __forceinline__ bool chunk_suspect(const uint4 *__restrict__ rec, size_t ci, int64_t w)
{
    return chunk_suspect1(rec, ci, w) || (w >= 1 && chunk_suspect1(rec, ci - 1, w - 1)) || (w >= 2 && chunk_suspect1(rec, ci - 2, w - 2));
}

// tables: uint32 [channel][block][128] = end state | bits << 16; headers: uint32 [channel][block] = window
__global__ void __launch_bounds__(32 * SPEC_WARPS) rlgr_table_kernel(Geometry G, const ChanInfo *__restrict__ info,
                                                                    const uint2 *__restrict__ ev, const uint4 *__restrict__ rec,
                                                                    const uint4 *__restrict__ lanerec,
                                                                    uint32_t *__restrict__ tables, uint32_t *__restrict__ headers)
{
    pdl_trigger();
    const int ch = blockIdx.y;
    const int lane = threadIdx.x & 31;
    const ChanInfo I = info[ch];
    if (channel_is_dense(I, G)) return;
    // (a grid-stride loop over the channel's blocks: the launch of one CTA per 4 blocks cost 66 us on the 1 M scene,
    // where every channel is dense and every CTA returned at once)
    for (int64_t b = (int64_t)blockIdx.x * SPEC_WARPS + (threadIdx.x >> 5); b < (int64_t)I.nblocks; b += (int64_t)gridDim.x * SPEC_WARPS) {
    const int64_t w = b / CHUNK_LANES;
    const int l = (int)(b - w * CHUNK_LANES) + 1;
    if (!chunk_suspect(rec, (size_t)ch * G.W + w, w)) {
        if (lane == 0) headers[(size_t)ch * G.nsb + b] = TAB_NO_WINDOW;
        continue;
    }
    LaneBlock B;
    B.base = (const uint4 *)(ev + ((size_t)ch * G.W + w) * CHUNK_SLOTS);
    B.slot = l;
    B.sym = nullptr; B.gap0 = 0; B.pos0 = 0; B.vec = 0; B.flag_signed = G.flag_signed; B.avail = 0;
    int64_t cnt = (int64_t)I.events - b * EV_BLOCK;
    B.cnt = (int)(cnt < 0 ? 0 : (cnt > EV_BLOCK ? EV_BLOCK : cnt));
    B.is_tail = b == (int64_t)I.nblocks - 1;
    B.tail_zeros = I.tail_zeros;
    // window around the first sweep's estimate of the state at the block's start (slot 1 of the lane's candidates)
    const uint4 lr0 = lanerec[(((size_t)ch * G.W + w) * 32 + l) * LANEREC_WORDS];
    const uint32_t window = tab_window(half_of(lr0, 1));
    State st[4];
    uint32_t bits[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) st[k] = unpack_state(tab_state(window, lane * 4 + k));
    count_lane<4>(B, st, bits);
    uint32_t *T = tables + ((size_t)ch * G.nsb + b) * TAB_STATES;
    uint4 out;
    out.x = pack_state(st[0]) | (bits[0] << 16); out.y = pack_state(st[1]) | (bits[1] << 16);
    out.z = pack_state(st[2]) | (bits[2] << 16); out.w = pack_state(st[3]) | (bits[3] << 16);
    ((uint4 *)T)[lane] = out;
    if (lane == 0) headers[(size_t)ch * G.nsb + b] = window;
    }
}

// chunk tables: uint2 [channel][chunk][128] = (end state or TAB_DEAD, code length of the chunk)
__global__ void __launch_bounds__(32 * SPEC_WARPS) rlgr_chunk_table_kernel(Geometry G, const ChanInfo *__restrict__ info,
                                                                          const uint32_t *__restrict__ tables,
                                                                          const uint32_t *__restrict__ headers,
                                                                          uint2 *__restrict__ ctab)
{
    pdl_enter();
    const int ch = blockIdx.y;
    const int64_t w = (int64_t)blockIdx.x * SPEC_WARPS + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    const ChanInfo I = info[ch];
    if (w >= (int64_t)I.nchunks || channel_is_dense(I, G)) return;
    const int64_t b0 = w * CHUNK_LANES;
    const int nb = (int)min((int64_t)CHUNK_LANES, (int64_t)I.nblocks - b0);
    const uint32_t *H = headers + (size_t)ch * G.nsb + b0;
    const uint32_t *T = tables + ((size_t)ch * G.nsb + b0) * TAB_STATES;
    uint32_t s[4], bits[4] = {0, 0, 0, 0};
    const uint32_t w0 = H[0];
    if (w0 == TAB_NO_WINDOW) return;                      // not tabulated: the resolve pass sees no window
#pragma unroll
    for (int k = 0; k < 4; ++k) s[k] = tab_state(w0, lane * 4 + k);
    for (int j = 0; j < nb; ++j) {
        const uint32_t win = H[j];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int idx = s[k] == TAB_DEAD ? -1 : tab_index(win, s[k]);
            if (idx < 0) { s[k] = TAB_DEAD; continue; }
            const uint32_t e = T[(size_t)j * TAB_STATES + idx];
            s[k] = e & 0xffffu;
            bits[k] += e >> 16;
        }
    }
    uint2 *C = ctab + ((size_t)ch * G.W + w) * TAB_STATES + lane * 4;
#pragma unroll
    for (int k = 0; k < 4; ++k) C[k] = make_uint2(s[k], bits[k]);
}

// lane records of the chunks the resolve pass settled through the tables (cstart = true state | 1 << 16)
__global__ void __launch_bounds__(32 * SPEC_WARPS) rlgr_table_lanes_kernel(Geometry G, const ChanInfo *__restrict__ info,
                                                                          const uint32_t *__restrict__ tables,
                                                                          const uint32_t *__restrict__ headers,
                                                                          const uint32_t *__restrict__ cstart, uint4 *__restrict__ lanerec)
{
    pdl_enter();
    const int ch = blockIdx.y;
    const int64_t w = (int64_t)blockIdx.x * SPEC_WARPS + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    const ChanInfo I = info[ch];
    if (w >= (int64_t)I.nchunks || channel_is_dense(I, G)) return;
    const size_t ci = (size_t)ch * G.W + w;
    const uint32_t cs = cstart[ci];
    if (!(cs >> 16)) return;
    const int64_t b0 = w * CHUNK_LANES;
    const int nb = (int)min((int64_t)CHUNK_LANES, (int64_t)I.nblocks - b0);
    const uint32_t *H = headers + (size_t)ch * G.nsb + b0;
    const uint32_t *T = tables + ((size_t)ch * G.nsb + b0) * TAB_STATES;
    uint32_t s = cs & 0xffffu, mine_s = s, mine_b = 0;
    for (int j = 0; j < nb; ++j) {                        // every lane walks; lane j + 1 keeps block j
        const uint32_t e = T[(size_t)j * TAB_STATES + tab_index(H[j], s)];
        if (lane == j + 1) { mine_s = s; mine_b = e >> 16; }
        s = e & 0xffffu;
    }
    if (lane > nb) mine_s = s;                            // lanes without a block: the state passes through
    uint4 *lr = lanerec + (ci * 32 + lane) * LANEREC_WORDS;
    lr[0] = make_uint4(mine_s, 0u, 0u, 0u);
    lr[1] = make_uint4(lane == 0 ? 0u : mine_b, 0u, 0u, 0u);
    lr[2] = make_uint4(0u, 0u, 0u, 0u);
}

// Exact walk of one chunk from its TRUE start state, block after block: lane l codes block l once the state has come
// down the lanes to it.  This is the repair of the sparse channels: their trajectories do not re-synchronise (that is why
// the chunk ended up here), so a pinned fixed point would advance the true state by one lane per ROUND anyway -- 29
// rounds of 8-candidate recodes with helper hand-overs, 0.8 ms per chunk on the 50 M-symbol position channel (97 such
// chunks: 78 of the 105 ms of that encode) -- where the walk pays 31 single-state block codings and nothing else.
// `stage` (shared memory, 32 x EV_BLOCK events; may be null): with one lane coding at a time, the event loads of
// count_lane are lone 32-byte requests that each pay a full round trip (190 SM clocks per event, 0.4 ms per chunk).  All
// lanes therefore copy their blocks into shared memory first -- together, coalesced -- and the serial pass reads them
// from there.
__device__ __forceinline__ void walk_chunk_exact(const LaneBlock &B, uint32_t start, uint4 *rec, uint4 *lanerec, uint32_t &end_state,
                                                 uint32_t &total_bits, uint2 *stage, bool use_record, int &blocks_coded)
{
    const int lane = threadIdx.x & 31;
    const bool has = lane > 0 && (B.cnt > 0 || B.is_tail);
    const bool staged = stage != nullptr && B.sym == nullptr;          // (warp-uniform: a chunk belongs to one channel)
    if (staged) {
        uint4 *mine = (uint4 *)(stage + (size_t)lane * EV_BLOCK);
        const int units = lane > 0 ? (B.cnt + EV_GROUP - 1) / EV_GROUP : 0;
        for (int j0 = 0; j0 < EV_BLOCK / EV_GROUP; j0 += 8) {
            uint4 a[8], b[8];
#pragma unroll
            for (int t = 0; t < 8; ++t)
                if (j0 + t < units) {
                    a[t] = __ldg(B.base + ((j0 + t) * 32 + B.slot) * 2);
                    b[t] = __ldg(B.base + ((j0 + t) * 32 + B.slot) * 2 + 1);
                }
#pragma unroll
            for (int t = 0; t < 8; ++t)
                if (j0 + t < units) {
                    mine[(j0 + t) * 2] = a[t];
                    mine[(j0 + t) * 2 + 1] = b[t];
                }
        }
        __syncwarp();
    }
    // The chunk's speculation record knows, for 8 start states of every block, where the block ends and how long its code
    // is.  None of lane 1's candidates was the true state (that is why the chunk is here), but the true trajectory usually
    // MERGES with a recorded one within a block or two (same class, section 4.2 of DESIGN.md): from then on every lane finds
    // the incoming state among its candidates and answers from the record.  Only the blocks before the merge are coded.
    uint4 *lr = lanerec + (size_t)lane * LANEREC_WORDS;
    const uint4 r_in = lr[0], r_bits = lr[1], r_out = lr[3];
    uint32_t s = start, my_in = start, my_bits = 0, total = 0, coded = 0;
    for (int l = 1; l < 32; ++l) {
        uint32_t out = s, b = 0;
        if (lane == l) {
            my_in = s;
            int hit = -1;
            if (has && use_record)
                for (int c = 7; c >= 0; --c)
                    if (half_of(r_in, c) == s && half_of(r_out, c) != STATE_INVALID) hit = c;
            if (hit >= 0) {
                out = half_of(r_out, hit);
                b = my_bits = half_of(r_bits, hit);
            } else if (has) {
                ++coded;
                State t[1] = {unpack_state(s)};
                uint32_t bb[1];
                if (staged) {
                    const uint2 *mine = stage + (size_t)lane * EV_BLOCK;
                    bool quirk = false;
                    bb[0] = 0;
                    for (int i = 0; i < B.cnt; ++i) {
                        const uint2 e = mine[i];
                        quirk = count_event(t[0], e.x, e.y, bb[0]);
                    }
                    if (B.is_tail) {
                        CountSink k{0};
                        code_tail(t[0], B.tail_zeros, quirk, k);
                        bb[0] += k.bits;
                    }
                } else {
                    count_lane<1>(B, t, bb);
                }
                out = pack_state(t[0]);
                b = bb[0];
                my_bits = b;
            }
        }
        s = __shfl_sync(~0u, out, l);
        total += __shfl_sync(~0u, b, l);
    }
    if (lane == 0) my_in = start;
    lr[0] = make_uint4(my_in & 0xffffu, 0u, 0u, 0u);
    lr[1] = make_uint4(lane == 0 ? 0u : (my_bits & 0xffffu), 0u, 0u, 0u);
    lr[2] = make_uint4(0u, 0u, 0u, 0u);
    lr[3] = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
    for (int o = 16; o; o >>= 1) coded += __shfl_xor_sync(~0u, coded, o);
    blocks_coded = (int)coded;
    if (lane == 0) {
        rec[0] = make_uint4((start & 0xffffu) | (STATE_INVALID << 16), ~0u, ~0u, ~0u);
        rec[1] = make_uint4(s & 0xffffu, 0u, 0u, 0u);
        rec[2] = make_uint4(total, 0u, 0u, 0u);
        rec[3] = make_uint4(0u, 0u, 0u, 0u);
    }
    end_state = s;
    total_bits = total;
}

// ---- B: resolve ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32 * SOLVE_WARPS) rlgr_resolve_kernel(Geometry G, const ChanInfo *__restrict__ info,
                                                          const uint2 *__restrict__ ev, uint4 *__restrict__ rec,
                                                          uint4 *__restrict__ lanerec, unsigned long long *__restrict__ chunkoff,
                                                          int64_t slot_bytes, int64_t *__restrict__ nbytes,
                                                          unsigned long long *__restrict__ totals, int32_t *status,
                                                          unsigned long long *stats, const uint32_t *__restrict__ headers,
                                                          const uint2 *__restrict__ ctab, uint32_t *__restrict__ cstart)
{
    pdl_trigger();
    const int ch = blockIdx.x;
    const int lane = threadIdx.x & 31;
    const ChanInfo I = info[ch];
    const int64_t nchunks = I.nchunks;
    const bool dense = channel_is_dense(I, G);
    __shared__ Helpers helpers;                          // warps 1..7: single-state coders for the re-solves of a sparse channel
    __shared__ uint2 walk_stage[32 * EV_BLOCK];          // the events of a chunk under repair (walk_chunk_exact)
    if (threadIdx.x >= 32) {
        if (!dense) helper_loop(&helpers, ev, G, ch, I);
        return;
    }
    unsigned long long table_hits = 0;
    uint32_t ps = pack_state(initial_state());
    unsigned long long off = 0;
    unsigned long long repairs = 0, repair_rounds = 0;
    for (int64_t base = 0; base < nchunks; base += 32) {
        // lane l holds the record of chunk base + l.  The only serial part is the chain of states: at step i every lane
        // looks the state up in its own record and lane i's answer is broadcast (one shuffle per chunk); which
        // candidate a chunk took, its code length and its start state stay in its lane, and the bit offsets of the
        // batch come from one prefix sum afterwards.
        const int64_t mine = base + lane < nchunks ? base + lane : nchunks - 1;
        const size_t cm = (size_t)ch * G.W + mine;
        const uint4 *mr = rec + cm * REC_WORDS;
        uint4 m0 = mr[0], m1 = mr[1], m2 = mr[2], m3 = mr[3];
        const int cnt = (int)min((int64_t)32, nchunks - base);
        // sparse channel: the table window of every chunk of the batch, one load per lane (inside the serial walk below it was
        // a dependent round trip in front of the table entry's, for every chunk that needs its table)
        const uint32_t my_window = dense ? 0u : headers[(size_t)ch * G.nsb + (size_t)mine * CHUNK_LANES];
        uint32_t my_sel = 0, my_bits = 0, my_start = 0;              // my_start: 0 = settled by a record, else state | 1 << 16 (table)
        for (int i = 0; i < cnt; ++i) {
            int b = -1;
#pragma unroll
            for (int c = 7; c >= 0; --c) b = half_of(m0, c) == ps ? c : b;
            uint32_t e = b >= 0 ? half_of(m1, b) : STATE_INVALID;
            uint32_t bits = b >= 0 ? word_of(m2, m3, b) : 0u;
            uint32_t e_i = __shfl_sync(~0u, e, i);
            uint32_t start_flag = 0;
            if (e_i == STATE_INVALID) {                              // (warp-uniform) no candidate of chunk i starts in the true state
                const int64_t w = base + i;
                const size_t ci = (size_t)ch * G.W + w;
                bool settled = false;
                if (!dense) {                                        // sparse channel: the chunk's transition table
                    const int idx = tab_index(__shfl_sync(~0u, my_window, i), ps);
                    uint2 te = make_uint2(TAB_DEAD, 0u);
                    if (idx >= 0) te = ctab[ci * TAB_STATES + idx];
                    if (te.x != TAB_DEAD) {
                        e_i = te.x;
                        if (lane == i) { b = 0; bits = te.y; start_flag = ps | (1u << 16); }
                        ++table_hits;
                        settled = true;
                    }
                }
                if (!settled) {                                      // solve this chunk again from the true state
                    const LaneBlock B = lane_block(ev, G, ch, w, I);
                    uint4 *lr = lanerec + ci * 32 * LANEREC_WORDS;
                    int rounds;
                    uint32_t pins[8];
                    for (int c = 0; c < 8; ++c) pins[c] = ps;
                    if (dense) {
                        ChunkSol<1> S;
                        solve_chunk<1>(B, true, pins, 1u, PIN_ROUNDS, S);
                        finish_chunk<1>(S, rec + ci * REC_WORDS, lr);
                        rounds = S.rounds;
                    } else {
                        uint32_t e_x, t_x;
                        walk_chunk_exact(B, ps, rec + ci * REC_WORDS, lr, e_x, t_x, walk_stage, true, rounds);
                    }
                    __syncwarp();
                    const uint4 *rr = rec + ci * REC_WORDS;          // (written by lane 0 just above): the pinned chain is candidate 0
                    const uint4 n1 = rr[1], n2 = rr[2];
                    e_i = half_of(n1, 0);
                    if (lane == i) { m0 = rr[0]; m1 = n1; m2 = n2; m3 = rr[3]; b = 0; bits = n2.x; }
                    ++repairs;
                    repair_rounds += (unsigned long long)rounds;
                }
            }
            if (lane == i) { my_sel = (uint32_t)b; my_bits = bits; my_start = start_flag; }
            ps = e_i;
        }
        // bit offsets of the batch
        uint32_t incl = lane < cnt ? my_bits : 0u;
        unsigned long long run = incl;
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long y = __shfl_up_sync(~0u, run, o);
            if (lane >= o) run += y;
        }
        if (lane < cnt) {
            chunkoff[cm] = ((off + run - incl) << 3) | (my_sel & 7u);
            if (!dense) cstart[cm] = my_start;
        }
        off += __shfl_sync(~0u, run, 31);
    }
    if (!dense) helpers_release(&helpers);
    if (lane == 0) {
        const int64_t nb = (int64_t)((off + 7) >> 3);
        nbytes[ch] = nb;
        totals[ch] = off;
        if (nb > slot_bytes && status) atomicOr(status, RAHT_ST_STREAM_OVERFLOW);
        if (stats) { atomicAdd(stats + 2, repairs); atomicAdd(stats + 4, repair_rounds); atomicAdd(stats + 5, table_hits); }
    }
}

// zero the words the emit pass will OR into: ceil(bits / 64) 8-byte words per channel, capped to the slot
__global__ void __launch_bounds__(256) rlgr_zero_kernel(uint8_t *__restrict__ streams, int64_t slot_bytes,
                                                        const unsigned long long *__restrict__ totals)
{
    pdl_enter();
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
    uint32_t cap_words;   // (a slot holds < 2^31 words: n < 2^31 symbols, 4 n + 64 bytes at most)
    uint32_t wi;          // index of the word being assembled
    uint64_t acc;
    unsigned nb;          // bits in acc (the first word starts with begin % 32 zero bits that are not ours)
    bool shared_word;     // the word being assembled has bits of the previous lane
    bool overflow;
    __device__ __forceinline__ void init(uint8_t *stream, int64_t slot_bytes, unsigned long long begin_bit)
    {
        words = (uint32_t *)stream;
        cap_words = (uint32_t)(slot_bytes >> 2);
        wi = (uint32_t)(begin_bit >> 5);
        nb = (unsigned)(begin_bit & 31);
        acc = 0;
        shared_word = nb != 0;
        overflow = false;
    }
    __device__ __forceinline__ unsigned long long bit_pos() const { return ((unsigned long long)wi << 5) + nb; }
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

__global__ void __launch_bounds__(32 * SPEC_WARPS) rlgr_emit_kernel(Geometry G, const ChanInfo *__restrict__ info,
                                                                   const uint2 *__restrict__ ev,
                                                                   const uint32_t *__restrict__ blockpos,
                                                                   const uint4 *__restrict__ lanerec,
                                                                   const unsigned long long *__restrict__ chunkoff,
                                                                   uint8_t *__restrict__ streams, int64_t slot_bytes,
                                                                   uint2 *__restrict__ sync_index, int64_t sync_entries,
                                                                   int32_t *status)
{
    pdl_enter();
    const int ch = blockIdx.y;
    const int64_t w = (int64_t)blockIdx.x * SPEC_WARPS + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    const ChanInfo I = info[ch];
    if (w >= (int64_t)I.nchunks) return;
    const size_t ci = (size_t)ch * G.W + w;
    const LaneBlock B = lane_block(ev, G, ch, w, I);
    const unsigned long long co = chunkoff[ci];
    const int cb = (int)(co & 7);
    const uint4 *lr = lanerec + (ci * 32 + lane) * LANEREC_WORDS;
    const int sel = (int)((lr[2].x >> (4 * cb)) & 7);      // this lane's candidate on the chosen chain
    State s = unpack_state(half_of(lr[0], sel));
    const uint32_t mybits = lane == 0 ? 0u : half_of(lr[1], sel);
    uint32_t x = mybits;                                   // inclusive prefix over the lanes
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(~0u, x, o);
        if (lane >= o) x += y;
    }
    const unsigned long long begin = (co >> 3) + (x - mybits);
    if (lane == 0 || mybits == 0) return;
    BitSink sink;
    sink.init(streams + (size_t)ch * slot_bytes, slot_bytes, begin);
    uint2 *sync = sync_index ? sync_index + (size_t)ch * sync_entries : nullptr;
    bool quirk = false;
    // first event of its 256-symbol range: where (bit offset, state) the coder stands before it
    auto sync_entry = [&](int64_t np) {
        const unsigned long long at = sink.bit_pos();
        const uint32_t hi = (uint32_t)((at >> 32) & 0x7f) | ((uint32_t)(s.kP & 0x7f) << 7) |
                            ((uint32_t)(s.kRP & 0xff) << 14) | ((uint32_t)(np & (SYNC_RANGE - 1)) << 22);
        sync[np >> SYNC_SHIFT] = make_uint2((uint32_t)at, hi);
    };
    if (B.sym) {                                           // dense channel: symbol by symbol, straight from q
        // sync index: the block's first event is the first of its 256-symbol range iff the zeros pending in front of the
        // block reach back to the start of the range (ranges start at block boundaries); the coder stands before that
        // event in the block's start state at the block's first bit
        bool want_sync = sync && B.cnt > 0 && B.gap0 >= (uint32_t)(B.pos0 & (SYNC_RANGE - 1));
        const unsigned long long at0 = begin;
        const State s0 = s;
        SymState t;
        t.kP = s.kP; t.kRP = s.kRP; t.m = 0;
        if (B.gap0 && B.cnt > 0) sym_begin(t, B.gap0, sink);
        walk_symbols(B, [&](int i, uint32_t u) {
            if (want_sync && u != 0) {
                const int64_t np = B.pos0 + i;
                const uint32_t hi = (uint32_t)((at0 >> 32) & 0x7f) | ((uint32_t)(s0.kP & 0x7f) << 7) |
                                    ((uint32_t)(s0.kRP & 0xff) << 14) | ((uint32_t)(np & (SYNC_RANGE - 1)) << 22);
                sync[np >> SYNC_SHIFT] = make_uint2((uint32_t)at0, hi);
                want_sync = false;
            }
            sym_emit(t, u, sink);
        });
        quirk = B.cnt > 0 && sym_tail_quirk(t, last_mapped(B));
        s.kP = t.kP; s.kRP = t.kRP;
    } else {
    int64_t pos = -1;                                      // symbol index of the previous event
    const int units = (B.cnt + EV_GROUP - 1) / EV_GROUP;
    uint4 a1, b1, a2, b2;                                   // two units requested ahead, as in count_lane
    a1 = b1 = a2 = b2 = make_uint4(0u, 0u, 0u, 0u);
    if (units > 0) { a1 = __ldg(B.base + lane * 2); b1 = __ldg(B.base + lane * 2 + 1); }
    if (units > 1) { a2 = __ldg(B.base + (32 + lane) * 2); b2 = __ldg(B.base + (32 + lane) * 2 + 1); }
    for (int j = 0; j < units; ++j) {
        const uint4 a = a1, b = b1;
        a1 = a2; b1 = b2;
        if (j + 2 < units) { a2 = __ldg(B.base + ((j + 2) * 32 + lane) * 2); b2 = __ldg(B.base + ((j + 2) * 32 + lane) * 2 + 1); }
        const uint32_t g[4] = {a.x, a.z, b.x, b.z}, u[4] = {a.y, a.w, b.y, b.w};
        if (j == 0) pos = (int64_t)blockpos[(size_t)ch * G.nsb + w * CHUNK_LANES - 1 + lane] - (int64_t)g[0] - 1;
#pragma unroll
        for (int t = 0; t < 4; ++t)
            if (j * EV_GROUP + t < B.cnt) {
                const int64_t np = pos + (int64_t)g[t] + 1;
                if (sync && (np >> SYNC_SHIFT) != (pos >> SYNC_SHIFT)) sync_entry(np);
                pos = np;
                quirk = emit_event(s, g[t], u[t], sink);
            }
    }
    }
    if (B.is_tail) code_tail(s, B.tail_zeros, quirk, sink);
    sink.finish();
    if (sink.overflow && status) atomicOr(status, RAHT_ST_STREAM_OVERFLOW);
}

// Golomb-Rice symbol whose whole code (p ones, a zero, kR bits) lies in the top 32 bits of the window: ~12 dependent
// instructions instead of the general reader's two refills, escape test and 64-bit arithmetic.  false (nothing
// consumed): an escape (p >= 32) or a code longer than 32 bits -- the caller takes the general path.
template <class Reader>
__device__ __forceinline__ bool gr_fast(Reader &br, unsigned kR, uint32_t &u, uint32_t &p_out, unsigned &len_out)
{
    br.fill();                                                             // >= 33 valid bits
    const uint32_t hi = (uint32_t)(br.acc >> 32), lo = (uint32_t)br.acc;
    const unsigned p = (unsigned)__clz((int)~hi);                          // 32 when the top word is all ones
    const unsigned len = p + 1u + kR;
    if (len > 32u) return false;
    const uint32_t rest = __funnelshift_lc(lo, hi, p + 1u);                // the window behind the prefix
    u = (p << kR) + __funnelshift_rc(rest, 0u, 32u - kR);                  // kR = 0: shift by 32 -> 0
    br.acc <<= len;
    br.nbits -= (int)len;
    p_out = p;
    len_out = len;
    return true;
}
__device__ __forceinline__ void adapt_krp_small(State &s, uint32_t p)      // adapt_krp for p < 32
{
    const int t = s.kRP + (p ? (int)p - 1 : -2);
    s.kRP = min(max(t, 0), RL_KRP_MAX);
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

// A stream without a sync index (written by the reference coder) in two passes: this kernel WALKS it once -- bit
// positions and coder states only, no symbol leaves the thread -- and writes the sync index the encoder would have
// written next to it; rlgr_decode_indexed_kernel then decodes every 256-symbol range in parallel.  The walk is the
// serial part the format forces (the state is carried from symbol to symbol, membuf.cpp:266-334), so it is made as
// cheap per symbol as it gets: a warp per channel, the lanes copy the stream into a ring in shared memory ahead of
// lane 0, which parses out of the ring (no global round trip per refill, no scattered stores).
constexpr uint32_t IDX_RING = 4096, IDX_MARGIN = 8;                       // words; one decode step reads < 8 words
struct RingWords {
    const uint32_t *ring;
    uint32_t pos, ahead;                                                   // pos: index of `ahead` = words handed out
    __device__ __forceinline__ void open(const uint32_t *r)
    {
        ring = r;
        pos = 0;
        ahead = ring[0];
    }
    __device__ __forceinline__ uint32_t next()
    {
        const uint32_t w = rl_bswap32(ahead);
        ++pos;
        ahead = ring[pos & (IDX_RING - 1)];
        return w;
    }
};
__global__ void __launch_bounds__(32) rlgr_index_kernel(const uint8_t *__restrict__ streams, int64_t slot_bytes,
                                                       const int64_t *__restrict__ nbytes, Geometry G,
                                                       uint2 *__restrict__ sync_index, int64_t entries)
{
    __shared__ uint32_t ring[IDX_RING];
    const int ch = blockIdx.x, lane = threadIdx.x;
    int64_t nb = nbytes ? nbytes[ch] : slot_bytes;
    if (nb > slot_bytes) nb = slot_bytes;
    if (nb < 0) nb = 0;
    const uint32_t n_words = (uint32_t)((nb + 3) >> 2);
    const uint32_t *in = (const uint32_t *)(streams + (size_t)ch * slot_bytes);
    uint2 *ent = sync_index + (size_t)ch * entries;
    for (int64_t e = lane; e < entries; e += 32) ent[e] = make_uint2(0u, SYNC_NONE << 22);
    BitReaderT<RingWords> br;
    State s = initial_state();
    const uint32_t n = (uint32_t)G.n;
    uint32_t pos = 0, loaded = 0, rd = 0;
    int64_t marked = -1;                                                   // last range that has its entry
    bool opened = false;
    for (;;) {
        const uint32_t hi = rd + IDX_RING - 32;
        for (uint32_t i = loaded + lane; i < hi; i += 32) ring[i & (IDX_RING - 1)] = i < n_words ? in[i] : 0u;
        loaded = hi;
        __syncwarp();
        int done = 0;
        if (lane == 0) {
            if (!opened) {
                br.src.open(ring);
                br.acc = 0;
                br.nbits = 0;
                opened = true;
            }
            while (pos < n && br.src.pos + IDX_MARGIN < loaded) {
                if (s.kP < RL_L) {                                         // no-run mode: one symbol, the lean way
                    uint32_t u, p;
                    unsigned len;
                    if (gr_fast(br, (unsigned)s.kRP / RL_L, u, p, len)) {
                        if (u) {
                            const int64_t r = (int64_t)(pos >> SYNC_SHIFT);
                            if (r != marked) {
                                marked = r;
                                const unsigned long long at = (unsigned long long)br.src.pos * 32ull - (unsigned long long)(br.nbits + (int)len);
                                ent[r] = make_uint2((uint32_t)at, (uint32_t)((at >> 32) & 0x7f) | ((uint32_t)s.kP << 7) |
                                                                      ((uint32_t)s.kRP << 14) | ((pos & (SYNC_RANGE - 1)) << 22));
                            }
                            s.kP = max(s.kP - RL_D0, 0);
                        } else {
                            s.kP += RL_U0;
                        }
                        adapt_krp_small(s, p);
                        ++pos;
                        continue;
                    }
                }
                const unsigned long long at_bit = (unsigned long long)br.src.pos * 32ull - (unsigned long long)br.nbits;
                const State before = s;
                int k = s.kP / RL_L;
                const unsigned kR = (unsigned)(s.kRP / RL_L);
                const bool run = k != 0;
                if (run) {
                    uint32_t m = 0;
                    bool bad = false;
                    for (;;) {
                        const int p = br.ones();
                        for (int j = 0; j < p; ++j) {
                            m += 1u << k;
                            s.kP += RL_U1;
                            k = s.kP / RL_L;
                            if (m >= (1u << 31) || k > 30) { bad = true; break; }
                        }
                        if (bad) break;
                        if (p < 32) { br.skip((unsigned)p + 1); break; }
                        br.skip(32);
                    }
                    if (!bad) m += br.get((unsigned)k);
                    if (bad || m >= n - pos) { pos = n; break; }           // zeros to the end (or a corrupt stream)
                    pos += m;
                }
                uint32_t u, pf;
                unsigned lf;
                if (gr_fast(br, kR, u, pf, lf)) {
                    adapt_krp_small(s, pf);
                } else {
                    u = br.gr(kR);
                    adapt_krp(s, (uint64_t)u >> kR);
                }
                if (run || u) {
                    const int64_t r = (int64_t)(pos >> SYNC_SHIFT);
                    if (r != marked) {
                        marked = r;
                        ent[r] = make_uint2((uint32_t)at_bit, (uint32_t)((at_bit >> 32) & 0x7f) | ((uint32_t)before.kP << 7) |
                                                                  ((uint32_t)before.kRP << 14) | ((pos & (SYNC_RANGE - 1)) << 22));
                    }
                    s.kP = s.kP < RL_D1 ? 0 : s.kP - RL_D1;
                } else {
                    s.kP += RL_U0;
                }
                ++pos;
            }
            done = pos >= n;
            rd = br.src.pos;
        }
        done = __shfl_sync(~0u, done, 0);
        rd = __shfl_sync(~0u, rd, 0);
        if (done) break;
    }
}

// with the encoder's sync index: one thread per entry (256 symbols)
constexpr int DEC_THREADS = 128;
// (Tried and dropped: feeding the window from a per-thread cp.async ring in shared memory instead of a register loaded
// one refill ahead.  It removes the long-scoreboard stalls -- a register loaded from global memory is a warp-wide
// scoreboard entry, so every refill waits for the load some OTHER lane issued an iteration earlier -- but the ring
// bookkeeping costs 34 % more instructions and the kernel got slower, 692 against 606 us.  Loading on a fixed beat --
// every second event all lanes merge the words they requested two events earlier into a 64-bit reserve and request
// the next two -- did not help either, 657 us: the compiler makes the loop's branch wait for the loads in flight.)
// One step of rlgr::decode_range (membuf.cpp:266-334): a zero run and its non-zero symbol in run mode, one symbol
// otherwise.  have = true: a non-zero symbol `val` at position `at` (relative to the range) came out of it.
__device__ __forceinline__ void decode_step(BitReader &br, State &s, bool &synced, uint32_t first, uint32_t &pos, uint32_t end,
                                            int flag_signed, bool &have, uint32_t &at, int32_t &val)
{
    int k = s.kP / RL_L;
    const unsigned kR = (unsigned)(s.kRP / RL_L);
    const bool run = k != 0;
    if (run) {
        uint32_t m = 0;
        bool bad = false;
        for (;;) {                                                         // complete runs: one 1 bit each
            const int p = br.ones();
            for (int j = 0; j < p; ++j) {
                m += 1u << k;
                s.kP += RL_U1;
                k = s.kP / RL_L;
                if (m >= (1u << 31) || k > 30) { bad = true; break; }      // corrupt stream guard
            }
            if (bad) break;
            if (p < 32) { br.skip((unsigned)p + 1); break; }
            br.skip(32);
        }
        if (!bad) m += br.get((unsigned)k);
        if (bad || (synced && m >= end - pos)) { pos = end; return; }      // the next non-zero lies past this range
        if (synced) pos += m;
        else { pos = first; synced = true; }
    }
    uint32_t u, pf;
    unsigned lf;
    if (gr_fast(br, kR, u, pf, lf)) {
        adapt_krp_small(s, pf);
    } else {
        u = br.gr(kR);
        adapt_krp(s, (uint64_t)u >> kR);
    }
    if (run || u) {
        if (!synced) { pos = first; synced = true; }
        const uint64_t v = (uint64_t)u + (run ? 1u : 0u);
        have = true;
        at = pos++;
        val = flag_signed ? u2s(v) : (int32_t)v;
        s.kP = s.kP < RL_D1 ? 0 : s.kP - RL_D1;
    } else {
        s.kP += RL_U0;
        if (synced) ++pos;
    }
}

// One thread per sync entry (256 symbols), but the symbols do not leave the thread one by one: 59 M four-byte stores,
// every lane of a warp to a different cache line, were what bound the kernel (53 M sectors through the L1 tag stage for
// 236 MB of output; doubling the threads did not make it faster).  The range is decoded in windows of 32 symbols into
// a row of shared memory per lane; after each window the warp writes its 32 rows out as 32 full lines -- zeros
// included, so the output needs no memset.
constexpr int DEC_WIN = 32, DEC_PITCH = DEC_WIN + 1;
__global__ void __launch_bounds__(DEC_THREADS) rlgr_decode_indexed_kernel(const uint8_t *__restrict__ streams, int64_t slot_bytes,
                                                                         const int64_t *__restrict__ nbytes,
                                                                         const uint2 *__restrict__ sync_index, int64_t entries,
                                                                         Geometry G, int32_t *__restrict__ q)
{
    __shared__ int32_t stage[DEC_THREADS / 32][32][DEC_PITCH];
    const int ch = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t e0 = (int64_t)blockIdx.x * DEC_THREADS + warp * 32;      // first entry of this warp
    const int64_t e = e0 + lane;
    const int64_t start = e << SYNC_SHIFT;
    const uint32_t end = e < entries ? (uint32_t)min((int64_t)SYNC_RANGE, G.n - start) : 0u;
    uint32_t first = SYNC_NONE;
    uint2 ent = make_uint2(0u, 0u);
    if (e < entries) {
        ent = sync_index[(size_t)ch * entries + e];
        first = ent.y >> 22;
    }
    BitReader br;
    State s;
    s.kP = (int)((ent.y >> 7) & 0x7f);
    s.kRP = (int)((ent.y >> 14) & 0xff);
    bool synced = false, have = false;
    uint32_t pos = 0, at = 0;
    int32_t val = 0;
    if (first != SYNC_NONE && first < end) {                               // else: nothing but zeros in this range
        int64_t nb = nbytes ? nbytes[ch] : slot_bytes;
        if (nb > slot_bytes) nb = slot_bytes;
        br.init(streams + (size_t)ch * slot_bytes, nb, (unsigned long long)ent.x | ((unsigned long long)(ent.y & 0x7f) << 32));
    } else {
        pos = end;
    }
    int32_t *row = stage[warp][lane];
    for (int i = 0; i < DEC_WIN; ++i) row[i] = 0;
    __syncwarp();
    int32_t *qch = q + (size_t)ch * G.ld;
    const int rows = (int)min((int64_t)32, entries - e0);                  // ranges of this warp that exist
    for (uint32_t w0 = 0; w0 < (uint32_t)SYNC_RANGE; w0 += DEC_WIN) {
        if (!__any_sync(~0u, w0 < end)) break;                             // past the end of the channel
        for (;;) {                                                         // one loop, one decode step per turn: the lanes stay together
            if (have && at < w0 + DEC_WIN) { row[at - w0] = val; have = false; }
            const bool more = !have && pos < end;
            if (!__any_sync(~0u, more)) break;
            if (more) decode_step(br, s, synced, first, pos, end, G.flag_signed, have, at, val);
        }
        __syncwarp();
        for (int r = 0; r < rows; ++r) {
            const int64_t g = ((e0 + r) << SYNC_SHIFT) + w0 + lane;
            const int32_t v = stage[warp][r][lane];
            stage[warp][r][lane] = 0;
            if (g < G.n) qch[g] = v;
        }
        __syncwarp();
    }
}

struct EncLayout {
    size_t off_lastnz, off_cnt, off_info, off_blockpos, off_rec, off_rec2, off_lanerec, off_chunkoff, off_totals, off_stats, off_ev, off_tables, off_headers, off_ctab, off_cstart, total;
};
EncLayout enc_layout(int64_t n, int nch)
{
    EncLayout L;
    const size_t nsb = (size_t)sym_blocks_of(n), W = (size_t)chunk_slots_of(n), c = (size_t)(nch > 0 ? nch : 0);
    size_t o = 0;
    L.off_lastnz = o; o += align_up(c * nsb * 4, 256);
    L.off_cnt = o; o += align_up(c * nsb * 4, 256);
    L.off_info = o; o += align_up(c * sizeof(ChanInfo), 256);
    L.off_blockpos = o; o += align_up(c * nsb * 4, 256);
    L.off_rec = o; o += align_up(c * W * REC_WORDS * 16, 256);
    L.off_rec2 = o; o += align_up(c * W * REC_WORDS * 16, 256);
    L.off_lanerec = o; o += align_up(c * W * 32 * LANEREC_WORDS * 16, 256);
    L.off_chunkoff = o; o += align_up(c * W * 8, 256);
    L.off_totals = o; o += align_up(c * 8, 256);
    L.off_stats = o; o += 256;
    L.off_ev = o; o += align_up(c * W * CHUNK_SLOTS * 8, 256);
    L.off_tables = o; o += align_up(c * nsb * TAB_STATES * 4, 256);
    L.off_headers = o; o += align_up(c * nsb * 4, 256);
    L.off_ctab = o; o += align_up(c * W * TAB_STATES * 8, 256);
    L.off_cstart = o; o += align_up(c * W * 4, 256);
    L.total = o + 256;
    return L;
}

}  // namespace

RAHT_API int64_t raht_rlgr_sync_entries(int64_t n) { return n <= 0 ? 0 : (n + SYNC_RANGE - 1) >> SYNC_SHIFT; }

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
    if (!workspace || workspace_bytes < L.total || ((uintptr_t)workspace & 15) != 0) {
        raht_set_error("raht_rlgr_encode: workspace too small or not 16-byte aligned (%zu < %zu)", workspace_bytes, L.total);
        return RAHT_E_WORKSPACE;
    }
    Geometry G;
    G.n = n; G.ld = ld; G.nsb = sym_blocks_of(n); G.W = chunk_slots_of(n); G.nch = nch; G.flag_signed = flag_signed;
    G.q = q;
    G.vec = (((uintptr_t)q & 15) == 0 && (ld % 4) == 0) ? 1 : 0;
    char *ws = (char *)workspace;
    int32_t *lastnz = (int32_t *)(ws + L.off_lastnz);
    G.lastnz = lastnz;
    uint32_t *cnt = (uint32_t *)(ws + L.off_cnt);
    ChanInfo *info = (ChanInfo *)(ws + L.off_info);
    uint32_t *blockpos = (uint32_t *)(ws + L.off_blockpos);
    uint4 *rec0 = (uint4 *)(ws + L.off_rec), *rec = (uint4 *)(ws + L.off_rec2);
    uint4 *lanerec = (uint4 *)(ws + L.off_lanerec);
    unsigned long long *chunkoff = (unsigned long long *)(ws + L.off_chunkoff);
    unsigned long long *totals = (unsigned long long *)(ws + L.off_totals);
    unsigned long long *stats = (unsigned long long *)(ws + L.off_stats);
    uint2 *ev = (uint2 *)(ws + L.off_ev);
    uint32_t *tables = (uint32_t *)(ws + L.off_tables), *headers = (uint32_t *)(ws + L.off_headers);
    uint2 *ctab = (uint2 *)(ws + L.off_ctab);
    uint32_t *cstart = (uint32_t *)(ws + L.off_cstart);
    RAHT_CUDA(cudaMemsetAsync(stats, 0, 256, st));
    const int64_t entries = raht_rlgr_sync_entries(n);
    if (sync_index) RAHT_CUDA(cudaMemsetAsync(sync_index, 0xff, (size_t)nch * entries * 8, st));   // "no event in this range"

    const unsigned groups = (unsigned)((G.nsb + 32 * K1_WARPS - 1) / (32 * K1_WARPS));
    (raht_count_launches(1), rlgr_count_kernel)<<<dim3(groups, nch), 32 * K1_WARPS, 0, st>>>(q, G, cnt, lastnz);
    RAHT_PDL_LAUNCH(PDL_RLGR, rlgr_scan_kernel, nch, 1024, 0, st, G, cnt, lastnz, info);
    RAHT_PDL_LAUNCH(PDL_RLGR, rlgr_scatter_kernel, dim3(groups, nch), 32 * K1_WARPS, 0, st, q, G, (const uint32_t *)cnt, (const int32_t *)lastnz, ev, blockpos,
                    (const ChanInfo *)info);
    const dim3 cgrid((unsigned)((G.W + SPEC_WARPS - 1) / SPEC_WARPS), nch);
    const dim3 sgrid((unsigned)G.W, nch, 2);                 // z = 0: a sparse chunk per CTA; z = 1: 8 dense chunks per CTA (the rest exit)
    RAHT_PDL_LAUNCH(PDL_RLGR, rlgr_speculate_kernel, sgrid, 32 * SOLVE_WARPS, 0, st, G, (const ChanInfo *)info, (const uint2 *)ev, 0, (const uint4 *)nullptr, rec0,
                    lanerec, stats);
    const dim3 bgrid((unsigned)min((int64_t)192, (G.nsb + SPEC_WARPS - 1) / SPEC_WARPS), nch);    // grid-stride over the blocks
    static const int use_side = getenv("RAHT_RLGR_SIDE") ? atoi(getenv("RAHT_RLGR_SIDE")) : 1;      // 0: everything on the caller's stream (A/B)
    bool forked = false;
    if (use_side) {
        // The table passes of the sparse channels (a few latency-bound warps) and the second speculation sweep touch disjoint
        // records -- sweep 1 re-solves dense chunks only and copies the records of the sparse channels unchanged (rec0 ->
        // rec) -- so the tables read rec0 and run on the device's side stream next to the sweep.
        cudaStream_t side = raht_side_fork(st);
        if (side) {
            (raht_count_launches(1), rlgr_table_kernel)<<<bgrid, 32 * SPEC_WARPS, 0, side>>>(G, info, ev, rec0, lanerec, tables, headers);
            RAHT_PDL_LAUNCH(PDL_RLGR, rlgr_chunk_table_kernel, cgrid, 32 * SPEC_WARPS, 0, side, G, (const ChanInfo *)info, (const uint32_t *)tables,
                            (const uint32_t *)headers, ctab);
            (raht_count_launches(1), rlgr_speculate_kernel)<<<sgrid, 32 * SOLVE_WARPS, 0, st>>>(G, info, ev, 1, rec0, rec, lanerec, stats);
            RAHT_TRY(raht_side_join(st));
            forked = true;
        }
    }
    if (!forked) {
        (raht_count_launches(1), rlgr_speculate_kernel)<<<sgrid, 32 * SOLVE_WARPS, 0, st>>>(G, info, ev, 1, rec0, rec, lanerec, stats);
        (raht_count_launches(1), rlgr_table_kernel)<<<bgrid, 32 * SPEC_WARPS, 0, st>>>(G, info, ev, rec, lanerec, tables, headers);
        (raht_count_launches(1), rlgr_chunk_table_kernel)<<<cgrid, 32 * SPEC_WARPS, 0, st>>>(G, info, tables, headers, ctab);
    }
    (raht_count_launches(1), rlgr_resolve_kernel)<<<nch, 32 * SOLVE_WARPS, 0, st>>>(G, info, ev, rec, lanerec, chunkoff, slot_bytes, nbytes,
                                                                    totals, status, stats, headers, ctab, cstart);
    RAHT_PDL_LAUNCH(PDL_RLGR, rlgr_table_lanes_kernel, cgrid, 32 * SPEC_WARPS, 0, st, G, (const ChanInfo *)info, (const uint32_t *)tables, (const uint32_t *)headers,
                    (const uint32_t *)cstart, lanerec);
    RAHT_PDL_LAUNCH(PDL_RLGR, rlgr_zero_kernel, dim3(64, nch), 256, 0, st, streams, slot_bytes, (const unsigned long long *)totals);
    RAHT_PDL_LAUNCH(PDL_RLGR, rlgr_emit_kernel, cgrid, 32 * SPEC_WARPS, 0, st, G, (const ChanInfo *)info, (const uint2 *)ev, (const uint32_t *)blockpos,
                    (const uint4 *)lanerec, (const unsigned long long *)chunkoff, streams, slot_bytes, (uint2 *)sync_index, entries, status);
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

// diagnostics of the last raht_rlgr_encode on this workspace: out[0] = chunk solves in the speculation
// pass, out[1] = their fixed-point rounds, out[2] = chunks re-solved by the resolve pass, out[3] = slowest
// chunk of the speculation pass (SM cycles << 24 | channel << 8 | rounds), out[4] = rounds spent in those re-solves,
// out[5] = chunks settled through the transition tables.  Synchronises.
RAHT_API int raht_rlgr_encode_stats(const void *workspace, int64_t n, int nch, unsigned long long *out5_host, void *stream)
{
    RAHT_REQUIRE(workspace && out5_host && n > 0, "raht_rlgr_encode_stats: bad arguments");
    const EncLayout L = enc_layout(n, nch);
    RAHT_CUDA(cudaMemcpyAsync(out5_host, (const char *)workspace + L.off_stats, 48, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    RAHT_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return RAHT_OK;
}

RAHT_API int raht_rlgr_build_index(const uint8_t *streams, int64_t slot_bytes, const int64_t *nbytes, int64_t n, int nch,
                                   void *sync_index, void *stream)
{
    RAHT_REQUIRE(n >= 0 && n < ((int64_t)1 << 31) - 256 && nch >= 0 && nch < 65536 && slot_bytes >= 8 && (slot_bytes % 8) == 0,
                 "raht_rlgr_build_index: bad sizes (slot_bytes must be a positive multiple of 8)");
    if (nch == 0 || n == 0) return RAHT_OK;
    RAHT_REQUIRE(streams && sync_index, "raht_rlgr_build_index: null pointer");
    RAHT_REQUIRE(((uintptr_t)streams & 7) == 0 && ((uintptr_t)sync_index & 7) == 0, "raht_rlgr_build_index: pointers must be 8-byte aligned");
    Geometry G;
    G.n = n; G.ld = n; G.nsb = sym_blocks_of(n); G.W = chunk_slots_of(n); G.nch = nch; G.flag_signed = 0;
    G.q = nullptr; G.lastnz = nullptr; G.vec = 0;
    (raht_count_launches(1), rlgr_index_kernel)<<<nch, 32, 0, (cudaStream_t)stream>>>(streams, slot_bytes, nbytes, G, (uint2 *)sync_index,
                                                                                       raht_rlgr_sync_entries(n));
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}

static thread_local bool g_dec_share = false;
void raht_rlgr_decode_share_gpu() { g_dec_share = true; }      // (internal, common.cuh) applies to this thread's next raht_rlgr_decode

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
    G.n = n; G.ld = ld; G.nsb = sym_blocks_of(n); G.W = chunk_slots_of(n); G.nch = nch; G.flag_signed = flag_signed;
    G.q = nullptr; G.lastnz = nullptr; G.vec = 0;
    if (sync_index) {                                        // (writes every symbol: no memset)
        const int64_t entries = raht_rlgr_sync_entries(n);
        const dim3 grid((unsigned)((entries + DEC_THREADS - 1) / DEC_THREADS), nch);
        raht_count_launches(1);
        // At most k CTAs of the decoder per SM (by dynamic shared memory it does not use) when the caller builds the tree on
        // the side stream at the same time (raht_rlgr_decode_share_gpu, scene.cu): the decoder's grid is little more than one
        // wave that fills every register file, and the tree's kernels found no room next to it until it drained.  9 of 10:
        // the two decoder legs 0.465 -> 0.427 ms at 1 M (8: 0.437, 7: 0.444, 6: 0.503).  RAHT_DEC_CTAS overrides (0 = no limit).
        static const int env_ctas = getenv("RAHT_DEC_CTAS") ? atoi(getenv("RAHT_DEC_CTAS")) : -1;
        // (only while the grid is a few waves: a long decode leaves gaps between its waves anyway, and the 6 M scene lost
        // 0.08 ms to the limit)
        const bool few_waves = (int64_t)grid.x * grid.y <= (int64_t)3 * 10 * sm_count_cached();
        const int dec_ctas = !g_dec_share ? 0 : env_ctas >= 0 ? env_ctas : (few_waves ? 9 : 0);
        g_dec_share = false;
        size_t pad = 0;
        if (dec_ctas > 0) {
            const size_t per = (size_t)227 * 1024 / (size_t)dec_ctas, fixed = sizeof(int32_t) * (DEC_THREADS / 32) * 32 * DEC_PITCH + 1024;
            pad = per > fixed + 1024 ? per - fixed - 512 : 0;
            if (pad > 200 * 1024) pad = 200 * 1024;
            RAHT_CUDA(cudaFuncSetAttribute(rlgr_decode_indexed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pad));
        }
        rlgr_decode_indexed_kernel<<<grid, DEC_THREADS, pad, st>>>(streams, slot_bytes, nbytes, (const uint2 *)sync_index, entries, G, q);
    } else {
        RAHT_CUDA(cudaMemsetAsync(q, 0, ((size_t)(nch - 1) * ld + n) * 4, st));  // the serial decoder stores non-zeros only
        (raht_count_launches(1), rlgr_decode_serial_kernel)<<<nch, 32, 0, st>>>(streams, slot_bytes, nbytes, G, q);
    }
    RAHT_LAUNCH_CHECK();
    return RAHT_OK;
}
