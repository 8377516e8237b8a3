// rlgr_emu.cpp -- host emulation of the chunk-parallel RLGR encoder / indexed decoder of
// raht-3dgs-codec_b200/csrc/rlgr.cu.  TEST INFRASTRUCTURE: it runs the same arithmetic
// (csrc/rlgr_core.cuh, compiled for the host) through the same passes -- last-nonzero scan,
// speculation sweeps with the in-warp fixed point, resolve, emit -- with the 32 lanes of a warp
// executed one after the other, so that the CPU test-suite can check the ALGORITHM (event alignment,
// tail rule, speculation + repair exactness, sync index) against the oracle without a GPU.
// The CUDA kernels themselves are checked on the GPU by tests/test_gpu_parity.py.
#include "../../raht-3dgs-codec_b200/csrc/rlgr_core.cuh"
#include <algorithm>
#include <cstring>
#include <vector>

using namespace rlgr;

namespace {

constexpr int SYNC_STRIDE = 4;                                                // as in rlgr.cu
constexpr int SYNC_PER_CHUNK = (CHUNK_LANES + SYNC_STRIDE - 1) / SYNC_STRIDE;
constexpr uint32_t SYNC_NONE = 1023;

struct Lane {
    std::vector<uint32_t> row;
    uint32_t mask[MASK_WORDS];
    int len;
    uint32_t zeros_before;
    bool is_tail;
    int last_nz;
    bool dense, single;
};

struct Chan {
    const int32_t *q;
    int64_t n;
    int flag_signed;
    int64_t nblk, nchunks;
    std::vector<int32_t> lastnz;
};

void load_chunk(const Chan &C, int64_t w, Lane L[32])
{
    for (int l = 0; l < 32; ++l) {
        const int64_t blk = w * CHUNK_LANES - 1 + l, start = blk * BLOCK_SYMS;
        Lane &B = L[l];
        B.row.assign(BLOCK_SYMS, 0);
        std::memset(B.mask, 0, sizeof(B.mask));
        for (int i = 0; i < BLOCK_SYMS; ++i) {
            const int64_t s = start + i;
            const int32_t v = (s >= 0 && s < C.n) ? C.q[s] : 0;
            const uint32_t u = C.flag_signed ? s2u(v) : (uint32_t)v;
            B.row[i] = u;
            if (u) B.mask[i / 32] |= 1u << (i % 32);
        }
        int64_t len = blk < 0 ? 0 : C.n - start;
        len = std::max<int64_t>(0, std::min<int64_t>(len, BLOCK_SYMS));
        B.len = (int)len;
        B.zeros_before = (blk <= 0 || len == 0) ? 0u : (uint32_t)(start - 1 - (int64_t)C.lastnz[blk - 1]);
        B.is_tail = len > 0 && start + len == C.n;
    }
    uint32_t nz = 0;
    for (int l = 0; l < 32; ++l) {
        L[l].last_nz = -1;
        for (int i = 0; i < BLOCK_SYMS; ++i)
            if (L[l].row[i]) { ++nz; L[l].last_nz = i; }
    }
    for (int l = 0; l < 32; ++l) {
        L[l].dense = nz * 4 >= (uint32_t)(CHUNK_LANES * BLOCK_SYMS);
        L[l].single = nz * 2 >= (uint32_t)(CHUNK_LANES * BLOCK_SYMS);
    }
}

struct LaneRes {
    State in[NCAND], out[NCAND];
    uint32_t bits[NCAND];
};

int solve_chunk(const Lane L[32], bool pin, const State first[NCAND], LaneRes R[32], bool *identity)
{
    bool need[32], has[32];
    int src[32];
    bool any_has = false;
    for (int l = 0; l < 32; ++l) {
        has[l] = L[l].is_tail;
        for (int w = 0; w < MASK_WORDS; ++w) has[l] |= L[l].mask[w] != 0;
        if (l == 0) has[l] = !pin && has[l];
        src[l] = 0;
        for (int k = 0; k < l; ++k) if (has[k]) src[l] = k;
        if (l > 0) any_has |= has[l];
        for (int c = 0; c < NCAND; ++c) { R[l].in[c] = first[c]; R[l].out[c] = first[c]; R[l].bits[c] = 0; }
        need[l] = has[l];
    }
    int r = 0;
    for (; r < 34; ++r) {
        for (int l = 0; l < 32; ++l) {
            if (!need[l]) continue;
            if (L[l].dense) {
                for (int c = 0; c < NCAND; ++c) R[l].out[c] = R[l].in[c];
                count_block_dense<NCAND>(L[l].row.data(), L[l].len, L[l].last_nz, L[l].zeros_before, L[l].is_tail, R[l].out,
                                         R[l].bits);
                continue;
            }
            for (int c = 0; c < NCAND; ++c) {
                State s = R[l].in[c];
                CountSink k{0};
                code_block(L[l].row.data(), L[l].len, L[l].mask, L[l].zeros_before, L[l].is_tail, s, k);
                R[l].out[c] = s;
                R[l].bits[c] = k.bits;
            }
        }
        bool any = false;
        State prev_out[32][NCAND];
        for (int l = 0; l < 32; ++l) for (int c = 0; c < NCAND; ++c) prev_out[l][c] = R[l].out[c];
        for (int l = 0; l < 32; ++l) {
            need[l] = false;
            for (int c = 0; c < NCAND; ++c)
                if (l > 0 && prev_out[src[l]][c] != R[l].in[c]) {
                    R[l].in[c] = prev_out[src[l]][c];
                    if (has[l]) need[l] = true;
                    else R[l].out[c] = R[l].in[c];
                }
            any |= need[l];
        }
        if (!any) break;
    }
    {
        State prev_out[32][NCAND];
        for (int l = 0; l < 32; ++l) for (int c = 0; c < NCAND; ++c) prev_out[l][c] = R[l].out[c];
        for (int l = 1; l < 32; ++l)
            if (!has[l]) for (int c = 0; c < NCAND; ++c) { R[l].in[c] = prev_out[src[l]][c]; R[l].out[c] = R[l].in[c]; }
    }
    R[0].bits[0] = R[0].bits[1] = 0;
    if (identity) *identity = !any_has;
    return r + 1;
}

constexpr uint32_t REC_IDENTITY = 0x80000000u;
struct ChunkRec {
    uint32_t cand, ends, bits0, bits1;
};
struct LaneRec {
    uint32_t starts, bits;
};

ChunkRec finish_chunk(const LaneRes R[32], bool identity, LaneRec *lr)
{
    ChunkRec rec{0, 0, 0, 0};
    for (int l = 0; l < 32; ++l) {
        lr[l].starts = pack_state(R[l].in[0]) | (pack_state(R[l].in[1]) << 16);
        lr[l].bits = R[l].bits[0] | (R[l].bits[1] << 16);
        rec.bits0 += R[l].bits[0];
        rec.bits1 += R[l].bits[1];
    }
    rec.cand = lr[1].starts;
    if (identity) rec.bits0 |= REC_IDENTITY;
    rec.ends = pack_state(R[31].out[0]) | (pack_state(R[31].out[1]) << 16);
    return rec;
}

struct HostSink {
    uint8_t *out;
    int64_t cap;
    uint64_t pos;      // bit position
    bool overflow;
    void put(uint64_t value, unsigned len)
    {
        for (int i = (int)len - 1; i >= 0; --i) {
            if ((int64_t)(pos >> 3) < cap) { if ((value >> i) & 1) out[pos >> 3] |= (uint8_t)(0x80u >> (pos & 7)); }
            else overflow = true;
            ++pos;
        }
    }
};

}  // namespace

extern "C" {

// One channel.  out must be zero-filled by the caller (cap bytes).  stats[0] chunk solves in the sweeps,
// [1] fixed-point rounds, [2] repairs in the resolve pass, [3] chunks.  sync (optional): 4 x uint32 per chunk.
// c1_kP: k_P of the second canonical start state.  sweeps: number of speculation sweeps (>= 1).
__attribute__((visibility("default"))) int64_t emu_rlgr_encode(const int32_t *q, int64_t n, int flag_signed, uint8_t *out,
                                                               int64_t cap, uint64_t *stats, uint32_t *sync, int sweeps, int c1_kP)
{
    if (n <= 0) return 0;
    Chan C;
    C.q = q; C.n = n; C.flag_signed = flag_signed;
    C.nblk = (n + BLOCK_SYMS - 1) / BLOCK_SYMS;
    C.nchunks = (C.nblk + CHUNK_LANES - 1) / CHUNK_LANES;
    C.lastnz.assign(C.nblk, -1);
    for (int64_t b = 0; b < C.nblk; ++b) {                       // K0 + max-scan
        int32_t last = b ? C.lastnz[b - 1] : -1;
        for (int64_t i = b * BLOCK_SYMS; i < std::min<int64_t>(n, (b + 1) * BLOCK_SYMS); ++i)
            if (q[i] != 0) last = (int32_t)i;
        C.lastnz[b] = last;
    }
    std::vector<ChunkRec> prev(C.nchunks), cur(C.nchunks);
    std::vector<LaneRec> lanerec(C.nchunks * 32);
    Lane L[32];
    LaneRes R[32];
    for (int sweep = 0; sweep < sweeps; ++sweep) {                // A
        for (int64_t w = 0; w < C.nchunks; ++w) {
            State first[NCAND];
            bool pin;
            if (w == 0) {
                first[0] = first[1] = initial_state();
                pin = true;
                if (sweep > 0) { cur[w] = prev[w]; continue; }
            } else if (sweep == 0) {
                first[0] = initial_state();
                first[1].kP = c1_kP; first[1].kRP = RL_KRP_INIT;
                pin = false;
            } else {
                const uint32_t guess = prev[w - 1].ends;
                if (guess == prev[w].cand || (prev[w].bits0 & REC_IDENTITY)) { cur[w] = prev[w]; continue; }
                first[0] = unpack_state(guess & 0xffff);
                first[1] = unpack_state(guess >> 16);
                pin = true;
            }
            load_chunk(C, w, L);
            if (sweep == 0 && L[0].single) first[1] = first[0];
            bool identity = false;
            const int rounds = solve_chunk(L, pin, first, R, &identity);
            cur[w] = finish_chunk(R, identity, &lanerec[w * 32]);
            if (stats) { stats[0]++; stats[1] += rounds; }
        }
        std::swap(prev, cur);
    }
    std::vector<ChunkRec> &rec = prev;
    std::vector<uint64_t> chunkoff(C.nchunks);
    State s = initial_state();                                    // B
    uint64_t off = 0;
    for (int64_t w = 0; w < C.nchunks; ++w) {
        const uint32_t ps = pack_state(s);
        int b = ps == (rec[w].cand & 0xffff) ? 0 : ps == (rec[w].cand >> 16) ? 1 : -1;
        if (rec[w].bits0 & REC_IDENTITY) { b = 0; rec[w].bits0 = 0; rec[w].ends = ps | (ps << 16); }
        if (b < 0) {
            load_chunk(C, w, L);
            State first[NCAND] = {s, s};
            solve_chunk(L, true, first, R, nullptr);
            rec[w] = finish_chunk(R, false, &lanerec[w * 32]);
            b = 0;
            if (stats) stats[2]++;
        }
        chunkoff[w] = (off << 1) | (unsigned)b;
        off += b ? rec[w].bits1 : rec[w].bits0;
        s = unpack_state(b ? (rec[w].ends >> 16) : (rec[w].ends & 0xffff));
    }
    if (stats) stats[3] = (uint64_t)C.nchunks;
    HostSink sink{out, cap, 0, false};                            // C
    for (int64_t w = 0; w < C.nchunks; ++w) {
        load_chunk(C, w, L);
        const int b = (int)(chunkoff[w] & 1);
        uint64_t begin = chunkoff[w] >> 1;
        for (int l = 1; l < 32; ++l) {
            const LaneRec &lr = lanerec[w * 32 + l];
            State st = unpack_state(b ? (lr.starts >> 16) : (lr.starts & 0xffff));
            const uint32_t mybits = b ? (lr.bits >> 16) : (lr.bits & 0xffff);
            if (sync && (l - 1) % SYNC_STRIDE == 0) {              // sync entry of lanes l .. l+3 (rlgr.cu, emit kernel)
                uint32_t first = SYNC_NONE;
                for (int o = 0; o < SYNC_STRIDE && l + o < 32 && first == SYNC_NONE; ++o)
                    for (int i = 0; i < L[l + o].len; ++i)
                        if (L[l + o].row[i]) { first = (uint32_t)(o * BLOCK_SYMS + i); break; }
                const int k = (l - 1) / SYNC_STRIDE;
                sync[2 * (w * SYNC_PER_CHUNK + k) + 0] = (uint32_t)begin;
                sync[2 * (w * SYNC_PER_CHUNK + k) + 1] = (uint32_t)((begin >> 32) & 0x7f) | ((uint32_t)(st.kP & 0x7f) << 7) |
                                                         ((uint32_t)(st.kRP & 0xff) << 14) | (first << 22);
            }
            sink.pos = begin;
            if (mybits) code_block(L[l].row.data(), L[l].len, L[l].mask, L[l].zeros_before, L[l].is_tail, st, sink);
            if (sink.pos != begin + mybits) return -1000 - l;     // length bookkeeping broken
            begin += mybits;
        }
    }
    const int64_t nb = (int64_t)((off + 7) >> 3);
    return sink.overflow ? -nb : nb;
}

// Decoder: out must be zero-filled (n int32).  sync == nullptr: sequential decode from the stream start;
// else every sync entry is decoded on its own, as the threads of rlgr_decode_indexed_kernel do.
__attribute__((visibility("default"))) void emu_rlgr_decode(const uint8_t *buf, int64_t nbytes, int64_t n, int flag_signed,
                                                            const uint32_t *sync, int32_t *out)
{
    if (n <= 0) return;
    std::vector<uint8_t> padded((size_t)((nbytes + 3) / 4 * 4 + 8), 0);       // the reader fetches whole words
    std::memcpy(padded.data(), buf, (size_t)nbytes);
    if (!sync) {
        BitReader br;
        br.init(padded.data(), nbytes, 0);
        decode_range(br, initial_state(), -1, n, flag_signed, out);
        return;
    }
    const int64_t nblk = (n + BLOCK_SYMS - 1) / BLOCK_SYMS, nchunks = (nblk + CHUNK_LANES - 1) / CHUNK_LANES;
    for (int64_t e = 0; e < nchunks * SYNC_PER_CHUNK; ++e) {
        const int64_t w = e / SYNC_PER_CHUNK;
        const int k = (int)(e - w * SYNC_PER_CHUNK);
        const int64_t start = (w * CHUNK_LANES + (int64_t)k * SYNC_STRIDE) * BLOCK_SYMS;
        if (start >= n) continue;
        const int nblocks = std::min(SYNC_STRIDE, CHUNK_LANES - k * SYNC_STRIDE);
        const int64_t end = std::min<int64_t>(start + (int64_t)nblocks * BLOCK_SYMS, n);
        const uint32_t x = sync[2 * e], y = sync[2 * e + 1];
        const uint32_t first = y >> 22;
        if (first == SYNC_NONE || start + first >= end) continue;
        State s;
        s.kP = (int)((y >> 7) & 0x7f);
        s.kRP = (int)((y >> 14) & 0xff);
        BitReader br;
        br.init(padded.data(), nbytes, (unsigned long long)x | ((unsigned long long)(y & 0x7f) << 32));
        decode_range(br, s, start + first, end, flag_signed, out);
    }
}

}  // extern "C"
