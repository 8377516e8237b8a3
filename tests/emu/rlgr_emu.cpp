// rlgr_emu.cpp -- host emulation of the chunk-parallel RLGR encoder / indexed decoder of
// raht-3dgs-codec_b200/csrc/rlgr.cu.  TEST INFRASTRUCTURE: it runs the same arithmetic
// (csrc/rlgr_core.cuh, compiled for the host) through the same passes -- event compaction into the
// lane-interleaved store, speculation with the in-warp fixed point, resolve, emit with the sync index
// written on the way -- with the 32 lanes of a warp executed one after the other, so that the CPU
// test-suite can check the ALGORITHM (event form, tail rule, speculation + repair exactness, sync
// index) against the oracle without a GPU.  The CUDA kernels themselves are checked on the GPU by
// tests/test_gpu_parity.py.  Not mirrored here: the transition tables of rlgr.cu (an extra way for the resolve pass
// to settle a chunk before it falls back on the pinned repair, which IS mirrored) and the one-candidate kernel
// instantiation for dense channels (here: 8 equal candidates).
#include "../../raht-3dgs-codec_b200/csrc/rlgr_core.cuh"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace rlgr;

namespace {

struct Chan {
    int64_t n;
    int64_t W;                    // chunk slots in the store
    uint32_t events, nblocks, nchunks, tail_zeros;
    std::vector<Event> store;     // W * CHUNK_SLOTS, lane-interleaved (ev_slot)
    std::vector<uint32_t> blockpos;
};

int64_t ev_blocks_of(int64_t e) { return e > 0 ? (e + EV_BLOCK - 1) / EV_BLOCK : 1; }
int64_t chunks_of_blocks(int64_t nb) { return (nb + CHUNK_LANES - 1) / CHUNK_LANES; }

// K1: compaction (count / scan / scatter of rlgr.cu, done serially)
void compact(const int32_t *q, int64_t n, int flag_signed, Chan &C)
{
    C.n = n;
    C.W = chunks_of_blocks(ev_blocks_of(n));
    C.store.assign((size_t)C.W * CHUNK_SLOTS, Event{0xdeadbeefu, 0xdeadbeefu});
    C.blockpos.assign((size_t)ev_blocks_of(n), 0);
    uint32_t e = 0;
    int64_t last = -1;
    for (int64_t i = 0; i < n; ++i) {
        if (q[i] == 0) continue;
        const uint32_t u = flag_signed ? s2u(q[i]) : (uint32_t)q[i];
        const Event E{(uint32_t)(i - last - 1), u};
        const uint32_t b = e / EV_BLOCK;
        const int k = (int)(e % EV_BLOCK);
        const uint32_t w = b / CHUNK_LANES;
        const int l = (int)(b - w * CHUNK_LANES) + 1;
        C.store[(size_t)w * CHUNK_SLOTS + ev_slot(l, k)] = E;
        if (l == CHUNK_LANES && (int64_t)w + 1 < C.W) C.store[(size_t)(w + 1) * CHUNK_SLOTS + ev_slot(0, k)] = E;
        if (k == 0) C.blockpos[b] = (uint32_t)i;
        ++e;
        last = i;
    }
    C.events = e;
    C.nblocks = (uint32_t)ev_blocks_of(e);
    C.nchunks = (uint32_t)chunks_of_blocks(C.nblocks);
    C.tail_zeros = (uint32_t)(n - 1 - last);
}

struct Lane {
    const Event *base;       // chunk in the store
    int lane, cnt;
    bool is_tail;
    uint32_t tail_zeros;
    Event at(int i) const { return base[ev_slot(lane, i)]; }
};

void lane_block(const Chan &C, int64_t w, int l, Lane &B)
{
    B.base = C.store.data() + (size_t)w * CHUNK_SLOTS;
    B.lane = l;
    const int64_t b = w * CHUNK_LANES - 1 + l;
    int64_t cnt = (b < 0 || b >= (int64_t)C.nblocks) ? 0 : (int64_t)C.events - b * EV_BLOCK;
    cnt = std::max<int64_t>(0, std::min<int64_t>(cnt, EV_BLOCK));
    B.cnt = (int)cnt;
    B.is_tail = b == (int64_t)C.nblocks - 1;
    B.tail_zeros = C.tail_zeros;
}

void count_lane(const Lane &B, State &s, uint32_t &bits)
{
    bits = 0;
    bool quirk = false;
    for (int i = 0; i < B.cnt; ++i) {
        const Event e = B.at(i);
        quirk = count_event(s, e.gap, e.u, bits);
    }
    if (B.is_tail) {
        CountSink k{0};
        code_tail(s, B.tail_zeros, quirk, k);
        bits += k.bits;
    }
}

constexpr uint32_t STATE_INVALID = 0xffffu;
constexpr int SPEC_ROUNDS = 4, RESPEC_ROUNDS = 6, REPAIR_ROUNDS = 34;      // as in rlgr.cu

struct LaneRes {
    uint32_t in[NCAND], out[NCAND];      // packed states
    uint32_t bits[NCAND];
    int sel[NCAND];
};
struct ChunkSol {
    LaneRes R[32];
    bool complete[NCAND];
    uint32_t end[NCAND], total[NCAND];
    int rounds;
};

// the chains that start at lane 1's candidates: sel[c] of every lane, completeness, code length, end state
void walk_chains(ChunkSol &S, const bool has[32], const int src[32])
{
    LaneRes *R = S.R;
    int from[32][NCAND];
    for (int l = 0; l < 32; ++l)
        for (int c = 0; c < NCAND; ++c) {
            from[l][c] = 15;
            for (int k = NCAND - 1; k >= 0; --k) if (R[src[l]].out[c] == R[l].in[k]) from[l][c] = k;
            if (!has[l]) from[l][c] = c;
        }
    for (int l = 0; l < 32; ++l) for (int c = 0; c < NCAND; ++c) R[l].sel[c] = (l == 1) ? c : 15;
    for (int l = 2; l < 32; ++l)
        for (int c = 0; c < NCAND; ++c) {
            const int p = R[src[l]].sel[c];
            R[l].sel[c] = p == 15 ? 15 : from[l][p];
        }
    for (int c = 0; c < NCAND; ++c) {
        S.complete[c] = true;
        S.total[c] = 0;
        for (int l = 1; l < 32; ++l) {
            if (R[l].sel[c] == 15) S.complete[c] = false;
            else if (l > 0) S.total[c] += R[l].bits[R[l].sel[c]];
        }
        S.end[c] = R[31].out[R[31].sel[c] & (NCAND - 1)];
    }
}

// solve_chunk of rlgr.cu with the lanes executed one after the other (every round: all lanes recode, then all
// lanes look at their predecessor's end states of THIS round)
// pin: lane 0 forwards `pinned` (the known state at the chunk start); else lane 0 codes the block before the chunk
// as warm-up.  Every other lane starts from the canonical states canon[0], canon[1].
void solve_chunk(const Lane L[32], bool pin, const uint32_t pins[NCAND], unsigned pin_need, const State canon[2], int max_rounds,
                 bool expand, ChunkSol &S)
{
    bool has[32], d[32][NCAND];
    int src[32];
    LaneRes *R = S.R;
    uint32_t guess[NCAND], start[NCAND];
    for (int c = 0; c < NCAND; ++c) {
        guess[c] = pack_state(canon[c < NCAND / 2 ? 0 : 1]);
        start[c] = pin ? pins[c] : guess[c];
    }
    for (int l = 0; l < 32; ++l) {
        has[l] = L[l].cnt > 0 || L[l].is_tail;
        if (l == 0) has[l] = !pin && has[l];
        src[l] = 0;
        for (int k = 0; k < l; ++k) if (has[k]) src[l] = k;
        for (int c = 0; c < NCAND; ++c) {
            R[l].in[c] = l >= 2 ? guess[c] : start[c];
            R[l].out[c] = R[l].in[c];
            R[l].bits[c] = 0;
            d[l][c] = has[l];
        }
    }
    int r = 0;
    for (;;) {
        for (int l = 0; l < 32; ++l)
            for (int c = 0; c < NCAND; ++c)
                if (d[l][c]) {
                    State s = unpack_state(R[l].in[c]);
                    count_lane(L[l], s, R[l].bits[c]);
                    R[l].out[c] = pack_state(s);
                }
        ++r;
        if (r >= 2) {                                             // enough: the pinned chain (repair) / any chain (speculation) is complete
            for (int l = 1; l < 32; ++l)
                if (!has[l]) for (int c = 0; c < NCAND; ++c) R[l].in[c] = R[l].out[c] = R[src[l]].out[c];
            walk_chains(S, has, src);
            bool done = pin;
            if (pin) { for (int c = 0; c < NCAND; ++c) if ((pin_need >> c & 1) && !S.complete[c]) done = false; }
            else for (int c = 0; c < NCAND; ++c) done |= S.complete[c];
            if (done) break;
        }
        uint32_t o[32][NCAND];
        for (int l = 0; l < 32; ++l) for (int c = 0; c < NCAND; ++c) o[l][c] = R[src[l]].out[c];
        bool any = false;
        for (int l = 0; l < 32; ++l) {
            for (int c = 0; c < NCAND; ++c) d[l][c] = false;
            if (l == 0) continue;
            if (!has[l]) {
                for (int c = 0; c < NCAND; ++c) R[l].in[c] = R[l].out[c] = o[l][c];
                continue;
            }
            if (r >= max_rounds) continue;
            uint32_t nw[NCAND];
            bool replace;
            if (pin && l == 1) continue;                          // lane 1 holds the pinned state
            if (expand && r == 1) {                               // after the warm-up round: the classes around the predecessor's end states
                expand_classes(unpack_state(o[l][0]), unpack_state(o[l][NCAND / 2]), nw);
                replace = true;
            } else {
                if (expand && l == 1) continue;                   // lane 1's candidates are the chunk's: fixed after the expansion
                replace = false;
                for (int c = 0; c < NCAND; ++c) {
                    bool cov = false;
                    for (int k = 0; k < NCAND; ++k) cov |= o[l][c] == R[l].in[k];
                    replace |= !cov;
                    nw[c] = o[l][c];
                }
            }
            if (!replace) continue;
            const LaneRes old = R[l];
            for (int c = 0; c < NCAND; ++c) {
                int m = -1;
                for (int k = 0; k < NCAND; ++k) if (nw[c] == old.in[k]) { m = k; break; }
                R[l].in[c] = nw[c];
                if (m >= 0) { R[l].out[c] = old.out[m]; R[l].bits[c] = old.bits[m]; }
                else { d[l][c] = true; any = true; }
            }
        }
        if (!any) break;
    }
    S.rounds = r;
    for (int c = 0; c < NCAND; ++c) R[0].bits[c] = 0;
    walk_chains(S, has, src);
}

struct ChunkRec {
    uint32_t cand[NCAND], ends[NCAND], bits[NCAND];
};
struct LaneRec {
    uint32_t starts[NCAND], bits[NCAND];
    uint32_t sel;            // a nibble per chunk candidate
};

ChunkRec finish_chunk(const ChunkSol &S, LaneRec *lr)
{
    ChunkRec rec;
    for (int l = 0; l < 32; ++l) {
        lr[l].sel = 0;
        for (int c = 0; c < NCAND; ++c) {
            lr[l].starts[c] = S.R[l].in[c];
            lr[l].bits[c] = S.R[l].bits[c];
            lr[l].sel |= (uint32_t)(S.R[l].sel[c] & 15) << (4 * c);
        }
    }
    for (int c = 0; c < NCAND; ++c) {
        rec.cand[c] = S.complete[c] ? lr[1].starts[c] : STATE_INVALID;
        rec.ends[c] = S.end[c];
        rec.bits[c] = S.total[c];
    }
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

// One channel.  out must be zero-filled by the caller (cap bytes).  stats[0] chunk solves in the speculation
// pass, [1] fixed-point rounds, [2] repairs in the resolve pass, [3] chunks, [4] events, [5] rounds spent in
// repairs, [6] chunks solved again in the second sweep, [7] repairs where the true state was a candidate whose chain was incomplete.
// sync (optional): 2 x uint32 per 256 symbols, pre-filled with 0xff by the caller.
// mode: 0 = the kernel's rule (one canonical state when >= 50 % non-zero), 1 = always two, 2 = always one.
// c1_kP: k_P of the second canonical start state.  max_spec_rounds: round cap of the speculation pass (0: the kernel's).
__attribute__((visibility("default"))) int64_t emu_rlgr_encode(const int32_t *q, int64_t n, int flag_signed, uint8_t *out,
                                                               int64_t cap, uint64_t *stats, uint32_t *sync, int mode, int c1_kP,
                                                               int max_spec_rounds)
{
    if (n <= 0) return 0;
    Chan C;
    compact(q, n, flag_signed, C);
    std::vector<ChunkRec> rec(C.nchunks);
    std::vector<LaneRec> lanerec((size_t)C.nchunks * 32);
    Lane L[32];
    ChunkSol S;
    const bool single = mode == 2 || (mode == 0 && (int64_t)C.events * 2 >= n);
    State canon[2];
    canon[0] = initial_state();
    canon[1].kP = c1_kP; canon[1].kRP = RL_KRP_INIT;
    if (single) canon[1] = canon[0];
    for (int64_t w = 0; w < C.nchunks; ++w) {                     // A
        for (int l = 0; l < 32; ++l) lane_block(C, w, l, L[l]);
        uint32_t pins[NCAND];
        for (int c = 0; c < NCAND; ++c) pins[c] = pack_state(initial_state());
        solve_chunk(L, w == 0, pins, 1u, canon, max_spec_rounds > 0 ? max_spec_rounds : (w == 0 ? REPAIR_ROUNDS : SPEC_ROUNDS), !single, S);
        rec[w] = finish_chunk(S, &lanerec[w * 32]);
        if (stats) { stats[0]++; stats[1] += S.rounds; }
    }
    {                                                             // A, sweep 1: candidates = the predecessor's end states
        const std::vector<ChunkRec> prev = rec;
        const int nce = single ? 1 : NCAND;                       // the dense kernel carries one candidate
        for (int64_t w = 1; w < C.nchunks; ++w) {
            bool again = false;
            uint32_t fill = STATE_INVALID;
            for (int c = 0; c < nce; ++c)
                if (prev[w - 1].cand[c] != STATE_INVALID) {
                    const uint32_t e = prev[w - 1].ends[c];
                    if (fill == STATE_INVALID) fill = e;
                    bool have = false;
                    for (int k = 0; k < nce; ++k) have |= prev[w].cand[k] == e;
                    again |= !have;
                }
            if (!again) continue;
            uint32_t pins[NCAND];
            unsigned need = 0;
            for (int c = 0; c < NCAND; ++c) {
                const bool ok = c < nce && prev[w - 1].cand[c] != STATE_INVALID;
                pins[c] = ok ? prev[w - 1].ends[c] : fill;
                bool dup = false;
                for (int k = 0; k < c; ++k) dup |= pins[k] == pins[c];
                if (ok && !dup) need |= 1u << c;
            }
            for (int l = 0; l < 32; ++l) lane_block(C, w, l, L[l]);
            solve_chunk(L, true, pins, need, canon, max_spec_rounds > 0 ? max_spec_rounds : RESPEC_ROUNDS, !single, S);
            rec[w] = finish_chunk(S, &lanerec[w * 32]);
            if (stats) { stats[0]++; stats[1] += S.rounds; stats[6]++; }
        }
    }
    std::vector<uint64_t> chunkoff(C.nchunks);
    State s = initial_state();                                    // B
    uint64_t off = 0;
    for (int64_t w = 0; w < C.nchunks; ++w) {
        const uint32_t ps = pack_state(s);
        int b = -1;
        for (int c = NCAND - 1; c >= 0; --c) if (rec[w].cand[c] == ps) b = c;
        if (b < 0) {
            if (stats) for (int c = 0; c < NCAND; ++c) if (lanerec[w * 32 + 1].starts[c] == ps) { stats[7]++; break; }   // present, chain incomplete
            if (getenv("RLGR_EMU_DEBUG")) {
                fprintf(stderr, "miss chunk %ld true (%d,%d) cands:", (long)w, s.kP, s.kRP);
                for (int c = 0; c < NCAND; ++c) {
                    const State t = unpack_state(lanerec[w * 32 + 1].starts[c]);
                    fprintf(stderr, " (%d,%d)%s", t.kP, t.kRP, rec[w].cand[c] == STATE_INVALID ? "x" : "");
                }
                fprintf(stderr, "\n");
            }
            for (int l = 0; l < 32; ++l) lane_block(C, w, l, L[l]);
            uint32_t pins[NCAND];
            for (int c = 0; c < NCAND; ++c) pins[c] = ps;
            solve_chunk(L, true, pins, 1u, canon, REPAIR_ROUNDS, !single, S);
            if (stats) stats[5] += S.rounds;
            if (!S.complete[0]) return -2000;                     // the exact chain must always complete
            rec[w] = finish_chunk(S, &lanerec[w * 32]);
            b = 0;
            if (stats) stats[2]++;
        }
        chunkoff[w] = (off << 3) | (unsigned)b;
        off += rec[w].bits[b];
        s = unpack_state(rec[w].ends[b]);
    }
    if (stats) { stats[3] = C.nchunks; stats[4] = C.events; }
    HostSink sink{out, cap, 0, false};                            // C
    for (int64_t w = 0; w < C.nchunks; ++w) {
        const int b = (int)(chunkoff[w] & 7);
        uint64_t begin = chunkoff[w] >> 3;
        for (int l = 1; l < 32; ++l) {
            lane_block(C, w, l, L[l]);
            const LaneRec &lr = lanerec[w * 32 + l];
            const int sel = (int)((lr.sel >> (4 * b)) & (NCAND - 1));
            State st = unpack_state(lr.starts[sel]);
            const uint32_t mybits = lr.bits[sel];
            if (mybits == 0) continue;
            sink.pos = begin;
            int64_t pos = -1;
            bool quirk = false;
            for (int i = 0; i < L[l].cnt; ++i) {
                const Event e = L[l].at(i);
                if (i == 0) pos = (int64_t)C.blockpos[w * CHUNK_LANES - 1 + l] - (int64_t)e.gap - 1;
                const int64_t np = pos + (int64_t)e.gap + 1;
                if (sync && (np >> SYNC_SHIFT) != (pos >> SYNC_SHIFT)) {     // first event of its 256-symbol range
                    const uint64_t at = sink.pos;
                    sync[2 * (np >> SYNC_SHIFT) + 0] = (uint32_t)at;
                    sync[2 * (np >> SYNC_SHIFT) + 1] = (uint32_t)((at >> 32) & 0x7f) | ((uint32_t)(st.kP & 0x7f) << 7) |
                                                       ((uint32_t)(st.kRP & 0xff) << 14) | ((uint32_t)(np & (SYNC_RANGE - 1)) << 22);
                }
                pos = np;
                quirk = emit_event(st, e.gap, e.u, sink);
            }
            if (L[l].is_tail) code_tail(st, L[l].tail_zeros, quirk, sink);
            if (sink.pos != begin + mybits) return -1000 - l;     // length bookkeeping broken
            begin += mybits;
        }
    }
    const int64_t nb = (int64_t)((off + 7) >> 3);
    return sink.overflow ? -nb : nb;
}

// Dense channels (rlgr.cu: channels with at least half of their symbols non-zero) are coded in SYMBOL space: lane
// blocks of 128 symbols walked by the per-symbol machine of rlgr_core.cuh (sym_begin / sym_count / sym_emit), a block's
// start state first guessed from a warm-up over the last `warm` symbols of the block before it.  This entry point runs
// that path block by block with the state carried exactly (the fixed-point rounds that repair wrong guesses are the
// same as in emu_rlgr_encode and are not repeated here): it checks the machine, the block-boundary convention
// (pending zeros are coded by the block that holds the non-zero symbol that ends them; a block stops at its last
// non-zero symbol), the tail rule, the length bookkeeping (sym_count == bits written by sym_emit) and the sync index.
// stats[0] = blocks, stats[1] = blocks whose warm-up guess equals the true start state.
__attribute__((visibility("default"))) int64_t emu_rlgr_encode_dense(const int32_t *q, int64_t n, int flag_signed, uint8_t *out,
                                                                     int64_t cap, uint32_t *sync, int warm, uint64_t *stats)
{
    if (n <= 0) return 0;
    const int64_t nsb = (n + SYM_BLOCK - 1) / SYM_BLOCK;
    int64_t lastnz = -1;
    for (int64_t i = n - 1; i >= 0; --i) if (q[i] != 0) { lastnz = i; break; }
    const int64_t tail_block = lastnz < 0 ? 0 : lastnz / SYM_BLOCK;
    const uint32_t tail_zeros = (uint32_t)(n - 1 - lastnz);
    HostSink sink{out, cap, 0, false};
    State s = initial_state();
    int64_t prev_nz = -1;                                         // last non-zero symbol before the block
    auto mapped = [&](int64_t i) { return q[i] == 0 ? 0u : (flag_signed ? s2u(q[i]) : (uint32_t)q[i]); };
    for (int64_t b = 0; b <= tail_block && b < nsb; ++b) {
        const int64_t pos0 = b * SYM_BLOCK;
        int64_t last = -1;                                        // last non-zero symbol of the block
        for (int64_t i = std::min<int64_t>(pos0 + SYM_BLOCK, n) - 1; i >= pos0; --i) if (q[i] != 0) { last = i; break; }
        const int cnt = last < 0 ? 0 : (int)(last - pos0 + 1);
        const uint32_t gap0 = (uint32_t)(pos0 - 1 - prev_nz);
        const bool is_tail = b == tail_block;
        if (stats) {                                              // would the warm-up over the previous block have guessed s?
            stats[0]++;
            if (b > 0) {
                int64_t plast = -1;
                for (int64_t i = pos0 - 1; i >= pos0 - SYM_BLOCK; --i) if (q[i] != 0) { plast = i; break; }
                const int pcnt = plast < 0 ? 0 : (int)(plast - (pos0 - SYM_BLOCK) + 1);
                int from = (pcnt - warm) & ~7;
                if (from < 0) from = 0;
                SymState g{RL_KP_INIT, RL_KRP_INIT, 0};
                uint32_t dummy = 0;
                for (int i = from; i < pcnt; ++i) sym_count(g, mapped(pos0 - SYM_BLOCK + i), dummy);
                if (pcnt > 0 ? (g.kP == s.kP && g.kRP == s.kRP) : false) stats[1]++;
            } else stats[1]++;
        }
        // lengths
        SymState t{s.kP, s.kRP, 0};
        uint32_t bits = 0;
        bool quirk = false;
        if (cnt > 0 || is_tail) {
            CountSink cs{0};
            if (cnt > 0) sym_begin(t, gap0, cs);
            bits = cs.bits;
            for (int i = 0; i < cnt; ++i) sym_count(t, mapped(pos0 + i), bits);
            if (cnt > 0 && t.m != 0) return -3000;                // a block ends at an event boundary
            quirk = cnt > 0 && sym_tail_quirk(t, mapped(pos0 + cnt - 1));
            if (is_tail) {
                State st{t.kP, t.kRP};
                CountSink k{0};
                code_tail(st, tail_zeros, quirk, k);
                bits += k.bits;
            }
        }
        // bits
        const uint64_t begin = sink.pos;
        SymState e{s.kP, s.kRP, 0};
        bool q2 = false;
        // sync index: a block's first event is the first of its 256-symbol range iff the zeros pending in front of the
        // block reach back to the start of the range (ranges start at block boundaries); the coder stands before that
        // event in the block's start state at the block's first bit
        bool want_sync = sync && cnt > 0 && gap0 >= (uint32_t)(pos0 & (SYNC_RANGE - 1));
        if (cnt > 0) sym_begin(e, gap0, sink);
        for (int i = 0; i < cnt; ++i) {
            const int64_t np = pos0 + i;
            if (want_sync && q[np] != 0) {
                sync[2 * (np >> SYNC_SHIFT) + 0] = (uint32_t)begin;
                sync[2 * (np >> SYNC_SHIFT) + 1] = (uint32_t)((begin >> 32) & 0x7f) | ((uint32_t)(s.kP & 0x7f) << 7) |
                                                   ((uint32_t)(s.kRP & 0xff) << 14) | ((uint32_t)(np & (SYNC_RANGE - 1)) << 22);
                want_sync = false;
            }
            sym_emit(e, mapped(np), sink);
        }
        q2 = cnt > 0 && sym_tail_quirk(e, mapped(pos0 + cnt - 1));
        if (cnt > 0 && (e.kP != t.kP || e.kRP != t.kRP || e.m != 0 || q2 != quirk)) return -3100;
        if (is_tail) {
            State st{e.kP, e.kRP};
            code_tail(st, tail_zeros, q2, sink);
        }
        if (sink.pos != begin + bits) return -3200;               // sym_count and sym_emit disagree on the length
        if (cnt > 0) { s.kP = t.kP; s.kRP = t.kRP; prev_nz = last; }
    }
    const int64_t nb = (int64_t)((sink.pos + 7) >> 3);
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
    const int64_t entries = (n + SYNC_RANGE - 1) >> SYNC_SHIFT;
    for (int64_t e = 0; e < entries; ++e) {
        const int64_t start = e << SYNC_SHIFT;
        const int64_t end = std::min<int64_t>(start + SYNC_RANGE, n);
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
