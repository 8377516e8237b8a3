// rlgr_core.cuh -- the arithmetic of the adaptive Run-Length / Golomb-Rice coder, written once for
// the device kernels (rlgr.cu) and for the host-side warp emulation used by the CPU tests
// (tests/emu/rlgr_emu.cpp).  Restates python/PyRLGR/src/libs/rlgr/membuf.cpp:340-423 (rlgrWrite),
// :242-256 (grWrite), :258-338 (rlgrRead) with constants membuf.h:18-22, in EVENT form:
//
//   an event = g zero symbols followed by one non-zero symbol u.
//
// Between events the coder's run counter m is 0 (membuf.cpp:379, :412), so the adaptive state at an
// event boundary is just (k_P, k_RP).  The coder output of an event and the state after it are a
// pure function of (state before, g, u) -- which is what lets the encoder cut a channel into blocks
// at event boundaries and run the blocks in parallel from speculated states (see rlgr.cu).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define RL_HD __host__ __device__ __forceinline__
#else
#define RL_HD inline
#endif

namespace rlgr {

constexpr int RL_L = 4, RL_U0 = 3, RL_D0 = 1, RL_U1 = 2, RL_D1 = 1;     // membuf.h:18-22
constexpr int RL_KRP_MAX = 32 * RL_L;
constexpr int RL_KP_INIT = 0, RL_KRP_INIT = 2 * RL_L;                   // membuf.cpp:343-345

constexpr int BLOCK_SYMS = 128;                    // symbols per lane block
constexpr int MASK_WORDS = BLOCK_SYMS / 32;
constexpr int CHUNK_LANES = 31;                    // lane blocks per warp chunk (lane 0 = warm-up block)
constexpr int CHUNK_SYMS = BLOCK_SYMS * CHUNK_LANES;
constexpr int NCAND = 2;                           // speculated start states carried side by side

struct State {
    int kP, kRP;
};
RL_HD bool operator==(const State &a, const State &b) { return a.kP == b.kP && a.kRP == b.kRP; }
RL_HD bool operator!=(const State &a, const State &b) { return !(a == b); }
RL_HD uint32_t pack_state(State s) { return (uint32_t)(s.kP & 0xff) | ((uint32_t)(s.kRP & 0xff) << 8); }
RL_HD State unpack_state(uint32_t v) { State s; s.kP = (int)(v & 0xff); s.kRP = (int)((v >> 8) & 0xff); return s; }
RL_HD State initial_state() { State s; s.kP = RL_KP_INIT; s.kRP = RL_KRP_INIT; return s; }

RL_HD uint32_t s2u(int32_t v)            // membuf.cpp:4-14 (_s2u); |v| < 2^31
{
    return v < 0 ? (((uint32_t)(-(int64_t)v)) << 1) - 1u : ((uint32_t)v) << 1;
}
RL_HD int32_t u2s(uint64_t u)
{
    int64_t d = (int64_t)(u >> 1);
    return (int32_t)((u & 1) ? -d - 1 : d);
}

RL_HD int rl_ctz(uint32_t x)
{
#if defined(__CUDA_ARCH__)
    return __ffs((int)x) - 1;
#else
    return __builtin_ctz(x);
#endif
}

RL_HD uint32_t rl_bswap32(uint32_t x)
{
#if defined(__CUDA_ARCH__)
    return __byte_perm(x, 0, 0x0123);
#else
    return __builtin_bswap32(x);
#endif
}
RL_HD int rl_clz64(uint64_t x)           // 64 for x == 0
{
#if defined(__CUDA_ARCH__)
    return __clzll((long long)x);
#else
    return x ? __builtin_clzll(x) : 64;
#endif
}

// Sink that only counts bits (passes that need code lengths, not codes)
struct CountSink {
    uint32_t bits;
    RL_HD void put(uint64_t, unsigned len) { bits += len; }
};

// k_RP adaptation after coding value u with parameter k_R (membuf.cpp:367-376, :395-404)
RL_HD void adapt_krp(State &s, uint64_t p)
{
    if (p) {
        uint64_t t = (uint64_t)s.kRP + p - 1;
        s.kRP = t > (uint64_t)RL_KRP_MAX ? RL_KRP_MAX : (int)t;
    } else {
        s.kRP = s.kRP < 2 ? 0 : s.kRP - 2;
    }
}

// grWrite(u, kr), membuf.cpp:242-256.  Returns p = u >> kr.
template <class Sink> RL_HD uint64_t gr_write(Sink &sink, uint64_t u, unsigned kr)
{
    const uint64_t p = u >> kr;
    if (p < 32) {
        sink.put((((uint64_t)1) << (p + 1)) - 2, (unsigned)p + 1);     // p ones and a zero
        sink.put(u & ((((uint64_t)1) << kr) - 1), kr);
    } else {
        sink.put(0xffffffffull, 32);                                   // escape: 32 ones, then u in 32 bits
        sink.put(u & 0xffffffffull, 32);
    }
    return p;
}

// g zero symbols starting from state s with an empty run counter.  Returns the partial run m left
// open (0 <= m < 2^k).  *closed_norun: the last zero was coded in no-run mode (k == 0), i.e. the
// tail rule of membuf.cpp:416 does not fire after it.
template <class Sink> RL_HD uint32_t code_zeros(State &s, uint32_t g, Sink &sink, bool *closed_norun)
{
    bool norun = false;
    while (g) {
        const int k = s.kP / RL_L;
        if (k == 0) {                                // membuf.cpp:391-413 with u == 0
            sink.put(0, 1);
            sink.put(0, (unsigned)(s.kRP / RL_L));
            s.kRP = s.kRP < 2 ? 0 : s.kRP - 2;
            s.kP += RL_U0;
            --g;
            norun = true;
        } else {                                     // :382-390
            const uint32_t run = 1u << k;
            norun = false;
            if (g < run) break;
            g -= run;
            sink.put(1, 1);
            s.kP += RL_U1;
        }
    }
    if (closed_norun) *closed_norun = norun && g == 0;
    return g;
}

// the non-zero symbol u (already mapped by _s2u) that ends an event; m = open partial run.
// Returns true when it was coded in run mode with u - 1 == 0 (tail quirk, membuf.cpp:416-419).
template <class Sink> RL_HD bool code_nonzero(State &s, uint32_t m, uint64_t u, Sink &sink)
{
    const int k = s.kP / RL_L;
    const unsigned kR = (unsigned)(s.kRP / RL_L);
    bool quirk = false;
    if (k) {                                         // :353-381
        u -= 1;
        sink.put(0, 1);
        sink.put(m, (unsigned)k);
        quirk = (u == 0);
        adapt_krp(s, gr_write(sink, u, kR));
        s.kP = s.kP < RL_D1 ? 0 : s.kP - RL_D1;
    } else {                                         // :391-413 with u != 0
        adapt_krp(s, gr_write(sink, u, kR));
        s.kP = s.kP < RL_D0 ? 0 : s.kP - RL_D0;
    }
    return quirk;
}

// What follows the last event of a channel: g trailing zeros (possibly none) and the tail rule
// `if (k && !u) { write(0); write(m, k_P / L); }` (membuf.cpp:416-419).  last_quirk: result of
// code_nonzero for the channel's last symbol when g == 0 (false when the channel has no symbols).
template <class Sink> RL_HD void code_tail(State &s, uint32_t g, bool last_quirk, Sink &sink)
{
    if (g) {
        bool closed_norun = false;
        const uint32_t m = code_zeros(s, g, sink, &closed_norun);
        if (!closed_norun) {
            sink.put(0, 1);
            sink.put(m, (unsigned)(s.kP / RL_L));
        }
    } else if (last_quirk) {
        sink.put(0, 1);
        sink.put(0, (unsigned)(s.kP / RL_L));
    }
}

// One lane block: the events whose non-zero symbol lies in row[0..len), in order.  row holds the
// mapped symbols u; mask bit i of word w is set iff row[32 w + i] != 0.  zeros_before = number of
// zero symbols between the previous non-zero of the channel (or its start) and row[0].
// is_tail: this block holds the channel's last symbol, so the trailing zeros and the tail rule are
// coded here too.
template <class Sink>
RL_HD void code_block(const uint32_t *row, int len, const uint32_t *mask, uint32_t zeros_before, bool is_tail, State &s,
                      Sink &sink)
{
    int prev = -1;
    bool quirk = false;
    for (int w = 0; w < MASK_WORDS; ++w) {
        uint32_t mm = mask[w];
        while (mm) {
            const int idx = 32 * w + rl_ctz(mm);
            mm &= mm - 1;
            uint32_t g = (uint32_t)(idx - prev - 1) + (prev < 0 ? zeros_before : 0u);
            uint32_t m = 0;
            if (g) m = code_zeros(s, g, sink, (bool *)nullptr);
            quirk = code_nonzero(s, m, (uint64_t)row[idx], sink);
            prev = idx;
        }
    }
    if (is_tail) {
        const uint32_t g = prev < 0 ? zeros_before + (uint32_t)len : (uint32_t)(len - 1 - prev);
        code_tail(s, g, quirk, sink);
    }
}

// ---- dense blocks: one branch-free step per SYMBOL (lengths only) -----------------------------------------------
// Same automaton as the event form, advanced one symbol at a time with selects instead of branches, for NC
// candidate states side by side (independent chains: instruction-level parallelism instead of divergence).
// Zeros cost as much as non-zeros here, so callers use it for blocks that are mostly non-zero.
struct SymState {
    int kP, kRP;
    uint32_t m;         // open partial run
    uint32_t bits;
};

// one symbol x (already mapped; 0 = zero symbol).  Returns true when the tail rule would fire if this were the
// channel's last symbol (membuf.cpp:416: k at the top of the iteration != 0 and u, post-decrement, == 0).
RL_HD bool sym_step(SymState &t, uint32_t x)
{
    const int k = t.kP >> 2;
    const unsigned kR = (unsigned)(t.kRP >> 2);
    const bool run = k != 0, z = x == 0;
    const bool coded = !(run && z);                               // a Golomb-Rice code is written for this symbol
    const uint32_t v = x - ((run && !z) ? 1u : 0u);               // run mode codes u - 1
    const uint32_t p = kR > 31 ? 0u : v >> kR;
    const uint32_t grlen = p < 32 ? p + 1 + kR : 64u;
    const uint32_t m1 = t.m + 1;
    const bool complete = !coded && m1 == (1u << k);
    t.bits += coded ? grlen + (run ? (uint32_t)(1 + k) : 0u) : (complete ? 1u : 0u);
    const int krp_up = t.kRP + (int)(p > 256u ? 256u : p) - 1;
    const int krp_new = p ? (krp_up > RL_KRP_MAX ? RL_KRP_MAX : krp_up) : (t.kRP < 2 ? 0 : t.kRP - 2);
    t.kRP = coded ? krp_new : t.kRP;
    const int kp_coded = z ? t.kP + RL_U0 : (t.kP < 1 ? 0 : t.kP - 1);   // D0 == D1 == 1
    t.kP = coded ? kp_coded : (complete ? t.kP + RL_U1 : t.kP);
    t.m = (coded || complete) ? 0u : m1;
    return run && v == 0;
}

// The block of code_block(), lengths only, for NC candidate states.  last_nz = index of the block's last non-zero
// symbol (-1: none).  s[] in/out, bits[] out.
template <int NC>
RL_HD void count_block_dense(const uint32_t *row, int len, int last_nz, uint32_t zeros_before, bool is_tail, State *s,
                             uint32_t *bits)
{
    SymState t[NC];
    for (int c = 0; c < NC; ++c) {
        CountSink k{0};
        State st = s[c];
        uint32_t m = 0;
        if (zeros_before && (last_nz >= 0 || is_tail)) m = code_zeros(st, zeros_before, k, (bool *)nullptr);
        t[c].kP = st.kP; t[c].kRP = st.kRP; t[c].m = m; t[c].bits = k.bits;
    }
    const int end = is_tail ? len : last_nz + 1;
    bool fire[NC];
    for (int c = 0; c < NC; ++c) fire[c] = false;
    for (int i = 0; i < end; ++i) {
        const uint32_t x = row[i];
        for (int c = 0; c < NC; ++c) fire[c] = sym_step(t[c], x);
    }
    for (int c = 0; c < NC; ++c) {
        if (is_tail && end > 0 && fire[c]) t[c].bits += 1 + (uint32_t)(t[c].kP >> 2);     // write(0); write(m, k_P / L)
        s[c].kP = t[c].kP; s[c].kRP = t[c].kRP;
        bits[c] = t[c].bits;
    }
}

// ---- decoder ---------------------------------------------------------------------------------------------
// MSB-first reader over big-endian 32-bit words; acc holds the next nbits bits left-aligned.
struct BitReader {
    const uint32_t *in;
    int64_t n_words, pos;
    uint64_t acc;
    int nbits;
    RL_HD void init(const uint8_t *stream, int64_t readable_bytes, unsigned long long begin_bit)
    {
        in = (const uint32_t *)stream;
        n_words = (readable_bytes + 3) >> 2;
        pos = (int64_t)(begin_bit >> 5);
        acc = 0;
        nbits = 0;
        const unsigned skip = (unsigned)(begin_bit & 31);
        if (skip) { fill(); acc <<= skip; nbits -= (int)skip; }
    }
    RL_HD void fill()                  // afterwards nbits >= 33
    {
        if (nbits <= 32) {
            const uint32_t w = pos < n_words ? rl_bswap32(in[pos]) : 0u;
            ++pos;
            acc |= (uint64_t)w << (32 - nbits);
            nbits += 32;
        }
    }
    RL_HD unsigned get1()
    {
        fill();
        const unsigned b = (unsigned)(acc >> 63);
        acc <<= 1;
        --nbits;
        return b;
    }
    RL_HD uint64_t get(unsigned len)   // len <= 32
    {
        if (len == 0) return 0;
        fill();
        const uint64_t v = acc >> (64 - len);
        acc <<= len;
        nbits -= (int)len;
        return v;
    }
    RL_HD uint64_t gr(unsigned kr)     // membuf.cpp:228-240 (grRead)
    {
        fill();
        const int p = rl_clz64(~acc);             // leading ones (bits past nbits read as zero)
        if (p >= 32) {
            acc <<= 32;
            nbits -= 32;
            return get(32);
        }
        acc <<= (p + 1);
        nbits -= p + 1;
        return (((uint64_t)p) << kr) + get(kr);
    }
};

// Events from reader br in state s.  dst is pre-zeroed: only non-zero symbols are stored, so zero runs cost
// nothing and the tail of the stream is never read.
//   first_pos < 0: plain sequential decode of symbols [0, end) (reader at the stream start);
//   first_pos >= 0: the reader stands before an event whose non-zero symbol is dst[first_pos] (a sync entry);
//   the zeros in front of it are decoded and dropped, then decoding continues up to `end`.
RL_HD void decode_range(BitReader &br, State s, int64_t first_pos, int64_t end, int flag_signed,
                                             int32_t *__restrict__ dst)
{
    bool synced = first_pos < 0;
    int64_t pos = 0;
    while (pos < end) {                                                    // membuf.cpp:266-334
        int k = s.kP / RL_L;
        const unsigned kR = (unsigned)(s.kRP / RL_L);
        if (k) {
            uint64_t m = 0;
            while (br.get1()) {                                            // complete runs
                m += ((uint64_t)1) << k;
                s.kP += RL_U1;
                k = s.kP / RL_L;
                if (m > (uint64_t)1 << 40) break;                          // corrupt stream guard
            }
            m += br.get((unsigned)k);
            if (synced) {
                if (m >= (uint64_t)(end - pos)) break;                     // the next non-zero lies past this range
                pos += (int64_t)m;
            } else {
                pos = first_pos;
                synced = true;
            }
            const uint64_t u = br.gr(kR);
            const uint64_t val = u + 1;
            dst[pos++] = flag_signed ? u2s(val) : (int32_t)val;
            adapt_krp(s, u >> kR);
            s.kP = s.kP < RL_D1 ? 0 : s.kP - RL_D1;
        } else {
            const uint64_t u = br.gr(kR);
            if (u) {
                if (!synced) { pos = first_pos; synced = true; }
                dst[pos] = flag_signed ? u2s(u) : (int32_t)u;
            }
            if (synced) ++pos;
            adapt_krp(s, u >> kR);
            if (u) s.kP = s.kP < RL_D0 ? 0 : s.kP - RL_D0;
            else s.kP += RL_U0;
        }
    }
}

}  // namespace rlgr
