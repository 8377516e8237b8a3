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

constexpr int EV_BLOCK = 128;                      // events per lane block
constexpr int EV_GROUP = 4;                        // events per 32-byte unit of the lane-interleaved event store
constexpr int CHUNK_LANES = 31;                    // lane blocks per warp chunk (lane 0 = warm-up block)
constexpr int CHUNK_EVENTS = EV_BLOCK * CHUNK_LANES;
constexpr int CHUNK_SLOTS = 32 * EV_BLOCK;         // event slots of a chunk in the store (32 lanes)
constexpr int NCAND = 8;                           // speculated start states of a block carried side by side
constexpr int SYM_BLOCK = 128;                     // symbols per block of the compaction passes
constexpr int SYNC_SHIFT = 8;                      // one sync entry per 2^8 = 256 symbols
constexpr int SYNC_RANGE = 1 << SYNC_SHIFT;
constexpr uint32_t SYNC_NONE = 1023;               // no non-zero symbol in the entry's range

// slot of event i (0..127) of lane l in a chunk of the store: units of 4 events, lane-interleaved, so that a warp
// reading unit j of all its lanes touches 1 KB of consecutive memory
RL_HD uint32_t ev_slot(int lane, int i) { return (uint32_t)((((i >> 2) * 32 + lane) << 2) | (i & 3)); }

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

// The state classes that do not re-synchronise: while a channel stays in run mode with k_R = 0 (sparse, small
// values), two trajectories keep the parity of their k_P difference (every event takes 1 from both, every completed
// run adds 2) and, unless a value of -1 (p = 0) comes by, their k_RP difference (k_RP moves in steps of 2 there
// except for p = 2).  Inside a class the trajectories contract (the lower k_P completes more runs) and merge within
// a few events.  The 8 candidates carried per block are therefore k_P in a-1 .. a+2 x k_RP in {a, a ^ 2} around a
// state a when it is in that regime, else a and b four times each.
RL_HD void expand_classes(State a, State b, uint32_t out[8])
{
    for (int c = 0; c < 8; ++c) {
        State s;
        if (a.kRP < 4) {
            s.kP = a.kP - 1 + (c & 3);
            if (s.kP < 0) s.kP = 0;
            s.kRP = (c >> 2) ? (a.kRP ^ 2) : a.kRP;
        } else s = c < 4 ? a : b;
        out[c] = pack_state(s);
    }
}

// ---- event space ------------------------------------------------------------------------------------------
// The encoder works on the compacted EVENT list of a channel (rlgr.cu): event e = (gap, u) = the number of zero
// symbols since the previous non-zero symbol (or the channel start) and the mapped non-zero symbol.  A lane
// block is EV_BLOCK consecutive events; work, warm-up length and chunk count are proportional to the number of
// events, whatever the zero runs between them.
struct Event {
    uint32_t gap, u;
};

// Code length of one event from state s (lengths only, lean form of code_zeros + code_nonzero).
// Returns the tail quirk of code_nonzero.
RL_HD bool count_event(State &s, uint32_t g, uint32_t u, uint32_t &bits)
{
    if (g) {
        CountSink z{0};
        code_zeros(s, g, z, (bool *)nullptr);
        bits += z.bits;
    }
    const int k = s.kP >> 2;
    const unsigned kR = (unsigned)(s.kRP >> 2);
    const uint32_t v = u - (k ? 1u : 0u);                               // run mode codes u - 1 (membuf.cpp:355)
    const uint32_t p = kR > 31 ? 0u : (v >> kR);
    bits += (p < 32 ? p + 1 + kR : 64u) + (k ? 1u + (uint32_t)k : 0u);  // grWrite (:242-256) + `0`, m in k bits (:357-358)
    const int up = s.kRP + (int)(p > 256u ? 256u : p) - 1;
    s.kRP = p ? (up > RL_KRP_MAX ? RL_KRP_MAX : up) : (s.kRP < 2 ? 0 : s.kRP - 2);
    s.kP = s.kP < 1 ? 0 : s.kP - 1;                                     // D0 == D1 == 1
    return k != 0 && v == 0;
}

// The same event with its bits written to sink (put(value, len <= 32)).
template <class Sink> RL_HD bool emit_event(State &s, uint32_t g, uint32_t u, Sink &sink)
{
    uint32_t m = 0;
    if (g) m = code_zeros(s, g, sink, (bool *)nullptr);
    const int k = s.kP >> 2;
    const unsigned kR = (unsigned)(s.kRP >> 2);
    uint32_t v = u;
    if (k) {
        v = u - 1;
        sink.put(m, 1u + (unsigned)k);                                  // `0` then m in k bits
    }
    const uint32_t p = kR > 31 ? 0u : (v >> kR);
    if (p + 1 + kR <= 32) {                                             // p ones, a zero, kR low bits: one put
        const uint32_t ones = ((1u << p) - 1u) << 1;                    // p <= 31 here
        const uint32_t low = kR ? (v & (0xffffffffu >> (32 - kR))) : 0u;
        sink.put(((uint64_t)ones << kR) | low, p + 1 + kR);
    } else if (p < 32) {
        sink.put((((uint64_t)1) << (p + 1)) - 2, p + 1);
        sink.put(kR > 31 ? (uint64_t)v : (uint64_t)(v & ((1u << kR) - 1u)), kR);
    } else {
        sink.put(0xffffffffull, 32);                                    // escape: 32 ones, then v in 32 bits
        sink.put((uint64_t)v, 32);
    }
    const int up = s.kRP + (int)(p > 256u ? 256u : p) - 1;
    s.kRP = p ? (up > RL_KRP_MAX ? RL_KRP_MAX : up) : (s.kRP < 2 ? 0 : s.kRP - 2);
    s.kP = s.kP < 1 ? 0 : s.kP - 1;
    return k != 0 && v == 0;
}

// ---- dense channels: the coder as a per-SYMBOL machine -----------------------------------------------------
// membuf.cpp:340-423 symbol by symbol, branch-free where the lanes of a warp would diverge: the state is
// (k_P, k_RP, m) with m = zeros of the open partial run.  A lane block of a dense channel is a range of SYMBOLS that
// ends with a non-zero symbol (an event boundary: m = 0 there, so blocks still hand on (k_P, k_RP) only); the zeros
// pending in front of the block are coded first (sym_begin), which leaves the run counter the block starts with.
// Every lane of a warp does the same work per symbol, whatever the symbol is -- the event form above serialises the
// lanes that sit in a zero run, in run mode or in an escape against those that do not.
struct SymState {
    int kP, kRP;
    uint32_t m;
};

template <class Sink> RL_HD void sym_begin(SymState &t, uint32_t gap0, Sink &sink)
{
    State s;
    s.kP = t.kP; s.kRP = t.kRP;
    t.m = gap0 ? code_zeros(s, gap0, sink, (bool *)nullptr) : 0u;
    t.kP = s.kP; t.kRP = s.kRP;
}

RL_HD uint32_t rl_shr_clamp(uint32_t v, uint32_t sh)                // v >> sh with sh in 0 .. 32 (32 -> 0)
{
#if defined(__CUDA_ARCH__)
    return __funnelshift_rc(v, 0u, sh);
#else
    return sh > 31 ? 0u : (v >> sh);
#endif
}
RL_HD int rl_max0(int x) { return x < 0 ? 0 : x; }

// One symbol, code length only.  u = the mapped symbol (_s2u; 0 for a zero).  Written as straight-line selects: every
// lane of a warp executes the same ~35 instructions whatever its symbol and mode are.
RL_HD void sym_count(SymState &t, uint32_t u, uint32_t &bits)
{
    const uint32_t k = (uint32_t)t.kP >> 2;
    const uint32_t kR = (uint32_t)t.kRP >> 2;                          // 0 .. 32
    const bool nz = u != 0, run = t.kP >= RL_L;
    const bool coded = nz || !run;                                      // a Golomb-Rice code is written (:359, :393)
    const uint32_t v = u - (run ? 1u : 0u);                             // run mode codes u - 1 (:355)
    const uint32_t p = rl_shr_clamp(v, kR);
    const uint32_t gr = p < 32 ? p + 1 + kR : 64u;                      // grWrite (:242-256)
    const uint32_t m1 = t.m + 1;
    const bool full = run && !nz && m1 == (1u << (k & 31));             // a complete run: one 1 bit (:383-389)
    bits += (coded ? gr + (run ? 1u + k : 0u) : 0u) + (full ? 1u : 0u);
    const int pc = (int)(p > 256u ? 256u : p);
    const int up = t.kRP + pc - 1;
    const int adapted = p ? (up > RL_KRP_MAX ? RL_KRP_MAX : up) : rl_max0(t.kRP - 2);
    t.kRP = coded ? adapted : t.kRP;
    const int grow = run ? (full ? t.kP + RL_U1 : t.kP) : t.kP + RL_U0;
    t.kP = nz ? rl_max0(t.kP - 1) : grow;                               // D0 == D1 == 1
    t.m = (run && !nz && !full) ? m1 : 0u;
}

// The tail quirk (membuf.cpp:416-419 looks at the k and the decremented u of the LAST symbol) for a block that ends
// with the channel's last non-zero symbol u_last, from the state AFTER it: the symbol was coded in run mode
// (k_P >= 4 before it, i.e. k_P >= 3 after its decrement) with u - 1 == 0.
RL_HD bool sym_tail_quirk(const SymState &after, uint32_t u_last) { return u_last == 1u && after.kP >= RL_L - 1; }

// The same symbol with its bits written to sink (put(value, len <= 32)).  The common case -- run prefix, unary part
// and remainder together fit 32 bits -- is one put.
template <class Sink> RL_HD void sym_emit(SymState &t, uint32_t u, Sink &sink)
{
    const uint32_t k = (uint32_t)t.kP >> 2;
    const uint32_t kR = (uint32_t)t.kRP >> 2;
    const bool run = t.kP >= RL_L;
    if (run && u == 0) {                                                // a zero in run mode (:382-390)
        const uint32_t m1 = t.m + 1;
        if (m1 == (1u << (k & 31))) {
            sink.put(1, 1);
            t.kP += RL_U1;
            t.m = 0;
        } else t.m = m1;
        return;
    }
    const uint32_t v = u - (run ? 1u : 0u);
    const uint32_t p = rl_shr_clamp(v, kR);
    const uint32_t plen = run ? 1u + k : 0u;                            // `0` then m in k bits (:357-358)
    const uint32_t glen = p + 1 + kR;                                   // p ones, a zero, kR low bits (:242-256)
    if (p < 32 && glen + plen <= 32) {                                  // (kR <= 31 here)
        const uint32_t ones = ((1u << p) - 1u) << 1;
        const uint32_t low = v & ~(0xffffffffu << kR);
        sink.put((uint64_t)((((run ? t.m : 0u) << (p + 1)) | ones) << kR | low), glen + plen);
    } else {
        if (run) sink.put(t.m, plen);
        if (p < 32) {
            sink.put((((uint64_t)1) << (p + 1)) - 2, p + 1);
            sink.put(kR > 31 ? (uint64_t)v : (uint64_t)(v & ((1u << kR) - 1u)), kR);
        } else {
            sink.put(0xffffffffull, 32);                                // escape: 32 ones, then v in 32 bits
            sink.put((uint64_t)v, 32);
        }
    }
    const int pc = (int)(p > 256u ? 256u : p);
    const int up = t.kRP + pc - 1;
    t.kRP = p ? (up > RL_KRP_MAX ? RL_KRP_MAX : up) : rl_max0(t.kRP - 2);
    t.kP = u ? rl_max0(t.kP - 1) : t.kP + RL_U0;                        // (u == 0 only in no-run mode here)
    t.m = 0;
}

// ---- decoder ---------------------------------------------------------------------------------------------
// MSB-first reader over big-endian 32-bit words; acc holds the next nbits bits left-aligned.
// Word sources: next() returns the following 32 bits of the stream (big-endian order restored).
// GlobalWords reads the stream in place, one word ahead of its use.
struct GlobalWords {
    const uint32_t *in;
    uint32_t n_words, pos;             // pos: index of `ahead`
    uint32_t ahead;                    // raw (still byte-swapped) so that nothing touches the loaded value before the next refill
    RL_HD uint32_t word(uint32_t i) const { return i < n_words ? in[i] : 0u; }
    RL_HD void start(const uint8_t *stream, int64_t readable_bytes, uint32_t first_word)
    {
        in = (const uint32_t *)stream;
        n_words = (uint32_t)((readable_bytes + 3) >> 2);
        pos = first_word;
        ahead = word(pos);
    }
    RL_HD uint32_t next()
    {
        const uint32_t w = rl_bswap32(ahead);
        ++pos;
        ahead = word(pos);
        return w;
    }
};

template <class Words> struct BitReaderT {
    // 64-bit window over a big-endian bit stream, refilled 32 bits at a time from a word source that runs ahead of
    // the window, so that a refill never waits for memory (on the GPU a warp refills on almost every symbol -- some
    // lane always runs dry).
    Words src;
    uint64_t acc;
    int nbits;
    RL_HD void init(const uint8_t *stream, int64_t readable_bytes, unsigned long long begin_bit)
    {
        src.start(stream, readable_bytes, (uint32_t)(begin_bit >> 5));
        acc = 0;
        nbits = 0;
        const unsigned skip_bits = (unsigned)(begin_bit & 31);
        if (skip_bits) { fill(); acc <<= skip_bits; nbits -= (int)skip_bits; }
    }
    RL_HD void fill()                  // afterwards nbits >= 33
    {
        if (nbits <= 32) {
            acc |= (uint64_t)src.next() << (32 - nbits);
            nbits += 32;
        }
    }
    RL_HD void skip(unsigned len)      // len <= nbits
    {
        acc <<= len;
        nbits -= (int)len;
    }
    RL_HD unsigned get1()
    {
        fill();
        const unsigned b = (unsigned)(acc >> 63);
        skip(1);
        return b;
    }
    RL_HD uint32_t get(unsigned len)   // len <= 32
    {
        if (len == 0) return 0;
        fill();
        const uint32_t v = (uint32_t)(acc >> (64 - len));
        skip(len);
        return v;
    }
    RL_HD int ones()                   // leading ones of the window after a refill, capped at 32
    {
        fill();
        const int p = rl_clz64(~acc);  // bits past nbits read as zero
        return p > 32 ? 32 : p;
    }
    RL_HD uint32_t gr(unsigned kr)     // membuf.cpp:228-240 (grRead); values are < 2^32 (written unmasked as 32 bits)
    {
        const int p = ones();
        if (p >= 32) {
            skip(32);
            return get(32);
        }
        skip((unsigned)p + 1);
        return (uint32_t)(((uint64_t)p << kr) + get(kr));        // kr can be 32 (k_RP = 128)
    }
};
using BitReader = BitReaderT<GlobalWords>;

// Events from reader br in state s.  dst is pre-zeroed: only non-zero symbols are stored, so zero runs cost
// nothing and the tail of the stream is never read.  Positions are 32-bit (the coder takes n < 2^31 symbols).
//   first_pos < 0: plain sequential decode of symbols [0, end) (reader at the stream start);
//   first_pos >= 0: the reader stands before an event whose non-zero symbol is dst[first_pos] (a sync entry);
//   the zeros in front of it are decoded and dropped, then decoding continues up to `end`.
// One loop body for both modes of the coder: the run prefix (k > 0 only) is the divergent part, the Golomb-Rice
// symbol, the adaptation and the store are shared (lanes of a warp sit in different modes all the time).
template <class Reader>
RL_HD void decode_range(Reader &br, State s, int64_t first_pos, int64_t end64, int flag_signed,
                                             int32_t *__restrict__ dst)
{
    bool synced = first_pos < 0;
    const uint32_t first = synced ? 0u : (uint32_t)first_pos, end = (uint32_t)end64;
    uint32_t pos = 0;
    while (pos < end) {                                                    // membuf.cpp:266-334
        int k = s.kP / RL_L;
        const unsigned kR = (unsigned)(s.kRP / RL_L);
        const bool run = k != 0;
        if (run) {
            uint32_t m = 0;
            bool bad = false;
            for (;;) {                                                     // complete runs: one 1 bit each
                const int p = br.ones();
                for (int j = 0; j < p; ++j) {
                    m += 1u << k;
                    s.kP += RL_U1;
                    k = s.kP / RL_L;
                    if (m >= (1u << 31) || k > 30) { bad = true; break; }  // corrupt stream guard
                }
                if (bad) break;
                if (p < 32) { br.skip((unsigned)p + 1); break; }
                br.skip(32);
            }
            if (bad) return;
            m += br.get((unsigned)k);
            if (synced) {
                if (m >= end - pos) break;                                 // the next non-zero lies past this range
                pos += m;
            } else {
                pos = first;
                synced = true;
            }
        }
        const uint32_t u = br.gr(kR);
        adapt_krp(s, (uint64_t)u >> kR);
        if (run || u) {                                                    // a non-zero symbol: u + 1 after a run
            if (!synced) { pos = first; synced = true; }
            const uint64_t val = (uint64_t)u + (run ? 1u : 0u);
            dst[pos] = flag_signed ? u2s(val) : (int32_t)val;
            s.kP = s.kP < RL_D1 ? 0 : s.kP - RL_D1;                        // D0 == D1
        } else {
            s.kP += RL_U0;
        }
        if (synced) ++pos;
    }
}

}  // namespace rlgr
