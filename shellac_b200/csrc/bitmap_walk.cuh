// bitmap_walk.cuh — framing decisions of HttpParser.parse read off two BITMAPS of a span of bytes.
//
// The route kernel (parse_kernel.cuh, phase A) classifies every byte pair of a group's span once,
// coalesced, and leaves two bits per byte in shared memory: "a CRLF starts here" and "a ': '
// starts here".  Everything HttpParser.parse decides about framing
// (/root/reference/src/python/shellac/server/HttpParser.py:150-186: find('\r\n'), find('\r\n\r\n'),
// and :326-328: every header line splits on ': ' into exactly two parts) is then bit arithmetic on
// those words, one message per lane:
//   * the first line ends at the first CRLF bit of the message;
//   * the header block ends at the first CRLF bit that has another CRLF bit two bytes on;
//   * "exactly one ': ' per header line" <=> CRLF and ': ' events ALTERNATE between those two
//     points.  With F = ~(cr | sep), the sum F + (cr << 1) carries a one from every CRLF through
//     the event-free bytes above it into the NEXT event; that event must be a ': ' and every ': '
//     must be reached that way: ((F + (cr << 1)) & (cr | sep)) == sep, one add per 32 bytes with
//     the carry chained from word to word.
// No compaction of the events into lists (round 1 spent 28 % of the kernel's instructions there).
//
// Coordinates: bit b of word w stands for the byte pair that STARTS at span position 32 w + b - 1
// ("q = p + 1": the classifier looks backwards, so the pair that straddles two words belongs to the
// upper one).  The code is plain C++ so that tests/bm_emu.cpp can run it on the CPU against the
// oracle (tests/test_bitmap_walk.py); the kernel includes the same file.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define SHL_BW __host__ __device__ __forceinline__
#else
#define SHL_BW inline
#endif

namespace shl {

#if defined(__CUDA_ARCH__)
SHL_BW int bw_ffs(uint32_t x) { return __ffs((int)x); }
SHL_BW int bw_popc(uint32_t x) { return __popc(x); }
#else
SHL_BW int bw_ffs(uint32_t x) { return __builtin_ffs((int)x); }
SHL_BW int bw_popc(uint32_t x) { return __builtin_popcount(x); }
#endif

// Bounds checks of our own (compute-sanitizer is not available on every pool): a build with
// -DSHL_BOUNDS_CHECK counts every shared-memory bitmap index and record slot that falls outside its
// array in g_shl_oob (read back with shl_debug_oob_count); the shipped build compiles them away.
#if defined(SHL_BOUNDS_CHECK) && defined(__CUDACC__)
__device__ unsigned int g_shl_oob;
#endif
#if defined(SHL_BOUNDS_CHECK) && defined(__CUDA_ARCH__)
#define SHL_CHK(cond) do { if (!(cond)) atomicAdd(&g_shl_oob, 1u); } while (0)
#else
#define SHL_CHK(cond) do { } while (0)
#endif
#ifndef SHL_BM_WORDS_FOR_CHECK
#define SHL_BM_WORDS_FOR_CHECK 0xffffffffu
#endif

struct alignas(8) bm2 { uint32_t x, y; };      // x: CRLF bits, y: ": " bits of 32 pair starts (layout of a uint2)

constexpr uint32_t kBmNone = 0xffffffffu;

struct BmWalk {
    int idx, T;            // message-relative position of the first line's CR / of the CR that starts the terminating CRLFCRLF (-1: none)
    uint32_t n_c;          // CRLFs from the first line's through T's (= header lines + 1) when T >= 0
    uint32_t q_idx;        // q coordinate of the first line's CR
    uint32_t bad;          // != 0: some header line does not hold exactly one ": "
};

// Message [s0, s0 + L) of the span.
SHL_BW void bm_walk(const bm2* bm, const uint32_t s0, const int L, BmWalk& o) {
    o.idx = o.T = -1; o.n_c = 0u; o.q_idx = 0u; o.bad = 0u;
    if (L < 2) return;
    const uint32_t qlo = s0 + 1u, qhi = s0 + (uint32_t)L - 1u;       // pair starts s0 .. s0 + L - 2
    const uint32_t wl = qhi >> 5;
    const uint32_t mhi = 0xffffffffu >> (31u - (qhi & 31u));        // bits <= qhi of word wl
    uint32_t w = qlo >> 5;
    bm2 cur = bm[w];
    uint32_t c = cur.x & (0xffffffffu << (qlo & 31u)) & (w == wl ? mhi : 0xffffffffu);
    SHL_CHK(wl < SHL_BM_WORDS_FOR_CHECK);
    while (!c && w < wl) { w++; cur = bm[w]; c = cur.x & (w == wl ? mhi : 0xffffffffu); }
    if (!c) return;                                                  // no CRLF at all
    const uint32_t b = (uint32_t)bw_ffs(c) - 1u;
    o.q_idx = (w << 5) + b;
    o.idx = (int)(o.q_idx - 1u - s0);
    // From the first line's CR on: count CRLFs, look for the first CRLF CRLF, check the alternation.
    const uint32_t from = 0xffffffffu << b;
    uint32_t cw = c;                                                 // (b is the lowest bit of c)
    uint32_t sw = cur.y & from & (w == wl ? mhi : 0xffffffffu);      // ": " inside the first line are not separators
    uint32_t cmpm = from, notfirst = ~(1u << b);                     // the block starts at idx + 2: a pair AT idx is no terminator
    uint32_t n = 0u, viol = 0u, carry = 0u, sc = 0u;
    for (;;) {
        bm2 nx; nx.x = nx.y = 0u;
        if (w < wl) { nx = bm[w + 1u]; if (w + 1u == wl) { nx.x &= mhi; nx.y &= mhi; } }
        const uint32_t E = cw | sw, F = ~E;
        const uint64_t sum = (uint64_t)F + (uint64_t)((cw << 1) | sc) + (uint64_t)carry;
        const uint32_t v = (((uint32_t)sum & E) ^ sw) & cmpm;        // events reached from a CRLF must be exactly the ": "
        const uint32_t d = cw & ((cw >> 2) | (nx.x << 30)) & notfirst;
        if (d) {
            const uint32_t t = (uint32_t)bw_ffs(d) - 1u;
            const uint32_t within = 0xffffffffu >> (31u - t);
            o.T = (int)((w << 5) + t - 1u - s0);
            o.n_c = n + (uint32_t)bw_popc(cw & within);
            o.bad = viol | (v & within);
            return;
        }
        n += (uint32_t)bw_popc(cw); viol |= v;
        carry = (uint32_t)(sum >> 32); sc = cw >> 31;
        if (w >= wl) { o.n_c = n; return; }                          // no terminator yet
        w++; cw = nx.x; sw = nx.y; cmpm = 0xffffffffu; notfirst = 0xffffffffu;
    }
}

// first set bit of component `sep ? y : x` at q' with q <= q' <= qend, or kBmNone
SHL_BW uint32_t bm_next(const bm2* bm, const bool sep, const uint32_t q, const uint32_t qend) {
    if (q > qend) return kBmNone;
    uint32_t w = q >> 5;
    const uint32_t wl = qend >> 5;
    uint32_t bits = (sep ? bm[w].y : bm[w].x) & (0xffffffffu << (q & 31u));
    while (!bits && w < wl) { w++; bits = sep ? bm[w].y : bm[w].x; }
    if (!bits) return kBmNone;
    const uint32_t r = (w << 5) + (uint32_t)bw_ffs(bits) - 1u;
    return r <= qend ? r : kBmNone;
}

// The header line that breaks the rule, for a block bm_walk flagged: lines before it become
// records (HttpParser.py:326-337 raises inside the loop, after having stored those).  Returns the
// line index; *no_sep tells a line without ": " from one with two or more.
SHL_BW uint32_t bm_offender(const bm2* bm, const uint32_t s0, const BmWalk& g, bool* no_sep) {
    const uint32_t q_T = s0 + 1u + (uint32_t)g.T;
    uint32_t prev = g.q_idx;
    uint32_t cur = bm_next(bm, true, prev + 1u, q_T);
    for (uint32_t line = 0; line + 1u < g.n_c; line++) {
        const uint32_t end = bm_next(bm, false, prev + 1u, q_T);
        const uint32_t nxt = cur != kBmNone ? bm_next(bm, true, cur + 1u, q_T) : kBmNone;
        if (cur > end) { *no_sep = true; return line; }
        if (nxt < end) { *no_sep = false; return line; }
        cur = nxt; prev = end;
    }
    *no_sep = true;
    return g.n_c - 1u;                 // not reached for a block that is really bad
}

// Forward cursor over one component: the caller knows the next bit exists (the block was
// validated), so there is no bound.
struct BmCursor { uint32_t w, bits; };
SHL_BW void bmc_init(BmCursor& c, const bm2* bm, const bool sep, const uint32_t q) {     // strictly after q
    c.w = q >> 5;
    c.bits = (sep ? bm[c.w].y : bm[c.w].x) & (0xfffffffeu << (q & 31u));
}
SHL_BW uint32_t bmc_next(BmCursor& c, const bm2* bm, const bool sep) {
    while (!c.bits) { c.w++; c.bits = sep ? bm[c.w].y : bm[c.w].x; }
    const uint32_t b = (uint32_t)bw_ffs(c.bits) - 1u;
    c.bits &= c.bits - 1u;
    return (c.w << 5) + b;
}


// Forward cursor over ALL events (cr | sep) of a validated block: there they alternate, so the
// separator and the end of line k are simply the next two events.  The word after the current one
// is already loaded when the current one runs dry (the shared-memory latency stays off the
// dependent chain).
struct EvCursor { uint32_t w, bits, nxt; };
SHL_BW void evc_init(EvCursor& c, const bm2* bm, const uint32_t q) {                    // strictly after q
    c.w = q >> 5;
    const bm2 a = bm[c.w], b = bm[c.w + 1u];
    c.bits = (a.x | a.y) & (0xfffffffeu << (q & 31u));
    c.nxt = b.x | b.y;
}
SHL_BW uint32_t evc_pop(EvCursor& c, const bm2* bm) {
    while (!c.bits) { c.w++; SHL_CHK(c.w + 1u < SHL_BM_WORDS_FOR_CHECK); c.bits = c.nxt; const bm2 b = bm[c.w + 1u]; c.nxt = b.x | b.y; }
    const uint32_t b = (uint32_t)bw_ffs(c.bits) - 1u;
    c.bits &= c.bits - 1u;
    return (c.w << 5) + b;
}

}  // namespace shl
