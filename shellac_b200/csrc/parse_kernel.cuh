// parse_kernel.cuh — K1: batched HTTP/1.1 tokeniser (+ optional fused key/MD5/ketama placement).
//
// Replaces HttpParser.parse / _parse_first_line / _parse_headers and the Content-Length framing
// decision (/root/reference/src/python/shellac/server/HttpParser.py:140-223, 304-337), driven
// over a whole message as the reference's own test loop does (test/HttpParserTests.py:36-39).
//
// Mapping (round 1, fifth design; DESIGN.md section 4 and profiles/r01/ have the history: a warp
// per message ran 960 warp-instructions per request, a lane per message 340, this one 277): one
// warp owns a GROUP of 32 consecutive messages.
//   phase A  the group's messages normally lie back to back, so the warp scans ONE span of bytes:
//            lane l classifies 32 bytes of every 1 KiB row (coalesced, next row in flight) for the
//            byte pairs "\r\n" and ": " (fp16x2 compares + HFMA2 gather, see pair_events_h) and
//            the set bits are compacted, in position order, into two flat u16 lists per group
//            with one warp prefix sum per row (group_list_events);
//   phase B  lane j == message j: binary search for the message's slice of the lists, walk its
//            CRLFs: first line end, CRLFCRLF, exactly one ": " per line (group_walk);
//   epilogue lane j == message j: first-line evaluation, one atomic record allocation per group,
//            value strip + record write, Content-Length / Transfer-Encoding decision,
//            struct-of-arrays stores (coalesced), MD5 + continuum search + shared-memory histogram.
// A group that is not a span (unordered or overlapping messages, bodies, too many events) is
// scanned message-per-lane (scan_lists) into lane-interleaved lists, same epilogue.  Anything off
// the canonical shape (responses' first line, whitespace around the header block, more lines than
// the per-lane lists hold) takes a per-lane scalar path with the same semantics.
#pragma once
#include <cuda_fp16.h>
#include "shl_device.cuh"
#ifndef SHL_OPT_SPANMAX
#define SHL_OPT_SPANMAX 20480         // bytes of span a warp keeps bitmaps for (config 2: 16.4 KiB +- 0.4)
#endif
#define SHL_BM_WORDS_FOR_CHECK (SHL_OPT_SPANMAX / 32u + 2u)
#include "bitmap_walk.cuh"

namespace shl {

// build-time switches kept for A/B timing (tools/build_variants.py)
#ifndef SHL_OPT_RINGDIR
#define SHL_OPT_RINGDIR 1             // continuum search through the 4 096-bucket directory
#endif
#ifndef SHL_OPT_MINBLOCKS
#define SHL_OPT_MINBLOCKS 6           // resident blocks per SM the register allocation aims at
#endif
#ifndef SHL_OPT_BATCH
#define SHL_OPT_BATCH 2               // header lines whose value peeks are in flight together
#endif
#ifndef SHL_OPT_WARPS
#define SHL_OPT_WARPS 4
#endif
constexpr int kWarpsPerBlock = SHL_OPT_WARPS;
constexpr int kGroup = 32;            // messages per warp (one per lane)
constexpr int kCrCap = 24;            // CRLF positions a lane can list (first line + 23 header lines)
constexpr int kSepCap = 28;           // ": " positions a lane can list

struct Msg {                          // per-lane result of one message
    uint32_t status, first_line_len, headers_len, sp1, sp2, fl_end, slash, n_headers;
    int32_t status_code;
    int64_t content_len;
};


// ----------------------------------------------------------------------------------- SWAR
// The replicated byte constants arrive as kernel arguments (RouteArgs::k), not as literals: as
// run-time values they stay in registers / the constant bank, so `(x ^ c) & m7` is ONE LOP3 and
// the `+ m7` is an IMAD (x * one + m7) on the FMA pipe.  The scan is bound by the ALU pipe
// (LOP3/SHF/IADD issue at half rate per scheduler), so every op moved to the FMA pipe counts.
struct SwarK { uint32_t m7, hi, cr, lf, col, sp, one; };

__host__ __device__ inline SwarK swar_host_consts() {
    SwarK k;
    k.m7 = 0x7f7f7f7fu; k.hi = 0x80808080u; k.cr = 0x0d0d0d0du; k.lf = 0x0a0a0a0au;
    k.col = 0x3a3a3a3au; k.sp = 0x20202020u; k.one = 1u;
    return k;
}
// 0x80 in every byte of x that equals the (7-bit) byte replicated in c
__device__ __forceinline__ uint32_t eqf(uint32_t x, uint32_t c, const SwarK& k) {
    const uint32_t t = ((x ^ c) & k.m7) * k.one + k.m7;
    return ~(t | x) & k.hi;
}
// ... restricted to the bytes flagged (0x80) in `also`
__device__ __forceinline__ uint32_t eqf_and(uint32_t x, uint32_t c, uint32_t also, const SwarK& k) {
    const uint32_t t = ((x ^ c) & k.m7) * k.one + k.m7;
    return ~(t | x) & also;
}
// gather the 0x80 flags of four words into a 16-bit position mask (bit j = byte j)
__device__ __forceinline__ uint32_t mask16(uint32_t f0, uint32_t f1, uint32_t f2, uint32_t f3) {
    const uint32_t h0 = __umulhi(f0, 0x02040810u) & 0xfu, h1 = __umulhi(f1, 0x02040810u) & 0xfu;
    const uint32_t h2 = __umulhi(f2, 0x02040810u), h3 = __umulhi(f3, 0x02040810u);   // stray bits land above bit 15
    return (h3 * 4096u + (h2 * 256u + (h1 * 16u + h0))) & 0xffffu;
}
__device__ __forceinline__ int clamp16(int v) { return min(max(v, 0), 16); }
__device__ __forceinline__ uint32_t below(int k) { return (1u << k) - 1u; }            // k in [0,16]
__device__ __forceinline__ uint32_t rng16(int lo, int hi) { return below(clamp16(hi)) & ~below(clamp16(lo)); }

struct RouteArgs {
    const uint8_t* arena; const uint64_t* off; const uint32_t* len; uint64_t n;
    shl_parse_out out; uint64_t* code_hist;
    int force_scalar;                 // test hook: take the scalar path for every message
    SwarK k;                          // SWAR byte constants, as run-time values (see SwarK)
    uint32_t hdr_base_bias;           // added to the stored hdr_base (out.hdr may be a slice of a larger arena)
    shl_msg_rec* crec;                // != null: compact output, one 32-byte record per message instead of the arrays of `out`
    uint32_t group;                   // messages per warp, 1..32 (small batches: fewer per warp, more warps, shorter chains)
    // fused placement (route); ring_values == nullptr -> parse only
    const uint32_t* ring_values; const uint32_t* ring_nodes; const uint32_t* ring_dir; uint32_t n_points, n_hosts;
    shl_place_out place;
};

// ----------------------------------------------------------------------------------- scalar
// HttpParser._parse_first_line (HttpParser.py:304-321) on m[0:idx+2).  Warp-uniform.
__device__ __noinline__ uint32_t first_line_scalar(const uint8_t* m, int idx, Msg& r, bool& get_or_head) {
    int e = idx + 2;
    while (e > 0 && is_ws(ldb(m + e - 1))) e--;
    int sp1 = -1, sp2 = -1;
    for (int i = 0; i < e; i++) if (ldb(m + i) == ' ') { sp1 = i; break; }
    if (sp1 >= 0) for (int i = sp1 + 1; i < e; i++) if (ldb(m + i) == ' ') { sp2 = i; break; }
    if (sp1 < 0 || sp2 < 0) return SHL_E_FIRSTLINE_PARTS;
    r.sp1 = (uint32_t)sp1; r.sp2 = (uint32_t)sp2; r.fl_end = (uint32_t)e;
    const bool is_resp = sp1 >= 2 && ldb(m) == 'H' && ldb(m + 1) == 'T';
    const int a = is_resp ? 0 : sp2 + 1, b = is_resp ? sp1 : e;
    int slash = -1, ns = 0;
    for (int i = a; i < b; i++) if (ldb(m + i) == '/') { if (!ns) slash = i; ns++; }
    if (is_resp) r.status &= ~SHL_F_IS_REQUEST;
    if (ns != 1) return SHL_E_VERSION_SPLIT;
    r.slash = (uint32_t)slash;
    if (!py_float_ok(m + slash + 1, b - slash - 1)) return SHL_E_VERSION_FLOAT;
    uint32_t vc = 2;
    if (span_eq(m, slash + 1, b, "1.1")) vc = 1; else if (span_eq(m, slash + 1, b, "1.0")) vc = 0;
    r.status = (r.status & ~SHL_F_VERSION_MASK) | (vc << SHL_F_VERSION_SHIFT);
    if (is_resp) {
        int64_t v = 0;
        const int k = py_int10(m + sp1 + 1, sp2 - sp1 - 1, &v);
        if (k == 0) return SHL_E_STATUS_INT;
        if (k == 2 || v > 2147483647ll || v < -2147483648ll) r.status |= SHL_F_STATUS_WIDE;
        else r.status_code = (int32_t)v;
    } else {
        get_or_head = span_eq(m, 0, sp1, "GET") || span_eq(m, 0, sp1, "HEAD");
    }
    r.status |= SHL_F_FIRST_LINE_DONE;
    return 0;
}

struct WellKnown { int n_te, n_cl; uint32_t te_vs, te_ve, cl_vs, cl_ve; };

// HttpParser._parse_headers (HttpParser.py:323-337) on m[h0:h1), h1 = one past CRLFCRLF.
// Records go to dst[k * stride].
__device__ __noinline__ uint32_t headers_scalar(const uint8_t* m, int h0, int h1, shl_hdr_span* dst, int stride,
                                                uint32_t cap, uint32_t* n_out, WellKnown* wk) {
    int s = h0, e = h1;
    while (s < e && is_ws(ldb(m + s))) s++;
    while (e > s && is_ws(ldb(m + e - 1))) e--;
    uint32_t n = 0;
    int ls = s;
    for (;;) {
        int le = -1;
        for (int i = ls; i + 1 < e; i++) if (ldb(m + i) == '\r' && ldb(m + i + 1) == '\n') { le = i; break; }
        const int lend = le < 0 ? e : le;
        int sep = -1, nsep = 0;
        for (int i = ls; i + 1 < lend; i++)
            if (ldb(m + i) == ':' && ldb(m + i + 1) == ' ') { if (!nsep) sep = i; nsep++; i++; }
        if (nsep != 1) { *n_out = n; return nsep == 0 ? SHL_E_HDR_NO_SEP : SHL_E_HDR_MULTI_SEP; }
        int vs = sep + 2, ve = lend;
        while (vs < ve && is_ws(ldb(m + vs))) vs++;
        while (ve > vs && is_ws(ldb(m + ve - 1))) ve--;
        if (n < cap) {
            shl_hdr_span rec;
            rec.key_off = (uint16_t)ls; rec.key_len = (uint16_t)(sep - ls);
            rec.val_off = (uint16_t)vs; rec.val_len = (uint16_t)(ve - vs);
            dst[(size_t)n * stride] = rec;
        }
        n++;
        if (span_eq_lower(m, ls, sep, "transfer-encoding")) { if (!wk->n_te) { wk->te_vs = vs; wk->te_ve = ve; } wk->n_te++; }
        if (span_eq_lower(m, ls, sep, "content-length")) { if (!wk->n_cl) { wk->cl_vs = vs; wk->cl_ve = ve; } wk->n_cl++; }
        if (le < 0) break;
        ls = le + 2;
    }
    *n_out = n;
    return 0;
}

// Header-state epilogue (HttpParser.py:187-200) + the Content-Length body call (:206-223).
__device__ __forceinline__ void finalize_headers(const uint8_t* m, int L, int body_off, bool get_or_head,
                                                 const WellKnown& wk, uint32_t& status, int64_t& content_len) {
    status |= SHL_F_HEADERS_COMPLETE | SHL_F_ON_BODY;
    bool complete = (status & SHL_F_IS_REQUEST) && get_or_head;
    const bool chunked = wk.n_te == 1 && span_eq(m, wk.te_vs, wk.te_ve, "chunked");
    if (chunked) status |= SHL_F_CHUNKED;
    else {
        if (wk.n_cl > 1) { status |= SHL_E_CLEN_LIST | (complete ? SHL_F_MESSAGE_COMPLETE : 0u); return; }
        if (wk.n_cl == 1) {
            int64_t v = 0;
            const int k = py_int10(m + wk.cl_vs, (int64_t)wk.cl_ve - wk.cl_vs, &v);
            if (k == 0) { status |= SHL_E_CLEN_INT | (complete ? SHL_F_MESSAGE_COMPLETE : 0u); return; }
            if (k == 2) { status |= SHL_F_CLEN_WIDE; v = ldb(m + wk.cl_vs) == '-' ? INT64_MIN : INT64_MAX; }
            content_len = v;
        }
        if (content_len == 0 && !(status & SHL_F_CLEN_WIDE)) { status &= ~SHL_F_ON_BODY; complete = true; }
    }
    if (complete) { status |= SHL_F_MESSAGE_COMPLETE | SHL_S_OK; return; }
    const int64_t avail = (int64_t)L - body_off;
    if (chunked) { status |= SHL_S_INCOMPLETE_BODY; return; }             // K2 (frame kernel) decides
    if (avail > 0 && avail >= content_len && !((status & SHL_F_CLEN_WIDE) && content_len > 0)) {
        status &= ~SHL_F_ON_BODY;
        status |= SHL_F_MESSAGE_COMPLETE | SHL_S_OK;
    } else status |= SHL_S_INCOMPLETE_BODY;
}

// ----------------------------------------------------------------------------------- scan: one message, one lane
struct LaneScan { int idx, T; uint32_t n_c, n_s, s_skip; bool ovf, bad; };

// cr[k * 32] / sp[k * 32]: this lane's slots of the warp's position lists (u16, lane-interleaved).
// 0xFFFF in each 16-bit half of x that equals the matching half of pat.  The halves are compared
// as fp16 numbers by ONE HSET2 on the FMA/half pipe; that is exact here because both patterns
// (0x0A0D = "\r\n", 0x203A = ": ", little-endian) are normal, non-zero fp16 values, so numeric
// equality is bit equality (NaN halves compare false, +-0 and denormals are not the patterns).
__device__ __forceinline__ uint32_t eq16x2(uint32_t x, uint32_t pat) {
    return __heq2_mask(*reinterpret_cast<const __half2*>(&x), *reinterpret_cast<const __half2*>(&pat));
}

// Event bits of one 32-bit word w (pv = the word BEFORE it in memory), position-ordered, for the
// four byte pairs that END inside w: bit 2j = a ": " starts at byte j-1 of w, bit 2j+1 = a CRLF
// starts there (j = 0 is the pair that straddles pv and w).  Looking backwards means the scan
// carries one word and never loads ahead of the byte it is classifying.
__device__ __forceinline__ uint32_t pair_events(uint32_t pv, uint32_t w) {
    const uint32_t sh = __funnelshift_r(pv, w, 24);                    // bytes -1..2: the odd alignment
    const uint32_t fc = __byte_perm(eq16x2(w, 0x0A0D0A0Du), eq16x2(sh, 0x0A0D0A0Du), 0x2604);   // 0xFF per pair that is a CRLF
    const uint32_t fs = __byte_perm(eq16x2(w, 0x203A203Au), eq16x2(sh, 0x203A203Au), 0x2604);   // ... that is a ": "
    return __umulhi((fc & 0x80808080u) | (fs & 0x40404040u), 0x04104100u);   // [s0 c0 s1 c1 s2 c2 s3 c3] + stray bits above bit 7
}

#ifndef SHL_OPT_HGATHER
#define SHL_OPT_HGATHER 1
#endif
// The same four pairs, gathered on the half-precision FMA pipe instead of with PRMT / LOP3 (the
// kernel is bound by the integer ALU pipe): HSET2.BF gives 1.0 per matching half, four HFMA2 add
// them up with power-of-two weights on top of 1024.0, whose fp16 mantissa then IS the bit field:
// low half = [s-1 c-1 s0 c0], high half = the pairs at bytes 1 and 2, weighted 16x.
// Returns 0x64hh64ll-style halves; (x * 0x10001) has lo + hi in byte 2.
__device__ __forceinline__ uint32_t pair_events_h(uint32_t pv, uint32_t w) {
    const uint32_t sh = __funnelshift_r(pv, w, 24);
    const uint32_t kCR = 0x0A0D0A0Du, kSEP = 0x203A203Au;
    const uint32_t k1 = 0x4C003C00u, k2 = 0x50004000u, k4 = 0x54004400u, k8 = 0x58004800u, kM = 0x64006400u;   // (1,16) (2,32) (4,64) (8,128) (1024,1024)
#define SHL_H2(x) (*reinterpret_cast<const __half2*>(&(x)))
    __half2 acc = __hfma2(__heq2(SHL_H2(sh), SHL_H2(kSEP)), SHL_H2(k1), SHL_H2(kM));
    acc = __hfma2(__heq2(SHL_H2(sh), SHL_H2(kCR)), SHL_H2(k2), acc);
    acc = __hfma2(__heq2(SHL_H2(w), SHL_H2(kSEP)), SHL_H2(k4), acc);
    acc = __hfma2(__heq2(SHL_H2(w), SHL_H2(kCR)), SHL_H2(k8), acc);
#undef SHL_H2
    return *reinterpret_cast<const uint32_t*>(&acc) * 0x00010001u;
}

// ... of one 16-byte vector (c = the word before it): bit pair t <-> the byte pair that starts at byte t - 1
__device__ __forceinline__ uint32_t vec_events(const uint32_t c, const uint4& v) {
#if SHL_OPT_HGATHER
    const uint32_t e0 = pair_events_h(c, v.x), e1 = pair_events_h(v.x, v.y), e2 = pair_events_h(v.y, v.z), e3 = pair_events_h(v.z, v.w);
    return __byte_perm(__byte_perm(e0, e1, 0x0062), __byte_perm(e2, e3, 0x0062), 0x5410);     // byte 2 of each
#else
    return (pair_events(c, v.x) & 0xffu) | ((pair_events(v.x, v.y) & 0xffu) << 8) |
           ((pair_events(v.y, v.z) & 0xffu) << 16) | (pair_events(v.z, v.w) << 24);
#endif
}

// Running state of one lane's scan.  Events are only LISTED here (positions of the CR of every
// CRLF, of the ':' of every ": "); which separator belongs to which line is worked out afterwards
// from the positions (count_lines), when the lanes are converged again.
struct ScanState {
    int T, last_cr;
    uint32_t n_c, n_s;
    uint32_t s_prev, s_skip, bad;      // ": " listed before the last CRLF / inside the first line; some line did not hold exactly one
};

// One 16-byte vector v at message position p (byte 0 of v), c = the word before it.
__device__ __forceinline__ void scan_step(const uint4& v, const uint32_t c, const int p, const int L,
                                          uint16_t* cr, uint16_t* sp, const SwarK& K, ScanState& st) {
    uint32_t ev = vec_events(c, v);
    // bit pair t <-> the byte pair (p + t - 1, p + t): it must start at >= 0 and end at <= L - 1
    if (p <= 0 || p + 16 > L) {
        const int lo = clamp16(1 - p), hi = clamp16(L - p);
        const uint32_t mhi = hi >= 16 ? 0xffffffffu : (1u << (2 * hi)) - 1u;
        const uint32_t mlo = lo >= 16 ? 0xffffffffu : (1u << (2 * lo)) - 1u;
        ev &= mhi & ~mlo;
    }
    uint32_t sb = ev & 0x55555555u, cb = ev & 0xaaaaaaaau;
    const int q = p - 1;
    const uint32_t sb0 = sb, n_s0 = st.n_s;
    if (sb) {                                      // the common vector holds at most one ": " ...
        const int b = __ffs(sb) - 1;
        sb &= sb - 1u;
        sp[min(st.n_s, (uint32_t)kSepCap - 1u) * 32] = (uint16_t)(q + (b >> 1));
        st.n_s++;
    }
    if (cb) {                                      // ... and at most one CRLF: straight-line code
        const int b = __ffs(cb) - 1;
        cb &= cb - 1u;
        const int pos = q + (b >> 1);
        if (pos == st.last_cr + 2 && st.n_c >= 2u) { st.T = st.last_cr; return; }     // empty line after the first line: CRLFCRLF
        const uint32_t seen = n_s0 + (uint32_t)__popc(sb0 & ((1u << b) - 1u));       // ": " before this CR
        if (st.n_c == 0u) st.s_skip = seen; else st.bad |= seen - st.s_prev - 1u;     // a header line must add exactly one
        st.s_prev = seen;
        cr[min(st.n_c, (uint32_t)kCrCap - 1u) * 32] = (uint16_t)pos;
        st.n_c++; st.last_cr = pos;
    }
    if (sb | cb) {                                 // the rest (a terminator inside one vector, short lines)
        while (sb) {
            const int b = __ffs(sb) - 1;
            sb &= sb - 1u;
            sp[min(st.n_s, (uint32_t)kSepCap - 1u) * 32] = (uint16_t)(q + (b >> 1));
            st.n_s++;
        }
        while (cb) {
            const int b = __ffs(cb) - 1;
            cb &= cb - 1u;
            const int pos = q + (b >> 1);
            if (pos == st.last_cr + 2 && st.n_c >= 2u) { st.T = st.last_cr; return; }
            const uint32_t seen = n_s0 + (uint32_t)__popc(sb0 & ((1u << b) - 1u));
            if (st.n_c == 0u) st.s_skip = seen; else st.bad |= seen - st.s_prev - 1u;
            st.s_prev = seen;
            cr[min(st.n_c, (uint32_t)kCrCap - 1u) * 32] = (uint16_t)pos;
            st.n_c++; st.last_cr = pos;
        }
    }
}

__device__ __forceinline__ void scan_lists(const uint8_t* m, const int L, uint16_t* cr, uint16_t* sp, const SwarK& K, LaneScan& o) {
    const int lead = (int)(reinterpret_cast<uintptr_t>(m) & 31);
    const uint4* vp = reinterpret_cast<const uint4*>(m - lead);
    const int nsec = (lead + L + 31) >> 5;         // 32-byte sectors that hold message bytes
    int p = -lead;                                 // message position of byte 0 of the current sector
    ScanState st; st.T = -1; st.last_cr = -4; st.n_c = st.n_s = st.s_prev = st.s_skip = st.bad = 0u;
    uint32_t c = 0u;                               // word before the current vector (masked off in sector 0)
    uint4 a0 = __ldg(vp), b0 = __ldg(vp + 1), a1, b1;
    a1 = b1 = make_uint4(0u, 0u, 0u, 0u);
    for (int s = 0; s < nsec; s++, p += 32, vp += 2) {
        if (s + 1 < nsec) { a1 = __ldg(vp + 2); b1 = __ldg(vp + 3); }
        scan_step(a0, c, p, L, cr, sp, K, st);
        if (st.T >= 0) break;
        scan_step(b0, a0.w, p + 16, L, cr, sp, K, st);
        if (st.T >= 0) break;
        c = b0.w; a0 = a1; b0 = b1;
    }
    o.T = st.T; o.n_c = st.n_c; o.n_s = st.n_s;
    o.idx = st.n_c ? (int)cr[0] : -1;
    o.s_skip = st.s_skip; o.bad = st.bad != 0u;
    o.ovf = st.n_c > (uint32_t)kCrCap || st.n_s > (uint32_t)kSepCap;       // the lists are incomplete: scalar path
}

// "Exactly one ': ' per header line" (HttpParser.py:326-328), from the two position lists: line k
// runs from cr[k] + 2 to cr[k + 1]; the separators are listed in position order, so line k must
// hold the next unused one and the one after that must lie beyond the line.  Returns the number
// of lines that become records (those before the first offender), the error code of the offender
// (or 0) in *err, and in *s_skip how many listed separators sit inside the first line.
__device__ __forceinline__ uint32_t count_lines(const uint16_t* cr, const uint16_t* sp, const LaneScan& sc, uint32_t* err, uint32_t* s_skip) {
    const uint32_t n_lines = sc.n_c - 1u;
    uint32_t j = 0;
    while (j < sc.n_s && (int)sp[j * 32] < sc.idx) j++;
    *s_skip = j;
    *err = 0u;
    uint32_t cur = j < sc.n_s ? (uint32_t)sp[j * 32] : 0xffffffffu;
    for (uint32_t k = 0; k < n_lines; k++) {
        const uint32_t end = cr[(k + 1u) * 32];
        const uint32_t nxt = j + 1u < sc.n_s ? (uint32_t)sp[(j + 1u) * 32] : 0xffffffffu;
        if (cur > end) { *err = SHL_E_HDR_NO_SEP; return k; }          // (cur == end cannot be: that byte is the CR)
        if (nxt < end) { *err = SHL_E_HDR_MULTI_SEP; return k; }
        cur = nxt; j++;
    }
    return n_lines;
}

// ----------------------------------------------------------------------------------- scan: one GROUP, one warp
// The 32 messages of a group usually lie back to back in the arena (that is how every packer in
// this repo, and a recv() loop, lays them out).  Then the warp does not scan 32 messages in 32
// lanes, it scans ONE span of bytes with 32 lanes:
//   phase A  lane l classifies bytes [1024 r + 32 l, +32) of the span in row r (fully coalesced,
//            every lane busy until the last row, no per-message edges) and stores the result as
//            two 32-bit words, "a CRLF starts here" / "a ': ' starts here", one bit per byte, in
//            shared memory: 2 bits per span byte, no compaction, no prefix sums;
//   phase B  lane j reads message j's framing off those bitmaps (bitmap_walk.cuh): first line end,
//            CRLFCRLF, and "one ': ' per line" as an alternation check with a carry chain.
// A group that does not qualify (unordered or overlapping messages, a span over kSpanMax) is
// scanned message-per-lane by scan_lists, with identical results.
constexpr uint32_t kSpanMax = SHL_OPT_SPANMAX;
constexpr uint32_t kBmWords = kSpanMax / 32u + 2u;         // + the straddling word + one zero word behind the last
static_assert(kBmWords * 8u >= (kCrCap * 32u + 2u + kSepCap * 32u + 2u) * 2u, "the per-lane lists share the bitmaps' storage");

// Event nibbles of one 32-bit word w (pv = the word BEFORE it in memory) for the four byte pairs
// that END inside w: bit j = a ": " starts at byte j - 1 of w, bit 4 + j = a CRLF starts there
// (j = 0 is the pair that straddles pv and w).  HSET2.BF gives 1.0 per matching half, four HFMA2
// add them up with power-of-two weights on top of 1024.0, whose fp16 mantissa then IS the bit
// field (low half: pairs -1 and 0, high half: pairs 1 and 2, disjoint bits); x * 0x10001 has
// lo + hi in byte 2.  The kernel is bound by the integer ALU pipe, so the gather runs on the
// half-precision FMA pipe on purpose.
__device__ __forceinline__ uint32_t pair_nibbles(uint32_t pv, uint32_t w) {
    const uint32_t sh = __funnelshift_r(pv, w, 24);
    const uint32_t kCR = 0x0A0D0A0Du, kSEP = 0x203A203Au;
    const uint32_t kS1 = 0x44003C00u, kS2 = 0x48004000u, kC1 = 0x54004C00u, kC2 = 0x58005000u, kM = 0x64006400u;   // (1,4) (2,8) (16,64) (32,128) (1024,1024)
#define SHL_H2(x) (*reinterpret_cast<const __half2*>(&(x)))
    __half2 acc = __hfma2(__heq2(SHL_H2(sh), SHL_H2(kSEP)), SHL_H2(kS1), SHL_H2(kM));
    acc = __hfma2(__heq2(SHL_H2(sh), SHL_H2(kCR)), SHL_H2(kC1), acc);
    acc = __hfma2(__heq2(SHL_H2(w), SHL_H2(kSEP)), SHL_H2(kS2), acc);
    acc = __hfma2(__heq2(SHL_H2(w), SHL_H2(kCR)), SHL_H2(kC2), acc);
#undef SHL_H2
    return *reinterpret_cast<const uint32_t*>(&acc) * 0x00010001u;
}

// The two bitmap words of the 32 bytes (va, vb), c = the word before them.  Words 0, 2, 4, 6 and
// 1, 3, 5, 7 are packed separately so that one shift and one LOP3 per bitmap interleave their
// nibbles back into position order.
__device__ __forceinline__ bm2 window_bits(const uint32_t c, const uint4& va, const uint4& vb) {
    const uint32_t e0 = pair_nibbles(c, va.x), e1 = pair_nibbles(va.x, va.y), e2 = pair_nibbles(va.y, va.z), e3 = pair_nibbles(va.z, va.w);
    const uint32_t e4 = pair_nibbles(va.w, vb.x), e5 = pair_nibbles(vb.x, vb.y), e6 = pair_nibbles(vb.y, vb.z), e7 = pair_nibbles(vb.z, vb.w);
    const uint32_t ev = __byte_perm(__byte_perm(e0, e2, 0x0062), __byte_perm(e4, e6, 0x0062), 0x5410);    // byte 2 of each
    const uint32_t od = __byte_perm(__byte_perm(e1, e3, 0x0062), __byte_perm(e5, e7, 0x0062), 0x5410);
    bm2 r;
    r.y = (ev & 0x0f0f0f0fu) | ((od << 4) & 0xf0f0f0f0u);
    r.x = ((ev >> 4) & 0x0f0f0f0fu) | (od & 0xf0f0f0f0u);
    return r;
}

// bitmaps of span[0:nbytes) (A0 is 32-byte aligned, nbytes <= kSpanMax)
#ifndef SHL_OPT_PREFETCH
#define SHL_OPT_PREFETCH 0
#endif
#ifndef SHL_OPT_ROWS_NOALLOC
#define SHL_OPT_ROWS_NOALLOC 0
#endif
#ifndef SHL_OPT_ROWS_AHEAD
#define SHL_OPT_ROWS_AHEAD 2          // rows of the span in flight ahead of the one being classified
#endif
__device__ __forceinline__ void row_load(const uint8_t* A0, const uint32_t off, const uint32_t nbytes, uint4& a, uint4& b) {
    a = b = make_uint4(0u, 0u, 0u, 0u);
    if (off < nbytes) {
#if SHL_OPT_ROWS_NOALLOC
        // the span streams through once: keep it out of L1, which the epilogue's first-line / URL / value-edge reads need
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w) : "l"(A0 + off));
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w) : "l"(A0 + off + 16));
#else
        a = __ldg(reinterpret_cast<const uint4*>(A0 + off)); b = __ldg(reinterpret_cast<const uint4*>(A0 + off + 16));
#endif
    }
}
__device__ __forceinline__ void group_bitmaps(const uint8_t* A0, const uint32_t nbytes, bm2* bm, const int lane) {
    uint32_t prev_last = 0u;
    uint32_t boff = (uint32_t)lane * 32u;
    uint4 va, vb, na, nb;
    row_load(A0, boff, nbytes, va, vb);
#if SHL_OPT_ROWS_AHEAD >= 2
    uint4 ma, mb;
    row_load(A0, boff + 1024u, nbytes, na, nb);
#endif
    uint32_t w = (uint32_t)lane;
    for (; boff - (uint32_t)lane * 32u < nbytes; boff += 1024u, w += 32u) {      // warp-uniform trip count
#if SHL_OPT_ROWS_AHEAD >= 2
        row_load(A0, boff + 2048u, nbytes, ma, mb);
#else
        row_load(A0, boff + 1024u, nbytes, na, nb);
#endif
        uint32_t c = __shfl_up_sync(SHL_FULL, vb.w, 1);
        if (lane == 0) c = prev_last;
        prev_last = __shfl_sync(SHL_FULL, vb.w, 31);
        SHL_CHK(w < kBmWords);
        bm[w] = window_bits(c, va, vb);
        va = na; vb = nb;
#if SHL_OPT_ROWS_AHEAD >= 2
        na = ma; nb = mb;
#endif
    }
    if (lane == 0) { bm2 z; z.x = z.y = 0u; bm[w] = z; }                        // the word behind the last row
}

// ----------------------------------------------------------------------------------- phase A, staged by TMA
// Build option (SHL_OPT_TMA=1): the span's rows do not come through the warps' registers with LDG,
// they are STAGED: one elected lane issues a bulk asynchronous copy (cp.async.bulk, the TMA unit;
// UBLKCP in SASS) of each 1 KiB row into a small per-warp ring in shared memory and arms an
// mbarrier with the byte count; the warp waits on the barrier's phase, every lane reads its 32
// bytes with two LDS.128, and the slot is refilled kTmaSlots rows ahead.  Same bitmaps out.
// Measured against the LDG path in profiles/r02/ (the ring costs shared memory, i.e. L1 and
// resident warps; the LDG path with two rows in flight already hides the row latency).
#ifndef SHL_OPT_TMA
#define SHL_OPT_TMA 0
#endif
constexpr int kTmaSlots = 3;
constexpr uint32_t kTmaRow = 1024u;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy (16-byte aligned, size a multiple of 16), completion counted on `bar`
__device__ __forceinline__ void tma_row(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// `it` is the warp's running row count (slot and barrier phase follow from it across groups)
__device__ __forceinline__ void group_bitmaps_tma(const uint8_t* A0, const uint32_t nbytes, bm2* bm, const int lane,
                                                  uint8_t* ring, uint64_t* bars, uint32_t& it) {
    const uint32_t rows = (nbytes + kTmaRow - 1u) / kTmaRow;
    auto issue = [&](uint32_t r) {
        const uint32_t left = nbytes - r * kTmaRow;
        const uint32_t bytes = min(kTmaRow, (left + 15u) & ~15u);      // past the span only inside the arena's padding
        const uint32_t slot = (it + r) % kTmaSlots;
        mbar_expect_tx(&bars[slot], bytes);
        tma_row(ring + slot * kTmaRow, A0 + (size_t)r * kTmaRow, bytes, &bars[slot]);
    };
    if (lane == 0) for (uint32_t r = 0; r < rows && r < (uint32_t)kTmaSlots; r++) issue(r);
    uint32_t prev_last = 0u, w = (uint32_t)lane;
    for (uint32_t r = 0; r < rows; r++, w += 32u) {
        const uint32_t slot = (it + r) % kTmaSlots, parity = ((it + r) / kTmaSlots) & 1u;
        mbar_wait(&bars[slot], parity);
        const uint4* t = reinterpret_cast<const uint4*>(ring + slot * kTmaRow + (uint32_t)lane * 32u);
        const uint4 va = t[0], vb = t[1];
        uint32_t c = __shfl_up_sync(SHL_FULL, vb.w, 1);
        if (lane == 0) c = prev_last;
        prev_last = __shfl_sync(SHL_FULL, vb.w, 31);
        SHL_CHK(w < kBmWords);
        bm[w] = window_bits(c, va, vb);
        __syncwarp();                                                   // every lane has read the slot: it may be refilled
        if (lane == 0 && r + (uint32_t)kTmaSlots < rows) issue(r + (uint32_t)kTmaSlots);
    }
    it += rows;
    if (lane == 0) { bm2 z; z.x = z.y = 0u; bm[w] = z; }
}

// ----------------------------------------------------------------------------------- epilogue helpers
// Canonical request line  METHOD SP URL SP "HTTP/1.x" CRLF  with a method of at most 7 bytes:
// everything is decided from the line's first 8 and last 9 bytes (five independent loads, one
// round trip), EXCEPT that the URL must not hold a space; the caller checks that while it fetches
// the URL's words for the MD5 (md5_key_t<true>) and falls back to first_line_scalar like for any
// other shape.  Fills the first-line fields and returns true when the shape holds.
__device__ __forceinline__ bool first_line_canon(const uint8_t* m, const int idx, Msg& r, bool& get_or_head) {
    if (idx < 14) return false;                                           // "X / HTTP/1.1" and shorter: the scalar path is exact
    const uint64_t head = ld64u(m), tail = ld64u(m + idx - 8);
    const uint32_t sp2c = ldb(m + idx - 9);
    const uint32_t f4 = (uint32_t)head, w0 = (uint32_t)tail, w1 = (uint32_t)(tail >> 32);
    if (w0 != 0x50545448u || (w1 != 0x312E312Fu && w1 != 0x302E312Fu) || sp2c != ' ') return false;
    const uint32_t z0 = has_zero_byte(f4 ^ 0x20202020u), z1 = has_zero_byte((uint32_t)(head >> 32) ^ 0x20202020u);
    if (!(z0 | z1)) return false;                                         // no space in the first 8 bytes
    const int sp1 = z0 ? ((__ffs(z0) - 1) >> 3) : 4 + ((__ffs(z1) - 1) >> 3);
    const int sp2 = idx - 9;
    if (sp2 <= sp1) return false;
    if ((f4 & 0xffffu) == 0x5448u && sp1 >= 2) return false;            // a.startswith('HT'): a response
    r.sp1 = (uint32_t)sp1; r.sp2 = (uint32_t)sp2; r.fl_end = (uint32_t)idx; r.slash = (uint32_t)(sp2 + 5);
    r.status = (r.status & ~SHL_F_VERSION_MASK) | ((w1 >> 24) == '1' ? 1u << SHL_F_VERSION_SHIFT : 0u) | SHL_F_FIRST_LINE_DONE;
    get_or_head = (sp1 == 3 && (f4 & 0xffffffu) == 0x544547u) || (sp1 == 4 && f4 == 0x44414548u);
    return true;
}

__device__ __noinline__ int find_crlf_scalar(const uint8_t* m, int from, int L) {
    for (int i = from; i + 1 < L; i++) if (ldb(m + i) == '\r' && ldb(m + i + 1) == '\n') return i;
    return -1;
}
__device__ __noinline__ int find_term_scalar(const uint8_t* m, int from, int L) {
    for (int i = from; i + 3 < L; i++)
        if (ldb(m + i) == '\r' && ldb(m + i + 1) == '\n' && ldb(m + i + 2) == '\r' && ldb(m + i + 3) == '\n') return i;
    return -1;
}

// ----------------------------------------------------------------------------------- records
// One header line [ls, end) with its separator at sep: strip the value (HttpParser.py:329; c0 / c1
// are the value's first and last byte, already loaded), store the record, remember the line if
// its key has the length of one of the two names the header state cares about (:194-197).
struct NameCand { uint32_t n14, n17, lo14, hi14, lo17, hi17; };       // first record of each length, as stored
__device__ __forceinline__ void put_record(const uint8_t* m, shl_hdr_span* out, const uint32_t k, const int ls, const int sep,
                                           const int end, const uint32_t c0, const uint32_t c1, NameCand& nc) {
    int vs = sep + 2, ve = end;
    if (is_ws(c0) || is_ws(c1)) {
        while (vs < ve && is_ws(ldb(m + vs))) vs++;
        while (ve > vs && is_ws(ldb(m + ve - 1))) ve--;
    }
    const int klen = sep - ls;
    const uint32_t lo = (uint32_t)ls | ((uint32_t)klen << 16), hi = (uint32_t)vs | ((uint32_t)(ve - vs) << 16);
    SHL_CHK(ls >= 0 && sep >= ls && end >= sep + 2 && end < 65536 && ve >= vs);
    *reinterpret_cast<uint2*>(out + k) = make_uint2(lo, hi);
    if (klen == 14) { if (!nc.n14) { nc.lo14 = lo; nc.hi14 = hi; } nc.n14++; }
    if (klen == 17) { if (!nc.n17) { nc.lo17 = lo; nc.hi17 = hi; } nc.n17++; }
}
// The candidates' names are compared after the loop (one out-of-line compare per candidate
// instead of a byte load in every iteration).  A second candidate of the same length is rare:
// then the records just written are read back.
__device__ __noinline__ void note_names_rescan(const uint8_t* m, const shl_hdr_span* out, const uint32_t nh, WellKnown& wk) {
    wk.n_te = wk.n_cl = 0;
    for (uint32_t k = 0; k < nh; k++) {
        const shl_hdr_span rec = out[k];
        const uint32_t vs = rec.val_off, ve = (uint32_t)rec.val_off + rec.val_len;
        if (rec.key_len == 14 && span_eq_lower(m, rec.key_off, rec.key_off + 14, "content-length")) { if (!wk.n_cl) { wk.cl_vs = vs; wk.cl_ve = ve; } wk.n_cl++; }
        if (rec.key_len == 17 && span_eq_lower(m, rec.key_off, rec.key_off + 17, "transfer-encoding")) { if (!wk.n_te) { wk.te_vs = vs; wk.te_ve = ve; } wk.n_te++; }
    }
}
__device__ __forceinline__ void note_names(const uint8_t* m, const shl_hdr_span* out, const uint32_t nh, const NameCand& nc, WellKnown& wk) {
    if (nc.n14 > 1u || nc.n17 > 1u) { note_names_rescan(m, out, nh, wk); return; }
    if (nc.n14) {
        const int ko = (int)(nc.lo14 & 0xffffu);
        if (span_eq_lower(m, ko, ko + 14, "content-length")) { wk.cl_vs = nc.hi14 & 0xffffu; wk.cl_ve = (nc.hi14 & 0xffffu) + (nc.hi14 >> 16); wk.n_cl = 1; }
    }
    if (nc.n17) {
        const int ko = (int)(nc.lo17 & 0xffffu);
        if (span_eq_lower(m, ko, ko + 17, "transfer-encoding")) { wk.te_vs = nc.hi17 & 0xffffu; wk.te_ve = (nc.hi17 & 0xffffu) + (nc.hi17 >> 16); wk.n_te = 1; }
    }
}

// Group-span path: line k's separator and end are the next two events of the message's block in
// the bitmaps (they alternate in a validated block).  kB lines per round; the round after it is
// located and its byte peeks are issued before this one is written (an L2 round trip per round
// would be exposed otherwise).
#ifndef SHL_OPT_EMIT
#define SHL_OPT_EMIT 1
#endif
__device__ __forceinline__ void emit_records_bm(const uint8_t* m, const bm2* bm, const uint32_t q_idx, const int qsub, const int h0,
                                                const uint32_t nh, shl_hdr_span* out, WellKnown& wk) {
    NameCand nc; nc.n14 = nc.n17 = nc.lo14 = nc.hi14 = nc.lo17 = nc.hi17 = 0u;
    constexpr int kB = SHL_OPT_EMIT;
    if (nh) {
        EvCursor ec;
        evc_init(ec, bm, q_idx);
        struct Round { int sep[kB], end[kB]; uint32_t c0[kB], c1[kB]; };
        auto fetch = [&](const uint32_t k, Round& p) {
#pragma unroll
            for (int j = 0; j < kB; j++) {
                p.sep[j] = p.end[j] = 0; p.c0[j] = p.c1[j] = 'x';
                if (k + (uint32_t)j < nh) {
                    p.sep[j] = (int)evc_pop(ec, bm) - qsub; p.end[j] = (int)evc_pop(ec, bm) - qsub;
#ifndef SHL_EXP_NOPEEK
                    if (p.sep[j] + 2 < p.end[j]) { p.c0[j] = ldb(m + p.sep[j] + 2); p.c1[j] = ldb(m + p.end[j] - 1); }
#endif
                }
            }
        };
#ifndef SHL_OPT_DEPTH
#define SHL_OPT_DEPTH 1               // rounds of byte peeks in flight ahead of the one being written
#endif
        Round cur, nxt;
#if SHL_OPT_DEPTH >= 2
        Round nx2;
        fetch(0u, cur);
        fetch((uint32_t)kB, nxt);
#else
        fetch(0u, cur);
#endif
        int ls = h0;
        for (uint32_t k = 0; k < nh; k += kB) {
#if SHL_OPT_DEPTH >= 2
            fetch(k + 2u * kB, nx2);
#else
            fetch(k + kB, nxt);
#endif
#pragma unroll
            for (int j = 0; j < kB; j++) {
                if (k + (uint32_t)j < nh) { put_record(m, out, k + j, ls, cur.sep[j], cur.end[j], cur.c0[j], cur.c1[j], nc); ls = cur.end[j] + 2; }
            }
            cur = nxt;
#if SHL_OPT_DEPTH >= 2
            nxt = nx2;
#endif
        }
    }
    note_names(m, out, nh, nc, wk);
}
// Per-lane path: the positions come from the lane's slots of the interleaved lists.
__device__ __noinline__ void emit_records_lists(const uint8_t* m, const uint16_t* cr, const uint16_t* spp, const int h0,
                                                const uint32_t nh, shl_hdr_span* out, WellKnown& wk) {
    NameCand nc; nc.n14 = nc.n17 = nc.lo14 = nc.hi14 = nc.lo17 = nc.hi17 = 0u;
    int ls = h0;
    for (uint32_t k = 0; k < nh; k++) {
        const int sep = (int)spp[k * 32], end = (int)cr[(k + 1u) * 32];
        uint32_t c0 = 'x', c1 = 'x';
        if (sep + 2 < end) { c0 = ldb(m + sep + 2); c1 = ldb(m + end - 1); }
        put_record(m, out, k, ls, sep, end, c0, c1, nc);
        ls = end + 2;
    }
    note_names(m, out, nh, nc, wk);
}

// ----------------------------------------------------------------------------------- kernel
template <bool kPlace, bool kFull>     // kFull: 32 messages per warp (compile-time), else a.group of them
__global__ void __launch_bounds__(kWarpsPerBlock * 32, SHL_OPT_MINBLOCKS) k_route(const RouteArgs a) {
    __shared__ bm2 s_bm[kWarpsPerBlock][kBmWords];             // the group's bitmaps; the per-lane path keeps its lists here
    __shared__ uint32_t s_codes[64];
#if SHL_OPT_TMA
    __shared__ __align__(128) uint8_t s_ring[kWarpsPerBlock][kTmaSlots * kTmaRow];
    __shared__ __align__(8) uint64_t s_bar[kWarpsPerBlock][kTmaSlots];
    uint32_t tma_it = 0u;
#endif
    extern __shared__ uint32_t s_hist[];                       // kPlace: n_hosts bins (when they fit)

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const SwarK K = a.k;
#if SHL_OPT_TMA
    if (lane == 0) {
        for (int k = 0; k < kTmaSlots; k++) mbar_init(&s_bar[warp][k], 1u);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
#endif
    if (threadIdx.x < 64) s_codes[threadIdx.x] = 0u;
    const bool smem_hist = kPlace && a.place.hist != nullptr && a.n_hosts <= 4096u;
    if (smem_hist) for (uint32_t i = threadIdx.x; i < a.n_hosts; i += blockDim.x) s_hist[i] = 0u;
    __syncthreads();

    // (A block-level counting sort of the messages by scan length, so that the lanes of a warp
    // finish together, was tried here: 1.6 % faster and one more thing to get wrong.  Not kept.)
    // A warp takes groups blockIdx * warps + warp, + gridDim * warps, ...: the grid is sized to what
    // the machine holds at once (api.cu), so the per-block set-up and the tail where finished warps
    // wait for the block's slowest are paid once per resident block, not once per four groups.
    const uint32_t gsz = kFull ? (uint32_t)kGroup : a.group;
    const uint64_t n_groups = (a.n + gsz - 1u) / gsz;
    for (uint64_t grp = (uint64_t)blockIdx.x * kWarpsPerBlock + warp; grp < n_groups; grp += (uint64_t)gridDim.x * kWarpsPerBlock) {
        __syncwarp();                                          // every lane is done with the previous group's bitmaps
        const uint64_t mine = grp * gsz + lane;
        const bool have = (kFull || (uint32_t)lane < gsz) && mine < a.n;
        const uint64_t my_off = have ? a.off[mine] : 0ull;
        const int L = have ? (int)min(a.len[mine], 0x7fffffc0u) : 0;
        const uint8_t* m = a.arena + my_off;
        const bm2* bm = s_bm[warp];
        uint16_t* cr = reinterpret_cast<uint16_t*>(s_bm[warp]) + lane;            // per-lane path: lane-interleaved lists
        uint16_t* sp = cr + kCrCap * 32 + 2;
        const bool force = a.force_scalar != 0;

        // ---- scan: the whole group as one span of bytes when its messages lie in order ...
        LaneScan sc; sc.idx = sc.T = -1; sc.n_c = sc.n_s = sc.s_skip = 0u; sc.ovf = sc.bad = false;
        BmWalk gw; gw.idx = gw.T = -1; gw.n_c = gw.q_idx = gw.bad = 0u;
        uint32_t s0 = 0u;                                      // my message's position in the span
        bool coal = false;
        if (!force && __shfl_sync(SHL_FULL, (int)have, 0)) {
            const uint8_t* p0 = a.arena + __shfl_sync(SHL_FULL, my_off, 0);
            const uint8_t* A0 = p0 - (reinterpret_cast<uintptr_t>(p0) & 31);
            const uint8_t* my_end = m + L;
            const uint8_t* prev_end = reinterpret_cast<const uint8_t*>(__shfl_up_sync(SHL_FULL, reinterpret_cast<unsigned long long>(my_end), 1));
            const bool ok = !have || ((lane == 0 || m >= prev_end) && (uint64_t)(my_end - A0) <= (uint64_t)kSpanMax);
            if (__all_sync(SHL_FULL, ok)) {
                const uint32_t nbytes = __reduce_max_sync(SHL_FULL, have ? (uint32_t)(my_end - A0) : 0u);
#if SHL_OPT_TMA
                group_bitmaps_tma(A0, nbytes, s_bm[warp], lane, s_ring[warp], s_bar[warp], tma_it);
#else
                group_bitmaps(A0, nbytes, s_bm[warp], lane);
#endif
                __syncwarp();
                coal = true;
                s0 = (uint32_t)(m - A0);
                if (have) bm_walk(bm, s0, L, gw);
#if SHL_OPT_PREFETCH
                // the epilogue touches every 128-byte line of the message again (first line, URL, value
                // edges): ask for them now, while the first-line work covers the round trip
                if (have && gw.T >= 0)
                    for (int o = 0; o < gw.T + 4; o += 128) asm volatile("prefetch.global.L1 [%0];" ::"l"(m + o));
#endif
                sc.idx = gw.idx; sc.T = gw.T; sc.n_c = gw.n_c;
            }
        }
        // ... else message per lane (lanes diverge on their own events; nothing heavy happens in here)
        if (!coal && L > 0) scan_lists(m, L, cr, sp, K, sc);

        // ---- first line, header block bounds, line count
        Msg r;
        r.status = (3u << SHL_F_VERSION_SHIFT) | SHL_F_IS_REQUEST;
        r.first_line_len = r.headers_len = r.sp1 = r.sp2 = r.fl_end = r.slash = r.n_headers = 0;
        r.status_code = 0; r.content_len = 0;
        WellKnown wk; wk.n_te = wk.n_cl = 0; wk.te_vs = wk.te_ve = wk.cl_vs = wk.cl_ve = 0;
        uint32_t n_hdr = 0, hdr_err = 0, s_skip = 0;
        uint32_t hash0 = 0u;
        bool hashed = false;                                   // the digest of this message's URL is stored, hash0 is its first word
        int h0 = 0, T = -1;
        bool get_or_head = false, slow = false, hdr_state = false;   // hdr_state: the header block will be tokenised
        if (L == 0) r.status |= SHL_S_EMPTY;
        else {
            int idx = sc.idx;
            if (sc.ovf && sc.n_c == 0u) idx = find_crlf_scalar(m, 0, L);
            if (idx < 0) r.status |= SHL_S_INCOMPLETE_LINE;
            else {
                r.first_line_len = (uint32_t)(idx + 2);
                uint32_t fl_code = 0;
                // canonical request line: fields from its two ends; its URL words are fetched once,
                // for the digest and for "no space inside" together (placement is speculative here:
                // it is committed at the end, if the message turns out routable)
                bool canon = !force && !sc.ovf;
                if (canon) {
                    Msg t = r; bool goh = false;
                    canon = first_line_canon(m, idx, t, goh);
                    if (canon) {
                        const uint8_t* url = m + t.sp1 + 1u;
                        const uint32_t ulen = t.sp2 - t.sp1 - 1u;
                        if (kPlace) {
                            uint32_t hh[4];
                            canon = md5_key_t<true>(url, ulen, hh) == 0u;
                            if (canon) {       // stored now, withdrawn at the end if the message is not routable after all
                                if (a.place.md5) reinterpret_cast<uint4*>(a.place.md5)[mine] = make_uint4(hh[0], hh[1], hh[2], hh[3]);
                                hash0 = hh[0]; hashed = true;
                            }
                        } else canon = !span_has_space(url, ulen);
                        if (canon) { r = t; get_or_head = goh; }
                    }
                }
                if (!canon) {
                    Msg t = r; bool goh = false;   // the cold path may keep its copy in local memory
                    fl_code = first_line_scalar(m, idx, t, goh);
                    r = t; get_or_head = goh;
                }
                if (fl_code) r.status |= fl_code;
                else {
                    h0 = idx + 2;
                    T = sc.ovf ? find_term_scalar(m, h0, L) : sc.T;
                    if (T < 0) r.status |= SHL_S_INCOMPLETE_HEADERS;
                    else if (T + 4 > 65536) r.status |= SHL_E_LIB_HDR_TOO_LARGE;
                    else {
                        r.headers_len = (uint32_t)(T + 4 - h0);
                        hdr_state = true;
                        slow = sc.ovf || force || is_ws(ldb(m + h0)) || is_ws(ldb(m + T - 1));   // strip() would move the block's ends
                        if (slow) {                // count (and classify) now, emit after the allocation
                            WellKnown w2 = wk; uint32_t n2 = 0;
                            hdr_err = headers_scalar(m, h0, T + 4, nullptr, 1, 0u, &n2, &w2);
                            wk = w2; n_hdr = n2;
                        } else {
                            // lines before the first one that does not hold exactly one ": " are the
                            // records (the scan kept a flag; only a block that tripped it is walked
                            // again to find the offender)
                            if (coal) {
                                n_hdr = gw.n_c - 1u;
                                if (gw.bad) {
                                    bool no_sep = false;
                                    n_hdr = bm_offender(bm, s0, gw, &no_sep);
                                    hdr_err = no_sep ? SHL_E_HDR_NO_SEP : SHL_E_HDR_MULTI_SEP;
                                }
                            } else if (sc.bad) n_hdr = count_lines(cr, sp, sc, &hdr_err, &s_skip);
                            else { n_hdr = sc.n_c - 1u; s_skip = sc.s_skip; }
                        }
                    }
                }
            }
        }

        // ---- one atomic allocates the group's header records
        uint32_t nh = have ? n_hdr : 0u;
        uint32_t inc = nh;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(SHL_FULL, inc, d); if (lane >= d) inc += t; }
        const uint32_t total = __shfl_sync(SHL_FULL, inc, 31);
        unsigned long long base = 0;
        if (lane == 0 && total) base = atomicAdd(reinterpret_cast<unsigned long long*>(a.out.hdr_used), (unsigned long long)total);
        base = __shfl_sync(SHL_FULL, base, 0);
        const unsigned long long my_base = base + (inc - nh);
        const bool fits = my_base + nh <= a.out.hdr_cap;
        if (!fits) {                                           // record arena full: library limit
            r.status = (r.status & ~0xffu) | SHL_E_LIB_RECORDS_FULL;
            nh = 0; hdr_state = false;
        }

        // ---- records: strip the value (HttpParser.py:329), note the two keys that matter
        if (hdr_state) {
            shl_hdr_span* out = a.out.hdr + my_base;
            if (slow) {
                WellKnown w2; uint32_t n2 = 0;
                w2.n_te = w2.n_cl = 0; w2.te_vs = w2.te_ve = w2.cl_vs = w2.cl_ve = 0;
                headers_scalar(m, h0, T + 4, out, 1, nh, &n2, &w2);
            } else if (coal) emit_records_bm(m, bm, gw.q_idx, (int)s0 + 1, h0, nh, out, wk);
            else emit_records_lists(m, cr, sp + s_skip * 32, h0, nh, out, wk);
            r.n_headers = nh;
            if (hdr_err) r.status |= hdr_err;
            else finalize_headers(m, L, T + 4, get_or_head, wk, r.status, r.content_len);
        }

        // ---- optional: the header names lower-cased (key.lower(), HttpParser.py:330), packed per message
        if (a.out.lkeys != nullptr) {
            const shl_hdr_span* recs = a.out.hdr + my_base;
            uint32_t kb = 0;
            for (uint32_t k = 0; k < r.n_headers; k++) kb += recs[k].key_len;
            uint32_t kinc = kb;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(SHL_FULL, kinc, d); if (lane >= d) kinc += t; }
            const uint32_t ktot = __shfl_sync(SHL_FULL, kinc, 31);
            unsigned long long kbase = 0;
            if (lane == 0 && ktot) kbase = atomicAdd(reinterpret_cast<unsigned long long*>(a.out.lkeys_used), (unsigned long long)ktot);
            kbase = __shfl_sync(SHL_FULL, kbase, 0) + (kinc - kb);
            if (have) {
                a.out.lkey_base[mine] = (uint32_t)kbase;
                if (kbase + kb <= a.out.lkeys_cap) {
                    uint8_t* dst = a.out.lkeys + kbase;
                    for (uint32_t k = 0; k < r.n_headers; k++) {
                        const shl_hdr_span rec = recs[k];
                        for (uint32_t j = 0; j < rec.key_len; j++) dst[j] = (uint8_t)lower(ldb(m + rec.key_off + j));
                        dst += rec.key_len;
                    }
                } else r.status = (r.status & ~0xffu) | SHL_E_LIB_RECORDS_FULL;
            }
        }
        if ((r.status & (SHL_F_IS_REQUEST | SHL_F_FIRST_LINE_DONE)) == (SHL_F_IS_REQUEST | SHL_F_FIRST_LINE_DONE) &&
            r.sp2 - r.sp1 - 1u > (uint32_t)SHL_MAX_KEY_LEN) r.status |= SHL_F_KEY_TOO_LONG;      // pylibmc raises before hashing (Server.py:335)

        // ---- per-message fixed fields: lane j == message j, so these are coalesced SoA stores
        if (have && a.crec == nullptr) {
            a.out.status[mine] = r.status;
            a.out.first_line_len[mine] = r.first_line_len;
            a.out.headers_len[mine] = r.headers_len;
            a.out.sp1[mine] = r.sp1; a.out.sp2[mine] = r.sp2;
            a.out.fl_end[mine] = r.fl_end; a.out.slash[mine] = r.slash;
            a.out.status_code[mine] = r.status_code;
            a.out.n_headers[mine] = r.n_headers;
            a.out.hdr_base[mine] = (uint32_t)my_base + a.hdr_base_bias;
            a.out.content_len[mine] = r.content_len;
        }
        if (have && a.crec != nullptr && (r.first_line_len > 65535u || (r.status & (SHL_F_STATUS_WIDE)) ||
                                          (!(r.status & SHL_F_IS_REQUEST) && (uint32_t)r.status_code > 65535u))) {
            // the compact record keeps these in 16 bits: a first line beyond 64 KiB is a library limit
            // there, a status code beyond 16 bits is handed to the host like any other wide one
            if (r.first_line_len > 65535u) { r.status = (r.status & ~0xffu) | SHL_E_LIB_HDR_TOO_LARGE; r.first_line_len = 0; }
            else r.status |= SHL_F_STATUS_WIDE;
        }
        if (have) atomicAdd(&s_codes[r.status & 63u], 1u);

        // ---- fused key + MD5 + ketama: the URL is the key, verbatim  (Server.py:325-335)
        uint32_t rec_node = 0xffffffffu;
        if (kPlace && have) {
            const bool routable = (r.status & SHL_F_MESSAGE_COMPLETE) && (r.status & SHL_F_IS_REQUEST) &&
                                  (r.status & SHL_F_FIRST_LINE_DONE) && (r.status & 0xffu) == SHL_S_OK &&
                                  !(r.status & SHL_F_KEY_TOO_LONG);
            uint32_t node = 0xffffffffu;
            if (routable) {
                if (!hashed) {
                    uint32_t hh[4];
                    md5_key(m + r.sp1 + 1u, r.sp2 - r.sp1 - 1u, hh);
                    if (a.place.md5) reinterpret_cast<uint4*>(a.place.md5)[mine] = make_uint4(hh[0], hh[1], hh[2], hh[3]);
                    hash0 = hh[0];
                }
                node = ring_lookup(a.ring_values, a.ring_nodes, SHL_OPT_RINGDIR ? a.ring_dir : nullptr, a.n_points, a.n_hosts, hash0);
                if (a.place.hist) {
                    if (smem_hist) atomicAdd(&s_hist[node], 1u);
                    else atomicAdd(reinterpret_cast<unsigned long long*>(a.place.hist) + node, 1ull);
                }
            } else {
                hash0 = 0u;
                if (a.place.md5) reinterpret_cast<uint4*>(a.place.md5)[mine] = make_uint4(0u, 0u, 0u, 0u);
            }
            if (a.place.hash) a.place.hash[mine] = hash0;
            if (a.place.node) a.place.node[mine] = node;
            rec_node = node;
        }
        // ---- compact output: ONE 32-byte record per message (two 16-byte stores, a whole sector)
        if (have && a.crec != nullptr) {
            const uint32_t tail = (r.status & SHL_F_IS_REQUEST) ? min(rec_node, 0xffffu) : ((uint32_t)r.status_code & 0xffffu);
            uint4 lo, hi;
            lo.x = r.status; lo.y = (uint32_t)my_base + a.hdr_base_bias;
            lo.z = (uint32_t)(uint64_t)r.content_len; lo.w = (uint32_t)((uint64_t)r.content_len >> 32);
            hi.x = r.first_line_len | (r.headers_len << 16); hi.y = r.sp1 | (r.sp2 << 16);
            hi.z = r.fl_end | (r.slash << 16); hi.w = r.n_headers | (tail << 16);
            uint4* dst = reinterpret_cast<uint4*>(a.crec + mine);
            dst[0] = lo; dst[1] = hi;
        }
    }
    __syncthreads();
    if (a.code_hist && threadIdx.x < 64 && s_codes[threadIdx.x])
        atomicAdd(reinterpret_cast<unsigned long long*>(a.code_hist) + threadIdx.x, (unsigned long long)s_codes[threadIdx.x]);
    if (smem_hist)
        for (uint32_t i = threadIdx.x; i < a.n_hosts; i += blockDim.x)
            if (s_hist[i]) atomicAdd(reinterpret_cast<unsigned long long*>(a.place.hist) + i, (unsigned long long)s_hist[i]);
}

}  // namespace shl
