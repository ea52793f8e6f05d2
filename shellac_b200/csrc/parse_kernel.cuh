// parse_kernel.cuh — K1: batched HTTP/1.1 tokeniser (+ optional fused key/MD5/ketama placement).
//
// Replaces HttpParser.parse / _parse_first_line / _parse_headers and the Content-Length framing
// decision (/root/reference/src/python/shellac/server/HttpParser.py:140-223, 304-337), driven
// over a whole message as the reference's own test loop does (test/HttpParserTests.py:36-39).
//
// Mapping (round 1, second design; the first one, a warp per message with ballot/shuffle structure
// extraction, measured 960 warp-instructions per request and was issue-bound at 6.6 % of the HBM
// roofline, see profiles/r01/): one warp owns a GROUP of 32 consecutive messages, LANE j owns
// MESSAGE j.
//   * each lane streams its own message in 16-byte vectors (aligned down, lead bytes masked,
//     next vector prefetched) and classifies the bytes with SWAR compares (CR, LF, ':', ' ');
//     "LF preceded by CR" and "' ' preceded by ':'" become 16-bit position masks, with the
//     previous vector's last-byte flags carried in two registers;
//   * a short per-lane loop walks the CRLF bits: the first one ends the request/status line, an
//     empty line ends the header block, every other one closes a header line whose ': ' count and
//     first position were accumulated from the mask; the line's key/value spans go to a
//     lane-interleaved (bank-conflict-free) shared-memory staging area;
//   * all per-message structure is lane-private: no ballots, shuffles or shared tables in the scan.
//     Lanes re-converge after the scan for the first-line evaluation, the Content-Length /
//     Transfer-Encoding decision, one atomic record allocation per group, the copy-out, the
//     struct-of-arrays stores (lane j == message j, so they are coalesced) and the fused
//     MD5 + continuum search + shared-memory histogram.
// Anything off the canonical shape (responses' first line, whitespace around the header block,
// more header lines than a lane's staging slots) takes a per-lane scalar path with the same
// semantics.
#pragma once
#include "shl_device.cuh"

namespace shl {

constexpr int kWarpsPerBlock = 4;
constexpr int kGroup = 32;            // messages per warp (one per lane)
constexpr int kLaneCap = 20;          // staged header records per lane (5 KB per warp)

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
    // fused placement (route); ring_values == nullptr -> parse only
    const uint32_t* ring_values; const uint32_t* ring_nodes; uint32_t n_points, n_hosts;
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

// unaligned little-endian 32-bit load (may touch up to 3 bytes past p+4, inside the arena padding)
__device__ __forceinline__ uint32_t ld32u(const uint8_t* p) {
    const uint32_t* a = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)3);
    const int sh = (int)(reinterpret_cast<uintptr_t>(p) & 3) * 8;
    const uint32_t lo = __ldg(a);
    const uint32_t hi = sh ? __ldg(a + 1) : 0u;
    return __funnelshift_r(lo, hi, sh);
}

// ----------------------------------------------------------------------------------- one message, one lane
// Per-lane scan state (all of it lives in registers once the steps below are inlined).
struct Scan {
    int p;                       // message position of byte 0 of the vector being scanned
    int idx, sp1, sp2, h0, T;    // first CRLF, first two spaces, header start, CRLFCRLF position
    int line_start, sep_first;
    uint32_t sep_cnt, n_ev, first_bad, bad_cnt, prev_cr, prev_co;
};

// One 16-byte vector.  Events (header line ends) are pushed to dst[k * stride] as
// {key_off = position of the line's first ": ", key_len = number of ": " (0, 1, 2+),
//  val_off = position of the line's CR}; they are turned into records after the scan.
__device__ __forceinline__ void scan_vec(Scan& s, const uint4 v, const int L, shl_hdr_span* dst, const int stride,
                                         const uint32_t cap, const SwarK& K) {
    const uint32_t cr0 = eqf(v.x, K.cr, K), cr1 = eqf(v.y, K.cr, K), cr2 = eqf(v.z, K.cr, K), cr3 = eqf(v.w, K.cr, K);
    const uint32_t co0 = eqf(v.x, K.col, K), co1 = eqf(v.y, K.col, K), co2 = eqf(v.z, K.col, K), co3 = eqf(v.w, K.col, K);
    // "LF preceded by CR" / "' ' preceded by ':'": the predecessor flags shifted up one byte
    uint32_t c16 = mask16(eqf_and(v.x, K.lf, __funnelshift_l(s.prev_cr, cr0, 8), K), eqf_and(v.y, K.lf, __funnelshift_l(cr0, cr1, 8), K),
                          eqf_and(v.z, K.lf, __funnelshift_l(cr1, cr2, 8), K), eqf_and(v.w, K.lf, __funnelshift_l(cr2, cr3, 8), K));
    uint32_t s16 = mask16(eqf_and(v.x, K.sp, __funnelshift_l(s.prev_co, co0, 8), K), eqf_and(v.y, K.sp, __funnelshift_l(co0, co1, 8), K),
                          eqf_and(v.z, K.sp, __funnelshift_l(co1, co2, 8), K), eqf_and(v.w, K.sp, __funnelshift_l(co2, co3, 8), K));
    s.prev_cr = cr3; s.prev_co = co3;
    const int p = s.p;
    if (p < 1 || p + 16 > L) { const uint32_t vm = rng16(1 - p, L - p); c16 &= vm; s16 &= vm; }
    if (s.idx < 0) {                               // still on the first line: note its first two spaces
        uint32_t z = mask16(eqf(v.x, K.sp, K), eqf(v.y, K.sp, K), eqf(v.z, K.sp, K), eqf(v.w, K.sp, K));
        if (p < 0 || p + 16 > L) z &= rng16(-p, L - p);
        if (c16) z &= below(__ffs(c16) - 1);
        if (z && s.sp1 < 0) { s.sp1 = p + __ffs(z) - 1; z &= z - 1u; }
        if (z && s.sp1 >= 0 && s.sp2 < 0) s.sp2 = p + __ffs(z) - 1;
    }
    int lo = 0;
    while (c16) {
        const int k = __ffs(c16) - 1;              // the LF; its CR sits at k - 1
        c16 &= c16 - 1u;
        const int cr_pos = p + k - 1;
        const uint32_t seg = s16 & below(k) & ~below(lo);
        lo = k + 1;
        if (s.idx < 0) { s.idx = cr_pos; s.h0 = s.line_start = cr_pos + 2; continue; }
        if (seg) { if (!s.sep_cnt) s.sep_first = p + __ffs(seg) - 2; s.sep_cnt += (uint32_t)__popc(seg); }
        if (cr_pos == s.line_start && s.line_start > s.h0) { s.T = s.line_start - 2; return; }   // empty line: CRLFCRLF at T
        if (s.sep_cnt != 1u && s.n_ev < s.first_bad) { s.first_bad = s.n_ev; s.bad_cnt = s.sep_cnt; }
        if (s.n_ev < cap) {
            shl_hdr_span ev;
            ev.key_off = (uint16_t)s.sep_first; ev.key_len = (uint16_t)min(s.sep_cnt, 2u);
            ev.val_off = (uint16_t)cr_pos; ev.val_len = 0;
            dst[(size_t)s.n_ev * stride] = ev;
        }
        s.n_ev++;
        s.line_start = cr_pos + 2; s.sep_cnt = 0;
    }
    if (s.idx >= 0) {
        const uint32_t seg = s16 & ~below(lo);
        if (seg) { if (!s.sep_cnt) s.sep_first = p + __ffs(seg) - 2; s.sep_cnt += (uint32_t)__popc(seg); }
    }
}

// Records go to dst[k * stride] for k < cap (shared staging: stride 32; global: stride 1);
// r.n_headers counts every tokenised line even when cap is smaller (the caller then re-runs the
// message with exactly that much room in the global record arena).
__device__ __forceinline__ void scan_message(const uint8_t* m, const int L, shl_hdr_span* dst, const int stride,
                                             const uint32_t cap, const SwarK& K, const bool force_scalar, Msg& r) {
    r.status = (3u << SHL_F_VERSION_SHIFT) | SHL_F_IS_REQUEST;
    r.first_line_len = r.headers_len = r.sp1 = r.sp2 = r.fl_end = r.slash = r.n_headers = 0;
    r.status_code = 0; r.content_len = 0;
    if (L == 0) { r.status |= SHL_S_EMPTY; return; }

    // Each lane streams its own message one 32-byte sector (two vectors) per iteration, the next
    // sector already in flight; the first sector is aligned down and its lead bytes are masked.
    const int lead = (int)(reinterpret_cast<uintptr_t>(m) & 31);
    const uint4* vp = reinterpret_cast<const uint4*>(m - lead);
    const int nsec = (lead + L + 31) >> 5;
    Scan s;
    s.p = -lead; s.idx = s.sp1 = s.sp2 = -1; s.h0 = 0; s.T = -1; s.line_start = 0; s.sep_first = 0;
    s.sep_cnt = 0; s.n_ev = 0; s.first_bad = 0xffffffffu; s.bad_cnt = 0; s.prev_cr = 0; s.prev_co = 0;
    uint4 a0 = __ldg(vp), a1 = __ldg(vp + 1);
    for (int i = 0; i < nsec; i++) {
        uint4 b0 = make_uint4(0u, 0u, 0u, 0u), b1 = b0;
        if (i + 1 < nsec) { b0 = __ldg(vp + 2 * i + 2); b1 = __ldg(vp + 2 * i + 3); }
        scan_vec(s, a0, L, dst, stride, cap, K);
        if (s.T >= 0) break;
        s.p += 16;
        if (s.p < L) {
            scan_vec(s, a1, L, dst, stride, cap, K);
            if (s.T >= 0) break;
        }
        s.p += 16;
        a0 = b0; a1 = b1;
    }

    // ---- lanes are converged again from here on
    const int idx = s.idx, sp1 = s.sp1, sp2 = s.sp2, h0 = s.h0, T = s.T;
    if (idx < 0) { r.status |= SHL_S_INCOMPLETE_LINE; return; }
    r.first_line_len = (uint32_t)(idx + 2);
    bool get_or_head = false;
    bool fast = !force_scalar && sp1 >= 0 && sp2 > sp1 && sp2 < idx && idx - sp2 - 1 == 8;
    if (fast) {                                    // canonical request line: METHOD SP URL SP "HTTP/1.x" CRLF
        const uint32_t w0 = ld32u(m + sp2 + 1), w1 = ld32u(m + sp2 + 5), f4 = ld32u(m);
        fast = w0 == 0x50545448u && (w1 == 0x312E312Fu || w1 == 0x302E312Fu) && !((f4 & 0xffffu) == 0x5448u && sp1 >= 2);
        if (fast) {
            r.sp1 = (uint32_t)sp1; r.sp2 = (uint32_t)sp2; r.fl_end = (uint32_t)idx; r.slash = (uint32_t)(sp2 + 5);
            r.status = (r.status & ~SHL_F_VERSION_MASK) | ((w1 >> 24) == '1' ? 1u << SHL_F_VERSION_SHIFT : 0u) | SHL_F_FIRST_LINE_DONE;
            get_or_head = (sp1 == 3 && (f4 & 0xffffffu) == 0x544547u) || (sp1 == 4 && f4 == 0x44414548u);
        }
    }
    if (!fast) {
        Msg t = r;                                 // the cold path may keep its copy in local memory
        bool goh = false;
        const uint32_t fl_code = first_line_scalar(m, idx, t, goh);
        r = t; get_or_head = goh;
        if (fl_code) { r.status |= fl_code; return; }
    }
    if (T < 0) { r.status |= SHL_S_INCOMPLETE_HEADERS; return; }
    if (T + 4 > 65536) { r.status |= SHL_E_LIB_HDR_TOO_LARGE; return; }
    r.headers_len = (uint32_t)(T + 4 - h0);
    WellKnown wk; wk.n_te = wk.n_cl = 0; wk.te_vs = wk.te_ve = wk.cl_vs = wk.cl_ve = 0;
    uint32_t n_hdr, hdr_err = 0;
    if (force_scalar || is_ws(ldb(m + h0)) || is_ws(ldb(m + T - 1))) {    // strip() would move the block's ends
        WellKnown w2 = wk; uint32_t n2 = 0;
        hdr_err = headers_scalar(m, h0, T + 4, dst, stride, cap, &n2, &w2);
        wk = w2; n_hdr = n2;
    } else {
        // turn the staged events into records, in place: strip the value, note the two keys that matter
        n_hdr = min(s.n_ev, s.first_bad);
        if (s.first_bad < s.n_ev) hdr_err = s.bad_cnt ? (uint32_t)SHL_E_HDR_MULTI_SEP : (uint32_t)SHL_E_HDR_NO_SEP;
        const uint32_t n_conv = min(n_hdr, cap);
        int ls = h0;
        for (uint32_t k = 0; k < n_conv; k++) {
            const shl_hdr_span ev = dst[(size_t)k * stride];
            const int sep = ev.key_off, cr = ev.val_off;
            int vs = sep + 2, ve = cr;
            const uint32_t c0 = vs < ve ? ldb(m + vs) : 'x', c1 = ve > vs ? ldb(m + ve - 1) : 'x';
            if (is_ws(c0) || is_ws(c1)) {
                while (vs < ve && is_ws(ldb(m + vs))) vs++;
                while (ve > vs && is_ws(ldb(m + ve - 1))) ve--;
            }
            const int klen = sep - ls;
            shl_hdr_span rec;
            rec.key_off = (uint16_t)ls; rec.key_len = (uint16_t)klen;
            rec.val_off = (uint16_t)vs; rec.val_len = (uint16_t)(ve - vs);
            dst[(size_t)k * stride] = rec;
            if (klen == 14) { if (span_eq_lower(m, ls, sep, "content-length")) { if (!wk.n_cl) { wk.cl_vs = vs; wk.cl_ve = ve; } wk.n_cl++; } }
            else if (klen == 17) { if (span_eq_lower(m, ls, sep, "transfer-encoding")) { if (!wk.n_te) { wk.te_vs = vs; wk.te_ve = ve; } wk.n_te++; } }
            ls = cr + 2;
        }
    }
    r.n_headers = n_hdr;
    if (hdr_err) { r.status |= hdr_err; return; }
    if (n_hdr > cap) return;                       // not all lines were staged: the caller re-runs this message
    finalize_headers(m, L, T + 4, get_or_head, wk, r.status, r.content_len);
}

// ----------------------------------------------------------------------------------- kernel
template <bool kPlace>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, 8) k_route(const RouteArgs a) {
    __shared__ shl_hdr_span stage[kWarpsPerBlock][kLaneCap * 32];
    __shared__ uint32_t s_codes[64];
    extern __shared__ uint32_t s_hist[];                       // kPlace: n_hosts bins (when they fit)

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const SwarK K = a.k;
    if (threadIdx.x < 64) s_codes[threadIdx.x] = 0u;
    const bool smem_hist = kPlace && a.place.hist != nullptr && a.n_hosts <= 4096u;
    if (smem_hist) for (uint32_t i = threadIdx.x; i < a.n_hosts; i += blockDim.x) s_hist[i] = 0u;
    __syncthreads();

    const uint64_t first = ((uint64_t)blockIdx.x * kWarpsPerBlock + warp) * kGroup;
    if (first < a.n) {
        const uint64_t mine = first + lane;
        const bool have = mine < a.n;
        const uint64_t my_off = have ? a.off[mine] : 0ull;
        const int L = have ? (int)min(a.len[mine], 0x7fffffc0u) : 0;
        const uint8_t* m = a.arena + my_off;
        shl_hdr_span* my_stage = &stage[warp][lane];
        Msg r;
        scan_message(m, L, my_stage, 32, (uint32_t)kLaneCap, K, a.force_scalar != 0, r);
        const bool ovf = r.n_headers > (uint32_t)kLaneCap;

        // ---- one atomic allocates the group's header records; lanes copy their own out
        uint32_t nh = have ? r.n_headers : 0u;
        uint32_t inc = nh;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(SHL_FULL, inc, d); if (lane >= d) inc += t; }
        const uint32_t total = __shfl_sync(SHL_FULL, inc, 31);
        unsigned long long base = 0;
        if (lane == 0 && total) base = atomicAdd(reinterpret_cast<unsigned long long*>(a.out.hdr_used), (unsigned long long)total);
        base = __shfl_sync(SHL_FULL, base, 0);
        const unsigned long long my_base = base + (inc - nh);
        if (my_base + nh > a.out.hdr_cap) {                    // record arena full: library limit
            r.status = (r.status & ~0xffu & ~SHL_F_MESSAGE_COMPLETE) | SHL_E_LIB_RECORDS_FULL;
            nh = 0; r.n_headers = 0;
        }
        if (!ovf) {
            uint2* g = reinterpret_cast<uint2*>(a.out.hdr + my_base);
            const uint2* sgl = reinterpret_cast<const uint2*>(my_stage);
#pragma unroll 4
            for (uint32_t k = 0; k < nh; k++) g[k] = sgl[k * 32];
        } else if (nh) {                                       // more lines than staging slots: redo, straight to global
            Msg r2;
            scan_message(m, L, a.out.hdr + my_base, 1, nh, K, a.force_scalar != 0, r2);
            r = r2;
        }

        // ---- per-message fixed fields: lane j == message j, so these are coalesced SoA stores
        if (have) {
            a.out.status[mine] = r.status;
            a.out.first_line_len[mine] = r.first_line_len;
            a.out.headers_len[mine] = r.headers_len;
            a.out.sp1[mine] = r.sp1; a.out.sp2[mine] = r.sp2;
            a.out.fl_end[mine] = r.fl_end; a.out.slash[mine] = r.slash;
            a.out.status_code[mine] = r.status_code;
            a.out.n_headers[mine] = r.n_headers;
            a.out.hdr_base[mine] = (uint32_t)my_base;
            a.out.content_len[mine] = r.content_len;
            atomicAdd(&s_codes[r.status & 63u], 1u);
        }

        // ---- fused key + MD5 + ketama: the URL is the key, verbatim  (Server.py:325-335)
        if (kPlace && have) {
            const bool routable = (r.status & SHL_F_MESSAGE_COMPLETE) && (r.status & SHL_F_IS_REQUEST) &&
                                  (r.status & SHL_F_FIRST_LINE_DONE) && (r.status & 0xffu) == SHL_S_OK;
            uint32_t h[4] = {0u, 0u, 0u, 0u}, node = 0xffffffffu;
            if (routable) {
                md5_key(m + r.sp1 + 1u, r.sp2 - r.sp1 - 1u, h);
                node = ring_lookup(a.ring_values, a.ring_nodes, a.n_points, a.n_hosts, h[0]);
                if (a.place.hist) {
                    if (smem_hist) atomicAdd(&s_hist[node], 1u);
                    else atomicAdd(reinterpret_cast<unsigned long long*>(a.place.hist) + node, 1ull);
                }
            }
            if (a.place.md5) reinterpret_cast<uint4*>(a.place.md5)[mine] = make_uint4(h[0], h[1], h[2], h[3]);
            if (a.place.hash) a.place.hash[mine] = h[0];
            if (a.place.node) a.place.node[mine] = node;
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
