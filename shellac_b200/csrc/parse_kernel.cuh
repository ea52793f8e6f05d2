// parse_kernel.cuh — K1: batched HTTP/1.1 tokeniser (+ optional fused key/MD5/ketama placement).
//
// Replaces HttpParser.parse / _parse_first_line / _parse_headers and the Content-Length framing
// decision (/root/reference/src/python/shellac/server/HttpParser.py:140-223, 304-337), driven
// over a whole message as the reference's own test loop does (test/HttpParserTests.py:36-39).
//
// Mapping: one warp owns a GROUP of 32 consecutive messages and walks them one at a time.
//   * message bytes are streamed in 512-byte chunks, one 16-byte vector per lane (coalesced,
//     alignment-agnostic: the first vector is aligned down and the lead bytes are masked);
//   * each lane classifies its 16 bytes with SWAR compares (CR, LF, ':', ' ') and folds them into
//     16-bit position masks for "CRLF here", "': ' here", "' ' here", "CRLFCRLF here";
//   * warp ballots locate the first line end, the two first-line spaces and the header
//     terminator; line ends and ': ' separators are scattered by rank into a 33-entry per-warp
//     shared-memory table, after which LANE j tokenises LINE j (key span, stripped value span);
//   * header records of the whole group are staged in shared memory and copied out with one
//     atomically-allocated, coalesced write; the per-message fixed fields are written lane-per-
//     message as struct-of-arrays.
// Anything off the canonical shape (responses' first line, whitespace around the header block,
// >32 lines in one chunk, ...) takes a warp-uniform scalar path with the same semantics.
#pragma once
#include "shl_device.cuh"

namespace shl {

constexpr int kWarpsPerBlock = 8;
constexpr int kGroup = 32;            // messages per warp
constexpr int kStageCap = 640;        // staged header records per warp (5 KB)
constexpr uint32_t kNoPos = 0x7fffffffu;

struct WarpScratch {
    uint32_t ends[34];
    uint32_t cnt[34];
    uint32_t first[34];
    uint32_t pad[2];
};

struct Msg {                          // warp-uniform result of one message
    uint32_t status, first_line_len, headers_len, sp1, sp2, fl_end, slash, n_headers;
    int32_t status_code;
    int64_t content_len;
};

struct RouteArgs {
    const uint8_t* arena; const uint64_t* off; const uint32_t* len; uint64_t n;
    shl_parse_out out; uint64_t* code_hist;
    int force_scalar;                 // test hook: take the scalar path for every message
    // fused placement (route); ring_values == nullptr -> parse only
    const uint32_t* ring_values; const uint32_t* ring_nodes; uint32_t n_points, n_hosts;
    shl_place_out place;
};

// ----------------------------------------------------------------------------------- SWAR
struct SwarK { uint32_t m7, hi, cr, lf, col, sp; };
__device__ __forceinline__ SwarK swar_consts() {
    SwarK k;   // opaque to the optimiser so LOP3 takes them as register operands (one LOP3 per term)
    asm("mov.u32 %0, 0x7f7f7f7f;" : "=r"(k.m7));
    asm("mov.u32 %0, 0x80808080;" : "=r"(k.hi));
    asm("mov.u32 %0, 0x0d0d0d0d;" : "=r"(k.cr));
    asm("mov.u32 %0, 0x0a0a0a0a;" : "=r"(k.lf));
    asm("mov.u32 %0, 0x3a3a3a3a;" : "=r"(k.col));
    asm("mov.u32 %0, 0x20202020;" : "=r"(k.sp));
    return k;
}
// 0x80 in every byte of x that equals the (7-bit) byte replicated in c
__device__ __forceinline__ uint32_t eqf(uint32_t x, uint32_t c, const SwarK& k) {
    uint32_t t = ((x ^ c) & k.m7) + k.m7;
    return ~(t | x) & k.hi;
}
// gather the four 0x80 flags of f into bits 0..3
__device__ __forceinline__ uint32_t nib(uint32_t f) { return __umulhi(f, 0x02040810u) & 0xfu; }
__device__ __forceinline__ uint32_t mask16(uint32_t f0, uint32_t f1, uint32_t f2, uint32_t f3) {
    return nib(f0) | (nib(f1) << 4) | (nib(f2) << 8) | (nib(f3) << 12);
}
__device__ __forceinline__ int clamp16(int v) { return min(max(v, 0), 16); }
__device__ __forceinline__ uint32_t below(int k) { return (1u << k) - 1u; }            // k in [0,16]
__device__ __forceinline__ uint32_t rng16(int lo, int hi) { return below(clamp16(hi)) & ~below(clamp16(lo)); }

__device__ __forceinline__ uint4 load_vec(const uint8_t* abase, int c, int lane, int qL) {
    const int q = c * 512 + lane * 16;
    if (q < qL) return __ldg(reinterpret_cast<const uint4*>(abase + q));
    return make_uint4(0u, 0u, 0u, 0u);
}

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
// Warp-uniform; lane 0 writes the records.
__device__ __noinline__ uint32_t headers_scalar(const uint8_t* m, int h0, int h1, shl_hdr_span* dst, uint32_t cap,
                                                int lane, uint32_t& n_out, WellKnown& wk) {
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
        if (nsep != 1) { n_out = n; return nsep == 0 ? SHL_E_HDR_NO_SEP : SHL_E_HDR_MULTI_SEP; }
        int vs = sep + 2, ve = lend;
        while (vs < ve && is_ws(ldb(m + vs))) vs++;
        while (ve > vs && is_ws(ldb(m + ve - 1))) ve--;
        if (n < cap && lane == 0) {
            shl_hdr_span rec;
            rec.key_off = (uint16_t)ls; rec.key_len = (uint16_t)(sep - ls);
            rec.val_off = (uint16_t)vs; rec.val_len = (uint16_t)(ve - vs);
            dst[n] = rec;
        }
        n++;
        if (span_eq_lower(m, ls, sep, "transfer-encoding")) { if (!wk.n_te) { wk.te_vs = vs; wk.te_ve = ve; } wk.n_te++; }
        if (span_eq_lower(m, ls, sep, "content-length")) { if (!wk.n_cl) { wk.cl_vs = vs; wk.cl_ve = ve; } wk.n_cl++; }
        if (le < 0) break;
        ls = le + 2;
    }
    n_out = n;
    return 0;
}

// Header-state epilogue (HttpParser.py:187-200) + the Content-Length body call (:206-223).
__device__ __forceinline__ void finalize_headers(const uint8_t* m, int L, int body_off, bool get_or_head,
                                                 const WellKnown& wk, Msg& r) {
    r.status |= SHL_F_HEADERS_COMPLETE | SHL_F_ON_BODY;
    bool complete = (r.status & SHL_F_IS_REQUEST) && get_or_head;
    const bool chunked = wk.n_te == 1 && span_eq(m, wk.te_vs, wk.te_ve, "chunked");
    if (chunked) r.status |= SHL_F_CHUNKED;
    else {
        if (wk.n_cl > 1) { r.status |= SHL_E_CLEN_LIST | (complete ? SHL_F_MESSAGE_COMPLETE : 0u); return; }
        if (wk.n_cl == 1) {
            int64_t v = 0;
            const int k = py_int10(m + wk.cl_vs, (int64_t)wk.cl_ve - wk.cl_vs, &v);
            if (k == 0) { r.status |= SHL_E_CLEN_INT | (complete ? SHL_F_MESSAGE_COMPLETE : 0u); return; }
            if (k == 2) { r.status |= SHL_F_CLEN_WIDE; v = ldb(m + wk.cl_vs) == '-' ? INT64_MIN : INT64_MAX; }
            r.content_len = v;
        }
        if (r.content_len == 0 && !(r.status & SHL_F_CLEN_WIDE)) { r.status &= ~SHL_F_ON_BODY; complete = true; }
    }
    if (complete) { r.status |= SHL_F_MESSAGE_COMPLETE | SHL_S_OK; return; }
    const int64_t avail = (int64_t)L - body_off;
    if (chunked) { r.status |= SHL_S_INCOMPLETE_BODY; return; }           // K2 (frame kernel) decides
    if (avail > 0 && avail >= r.content_len && !((r.status & SHL_F_CLEN_WIDE) && r.content_len > 0)) {
        r.status &= ~SHL_F_ON_BODY;
        r.status |= SHL_F_MESSAGE_COMPLETE | SHL_S_OK;
    } else r.status |= SHL_S_INCOMPLETE_BODY;
}

// ----------------------------------------------------------------------------------- one message
// Warp-cooperative.  Records go to dst[0..cap) (shared staging or global); r.n_headers counts all
// lines tokenised even when cap is smaller (the caller then re-runs with the exact room).
__device__ __forceinline__ void process_message(const uint8_t* m, const int L, shl_hdr_span* dst, const uint32_t cap,
                                                WarpScratch& ws, const int lane, const SwarK& K,
                                                const bool force_scalar, Msg& r) {
    r.status = (3u << SHL_F_VERSION_SHIFT) | SHL_F_IS_REQUEST;
    r.first_line_len = r.headers_len = r.sp1 = r.sp2 = r.fl_end = r.slash = r.n_headers = 0;
    r.status_code = 0; r.content_len = 0;
    if (L == 0) { r.status |= SHL_S_EMPTY; return; }

    const int lead = (int)(reinterpret_cast<uintptr_t>(m) & 15);
    const uint8_t* abase = m - lead;
    const int qL = lead + L;
    const int nchunks = (qL + 511) >> 9;
    const uint32_t lt_mask = (1u << lane) - 1u;

    int phase = 0;                     // 0: looking for the first CRLF, 1: header lines
    int idx = -1, h0 = 0, T = -1;
    bool fl_fast = !force_scalar, hdr_slow = force_scalar, get_or_head = false;
    uint32_t fl_code = 0, hdr_err = 0, n_hdr = 0;
    int open_start = 0;
    uint32_t carry_cnt = 0, carry_first = kNoPos;
    WellKnown wk; wk.n_te = wk.n_cl = 0; wk.te_vs = wk.te_ve = wk.cl_vs = wk.cl_ve = 0;

    uint4 cur = load_vec(abase, 0, lane, qL);
    for (int c = 0; c < nchunks; c++) {
        const uint4 nxt = (c + 1 < nchunks) ? load_vec(abase, c + 1, lane, qL) : make_uint4(0u, 0u, 0u, 0u);
        uint32_t nw = __shfl_down_sync(SHL_FULL, cur.x, 1);
        const uint32_t n0 = __shfl_sync(SHL_FULL, nxt.x, 0);
        if (lane == 31) nw = n0;
        const int pb = c * 512 + lane * 16 - lead;       // message position of my byte 0

        // ---- stage 1: byte classes -> position masks
        const uint32_t lf0 = eqf(cur.x, K.lf, K), lf1 = eqf(cur.y, K.lf, K), lf2 = eqf(cur.z, K.lf, K),
                       lf3 = eqf(cur.w, K.lf, K), lfn = eqf(nw, K.lf, K);
        const uint32_t sp0 = eqf(cur.x, K.sp, K), sp1f = eqf(cur.y, K.sp, K), sp2f = eqf(cur.z, K.sp, K),
                       sp3 = eqf(cur.w, K.sp, K), spn = eqf(nw, K.sp, K);
        const uint32_t crn = eqf(nw, K.cr, K);
        uint32_t crlf16 = mask16(eqf(cur.x, K.cr, K) & __funnelshift_r(lf0, lf1, 8),
                                 eqf(cur.y, K.cr, K) & __funnelshift_r(lf1, lf2, 8),
                                 eqf(cur.z, K.cr, K) & __funnelshift_r(lf2, lf3, 8),
                                 eqf(cur.w, K.cr, K) & __funnelshift_r(lf3, lfn, 8));
        uint32_t cs16 = mask16(eqf(cur.x, K.col, K) & __funnelshift_r(sp0, sp1f, 8),
                               eqf(cur.y, K.col, K) & __funnelshift_r(sp1f, sp2f, 8),
                               eqf(cur.z, K.col, K) & __funnelshift_r(sp2f, sp3, 8),
                               eqf(cur.w, K.col, K) & __funnelshift_r(sp3, spn, 8));
        const uint32_t xn = crn & (lfn >> 8);                                    // CRLF at rel 16 / 17
        const uint32_t ext = crlf16 | (((xn >> 7) & 1u) << 16) | (((xn >> 15) & 1u) << 17);
        uint32_t term16 = ext & (ext >> 2) & 0xffffu;
        const int lo = -pb, hi = L - pb;
        crlf16 &= rng16(lo, hi - 1);
        cs16 &= rng16(lo, hi - 1);
        term16 &= rng16(lo, hi - 3);

        if (phase == 0) {
            const uint32_t b = __ballot_sync(SHL_FULL, crlf16 != 0u);
            if (b == 0u) { fl_fast = false; cur = nxt; continue; }
            const int f = __ffs(b) - 1;
            const uint32_t m16 = __shfl_sync(SHL_FULL, crlf16, f);
            idx = c * 512 + f * 16 + (__ffs(m16) - 1) - lead;
            r.first_line_len = (uint32_t)(idx + 2);
            h0 = idx + 2;
            bool done = false;
            if (fl_fast && c == 0) {
                // canonical request line:  METHOD SP URL SP "HTTP/1.x" CRLF, all inside chunk 0
                const uint32_t sp16 = mask16(sp0, sp1f, sp2f, sp3) & rng16(lo, hi);
                const uint32_t sb = __ballot_sync(SHL_FULL, sp16 != 0u);
                if (sb != 0u) {
                    const int f1 = __ffs(sb) - 1;
                    const uint32_t s16 = __shfl_sync(SHL_FULL, sp16, f1);
                    const int p1 = f1 * 16 + (__ffs(s16) - 1) - lead;
                    const uint32_t s16b = s16 & (s16 - 1u);
                    const uint32_t sb2 = sb & ~((2u << f1) - 1u);
                    const int f2 = s16b ? f1 : (sb2 ? __ffs(sb2) - 1 : 0);
                    const uint32_t t16 = __shfl_sync(SHL_FULL, sp16, f2);
                    const uint32_t u16 = s16b ? s16b : t16;
                    const int p2 = (s16b || sb2) ? f2 * 16 + (__ffs(u16) - 1) - lead : -1;
                    if (p2 > p1 && p2 < idx && idx - p2 - 1 == 8) {
                        // lanes 0..7 peek c = m[p2+1 : idx), lanes 8..11 peek m[0:4)
                        uint32_t byte = 0;
                        if (lane < 8) byte = ldb(m + p2 + 1 + lane);
                        else if (lane < 12 && lane - 8 < L) byte = ldb(m + lane - 8);
                        // "HTTP/1.1" and "GET HEAD" as little-endian byte strings, indexed by lane
                        const unsigned long long lit = lane < 8 ? 0x312E312F50545448ull : 0x4441454820544547ull;
                        const uint32_t want = (uint32_t)(lit >> (8 * (lane & 7))) & 0xffu;
                        const uint32_t want_head = (uint32_t)(lit >> (8 * ((lane + 4) & 7))) & 0xffu;
                        const uint32_t okb = __ballot_sync(SHL_FULL, lane < 12 && byte == want);
                        const uint32_t hb = __ballot_sync(SHL_FULL, lane >= 8 && lane < 12 && byte == want_head);
                        const uint32_t b7 = __shfl_sync(SHL_FULL, byte, 7);
                        const uint32_t b8 = __shfl_sync(SHL_FULL, byte, 8), b9 = __shfl_sync(SHL_FULL, byte, 9);
                        const bool ver_ok = ((okb & 0x7fu) == 0x7fu) && (b7 == '0' || b7 == '1');
                        const bool is_ht = p1 >= 2 && b8 == 'H' && b9 == 'T';
                        if (ver_ok && !is_ht) {
                            r.sp1 = (uint32_t)p1; r.sp2 = (uint32_t)p2; r.fl_end = (uint32_t)idx;
                            r.slash = (uint32_t)(p2 + 5);
                            r.status = (r.status & ~SHL_F_VERSION_MASK) | ((b7 == '1' ? 1u : 0u) << SHL_F_VERSION_SHIFT);
                            r.status |= SHL_F_FIRST_LINE_DONE;
                            get_or_head = (p1 == 3 && ((okb >> 8) & 7u) == 7u) || (p1 == 4 && ((hb >> 8) & 0xfu) == 0xfu);
                            done = true;
                        }
                    }
                }
            }
            if (!done) fl_code = first_line_scalar(m, idx, r, get_or_head);
            phase = 1;
            if (fl_code) break;                                                    // the reference raises here
            open_start = h0;
            if (fl_code == 0u && h0 < L && is_ws(ldb(m + h0))) hdr_slow = true;    // strip() would move the start
        }

        // ---- phase 1: header lines of this chunk
        const uint32_t ge_h0 = ~below(clamp16(h0 - pb));
        uint32_t le16 = crlf16 & ge_h0;
        uint32_t se16 = cs16 & ge_h0;
        const uint32_t tb = __ballot_sync(SHL_FULL, (term16 & ge_h0) != 0u);
        if (tb != 0u) {
            const int ft = __ffs(tb) - 1;
            const uint32_t t16 = __shfl_sync(SHL_FULL, term16 & ge_h0, ft);
            const int kt = __ffs(t16) - 1;
            T = c * 512 + ft * 16 + kt - lead;
            const uint32_t keep = lane < ft ? 0xffffu : (lane == ft ? below(kt + 1) : 0u);
            le16 &= keep;
            se16 &= keep;
        }
        if (fl_code == 0u && !hdr_slow && hdr_err == 0u) {
            const int nle = __popc(le16);
            const uint32_t b1 = __ballot_sync(SHL_FULL, nle >= 1), b2 = __ballot_sync(SHL_FULL, nle >= 2);
            int pre, total;
            if (b2 == 0u) { pre = __popc(b1 & lt_mask); total = __popc(b1); }
            else {
                int inc = nle;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(SHL_FULL, inc, d); if (lane >= d) inc += t; }
                pre = inc - nle; total = __shfl_sync(SHL_FULL, inc, 31);
            }
            if (total > 32) hdr_slow = true;
            else {
                ws.cnt[lane] = lane == 0 ? carry_cnt : 0u;
                ws.first[lane] = lane == 0 ? carry_first : kNoPos;
                if (lane == 0) { ws.cnt[32] = 0u; ws.first[32] = kNoPos; }
                __syncwarp();
                { uint32_t mm = le16; int rk = pre; while (mm) { const int k = __ffs(mm) - 1; ws.ends[rk++] = (uint32_t)(pb + k); mm &= mm - 1u; } }
                { uint32_t mm = se16; while (mm) { const int k = __ffs(mm) - 1; const int rk = pre + __popc(le16 & below(k));
                      atomicAdd(&ws.cnt[rk], 1u); atomicMin(&ws.first[rk], (uint32_t)(pb + k)); mm &= mm - 1u; } }
                __syncwarp();
                bool bad = false; int start = 0, end = 0; uint32_t cn = 0, fs = 0;
                if (lane < total) {
                    start = lane == 0 ? open_start : (int)ws.ends[lane - 1] + 2;
                    end = (int)ws.ends[lane];
                    cn = ws.cnt[lane]; fs = ws.first[lane];
                    bad = cn != 1u;
                }
                const uint32_t eb = __ballot_sync(SHL_FULL, bad);
                const int nvalid = eb ? __ffs(eb) - 1 : total;
                if (eb) { const uint32_t ec = __shfl_sync(SHL_FULL, cn, nvalid); hdr_err = ec == 0u ? SHL_E_HDR_NO_SEP : SHL_E_HDR_MULTI_SEP; }
                int wkk = 0, vs = 0, ve = 0;
                if (lane < nvalid) {
                    vs = (int)fs + 2; ve = end;
                    if (vs < ve && is_ws(ldb(m + vs))) { do vs++; while (vs < ve && is_ws(ldb(m + vs))); }
                    if (ve > vs && is_ws(ldb(m + ve - 1))) { do ve--; while (ve > vs && is_ws(ldb(m + ve - 1))); }
                    const int klen = (int)fs - start;
                    if (n_hdr + (uint32_t)lane < cap) {
                        shl_hdr_span rec;
                        rec.key_off = (uint16_t)start; rec.key_len = (uint16_t)klen;
                        rec.val_off = (uint16_t)vs; rec.val_len = (uint16_t)(ve - vs);
                        dst[n_hdr + lane] = rec;
                    }
                    if (klen == 14) { if (span_eq_lower(m, start, (int)fs, "content-length")) wkk = 1; }
                    else if (klen == 17) { if (span_eq_lower(m, start, (int)fs, "transfer-encoding")) wkk = 2; }
                }
                const uint32_t clb = __ballot_sync(SHL_FULL, wkk == 1), teb = __ballot_sync(SHL_FULL, wkk == 2);
                if (clb) {
                    if (!wk.n_cl) { const int s = __ffs(clb) - 1; wk.cl_vs = __shfl_sync(SHL_FULL, vs, s); wk.cl_ve = __shfl_sync(SHL_FULL, ve, s); }
                    wk.n_cl += __popc(clb);
                }
                if (teb) {
                    if (!wk.n_te) { const int s = __ffs(teb) - 1; wk.te_vs = __shfl_sync(SHL_FULL, vs, s); wk.te_ve = __shfl_sync(SHL_FULL, ve, s); }
                    wk.n_te += __popc(teb);
                }
                n_hdr += (uint32_t)nvalid;
                if (total) open_start = (int)ws.ends[total - 1] + 2;
                carry_cnt = ws.cnt[total]; carry_first = ws.first[total];
                __syncwarp();
            }
        }
        if (T >= 0) break;
        cur = nxt;
    }

    if (phase == 0) { r.status |= SHL_S_INCOMPLETE_LINE; return; }
    if (fl_code) { r.status |= fl_code; return; }
    if (T < 0) { r.status |= SHL_S_INCOMPLETE_HEADERS; return; }
    if (T + 4 > 65536) { r.status |= SHL_E_LIB_HDR_TOO_LARGE; return; }
    r.headers_len = (uint32_t)(T + 4 - h0);
    if (!hdr_slow && is_ws(ldb(m + T - 1))) hdr_slow = true;              // strip() would move the end
    if (hdr_slow) {
        wk.n_te = wk.n_cl = 0; hdr_err = 0; n_hdr = 0;
        hdr_err = headers_scalar(m, h0, T + 4, dst, cap, lane, n_hdr, wk);
    }
    r.n_headers = n_hdr;
    if (hdr_err) { r.status |= hdr_err; return; }
    finalize_headers(m, L, T + 4, get_or_head, wk, r);
}

// ----------------------------------------------------------------------------------- kernel
template <bool kPlace>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) k_route(const RouteArgs a) {
    __shared__ shl_hdr_span stage[kWarpsPerBlock][kStageCap];
    __shared__ WarpScratch scratch[kWarpsPerBlock];
    __shared__ uint32_t s_codes[64];
    extern __shared__ uint32_t s_hist[];                       // kPlace: n_hosts bins (when they fit)

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const SwarK K = swar_consts();
    if (threadIdx.x < 64) s_codes[threadIdx.x] = 0u;
    const bool smem_hist = kPlace && a.place.hist != nullptr && a.n_hosts <= 4096u;
    if (smem_hist) for (uint32_t i = threadIdx.x; i < a.n_hosts; i += blockDim.x) s_hist[i] = 0u;
    __syncthreads();

    const uint64_t group = (uint64_t)blockIdx.x * kWarpsPerBlock + warp;
    const uint64_t first = group * kGroup;
    if (first < a.n) {
        const uint64_t mine = first + lane;
        const bool have = mine < a.n;
        const uint64_t my_off = have ? a.off[mine] : 0ull;
        const uint32_t my_len = have ? a.len[mine] : 0u;
        const int count = (int)min((uint64_t)kGroup, a.n - first);

        Msg keep; keep.status = 0; keep.first_line_len = keep.headers_len = keep.sp1 = keep.sp2 = 0;
        keep.fl_end = keep.slash = keep.n_headers = 0; keep.status_code = 0; keep.content_len = 0;
        uint32_t staged = 0;          // records currently staged by this warp
        uint32_t my_stage_at = 0;     // where my message's records start in the staging buffer
        bool my_ovf = false;
        bool any_ovf = false;

        for (int j = 0; j < count; j++) {
            const uint64_t o = __shfl_sync(SHL_FULL, my_off, j);
            const int L = (int)min(__shfl_sync(SHL_FULL, my_len, j), 0x7fffffc0u);
            Msg r;
            process_message(a.arena + o, L, &stage[warp][staged], kStageCap - staged, scratch[warp], lane, K,
                            a.force_scalar != 0, r);
            const bool ovf = r.n_headers > kStageCap - staged;
            if (lane == j) { keep = r; my_stage_at = staged; my_ovf = ovf; }
            if (ovf) any_ovf = true; else staged += r.n_headers;
            __syncwarp();
        }

        // ---- allocate the group's header records, copy the staged ones out
        uint32_t nh = have ? keep.n_headers : 0u;
        uint32_t inc = nh;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(SHL_FULL, inc, d); if (lane >= d) inc += t; }
        const uint32_t total = __shfl_sync(SHL_FULL, inc, 31);
        unsigned long long base = 0;
        if (lane == 0 && total) base = atomicAdd(reinterpret_cast<unsigned long long*>(a.out.hdr_used), (unsigned long long)total);
        base = __shfl_sync(SHL_FULL, base, 0);
        unsigned long long my_base = base + (inc - nh);
        const bool fits = my_base + nh <= a.out.hdr_cap;
        if (!fits) {                                           // record arena full: library limit
            keep.status = (keep.status & ~0xffu & ~SHL_F_MESSAGE_COMPLETE) | SHL_E_LIB_RECORDS_FULL;
            nh = 0; keep.n_headers = 0;
        }
        const uint32_t all_fit = __ballot_sync(SHL_FULL, fits || !have);
        __syncwarp();
        if (!any_ovf && all_fit == SHL_FULL) {
            // common case: staging order == global order, one coalesced copy (8 B per lane)
            const uint2* src = reinterpret_cast<const uint2*>(&stage[warp][0]);
            uint2* dstp = reinterpret_cast<uint2*>(a.out.hdr + base);
            for (uint32_t i = lane; i < total; i += 32) dstp[i] = src[i];
        } else {
            for (int j = 0; j < count; j++) {
                const uint32_t n_j = __shfl_sync(SHL_FULL, nh, j);
                const unsigned long long b_j = __shfl_sync(SHL_FULL, my_base, j);
                const uint32_t at_j = __shfl_sync(SHL_FULL, my_stage_at, j);
                const bool ovf_j = __shfl_sync(SHL_FULL, (int)my_ovf, j) != 0;
                if (n_j == 0u) continue;
                if (!ovf_j) {
                    for (uint32_t i = lane; i < n_j; i += 32) a.out.hdr[b_j + i] = stage[warp][at_j + i];
                } else {                                      // did not fit the staging buffer: redo, straight to global
                    const uint64_t o = __shfl_sync(SHL_FULL, my_off, j);
                    const int L = (int)min(__shfl_sync(SHL_FULL, my_len, j), 0x7fffffc0u);
                    Msg r2;
                    process_message(a.arena + o, L, a.out.hdr + b_j, n_j, scratch[warp], lane, K, a.force_scalar != 0, r2);
                }
            }
        }

        // ---- per-message fixed fields, lane per message (coalesced struct-of-arrays stores)
        if (have) {
            a.out.status[mine] = keep.status;
            a.out.first_line_len[mine] = keep.first_line_len;
            a.out.headers_len[mine] = keep.headers_len;
            a.out.sp1[mine] = keep.sp1; a.out.sp2[mine] = keep.sp2;
            a.out.fl_end[mine] = keep.fl_end; a.out.slash[mine] = keep.slash;
            a.out.status_code[mine] = keep.status_code;
            a.out.n_headers[mine] = keep.n_headers;
            a.out.hdr_base[mine] = (uint32_t)my_base;
            a.out.content_len[mine] = keep.content_len;
            atomicAdd(&s_codes[keep.status & 63u], 1u);
        }

        // ---- fused key + MD5 + ketama: lane j hashes message j's URL  (Server.py:325-335)
        if (kPlace && have) {
            const bool routable = (keep.status & SHL_F_MESSAGE_COMPLETE) && (keep.status & SHL_F_IS_REQUEST) &&
                                  (keep.status & SHL_F_FIRST_LINE_DONE) && (keep.status & 0xffu) == SHL_S_OK;
            uint32_t h[4] = {0u, 0u, 0u, 0u}, node = 0xffffffffu;
            if (routable) {
                md5_key(a.arena + my_off + keep.sp1 + 1u, keep.sp2 - keep.sp1 - 1u, h);
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
