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
// Records go to dst[k * stride] for k < cap (shared staging: stride 32; global: stride 1);
// r.n_headers counts every tokenised line even when cap is smaller (the caller then re-runs the
// message with exactly that much room in the global record arena).
__device__ __forceinline__ void scan_message(const uint8_t* m, const int L, shl_hdr_span* dst, const int stride,
                                             const uint32_t cap, const SwarK& K, const bool force_scalar, Msg& r) {
    r.status = (3u << SHL_F_VERSION_SHIFT) | SHL_F_IS_REQUEST;
    r.first_line_len = r.headers_len = r.sp1 = r.sp2 = r.fl_end = r.slash = r.n_headers = 0;
    r.status_code = 0; r.content_len = 0;
    if (L == 0) { r.status |= SHL_S_EMPTY; return; }

    const int lead = (int)(reinterpret_cast<uintptr_t>(m) & 15);
    const uint4* vp = reinterpret_cast<const uint4*>(m - lead);
    const int nvec = (lead + L + 15) >> 4;
    int p = -lead;                                 // message position of byte 0 of the current vector
    int idx = -1, sp1 = -1, sp2 = -1, h0 = 0, T = -1;
    int line_start = 0, sep_first = 0;
    uint32_t sep_cnt = 0, n_hdr = 0, hdr_err = 0, prev_cr = 0, prev_co = 0;
    bool suspect = false;

    uint4 cur = __ldg(vp);
    for (int v = 0; v < nvec; v++, p += 16) {
        const uint4 nxt = (v + 1 < nvec) ? __ldg(vp + v + 1) : make_uint4(0u, 0u, 0u, 0u);
        // ---- byte classes -> "pair ends here" position masks
        const uint32_t cr0 = eqf(cur.x, K.cr, K), cr1 = eqf(cur.y, K.cr, K), cr2 = eqf(cur.z, K.cr, K), cr3 = eqf(cur.w, K.cr, K);
        const uint32_t co0 = eqf(cur.x, K.col, K), co1 = eqf(cur.y, K.col, K), co2 = eqf(cur.z, K.col, K), co3 = eqf(cur.w, K.col, K);
        const uint32_t sx0 = eqf(cur.x, K.sp, K), sx1 = eqf(cur.y, K.sp, K), sx2 = eqf(cur.z, K.sp, K), sx3 = eqf(cur.w, K.sp, K);
        uint32_t c16 = mask16(eqf(cur.x, K.lf, K) & __funnelshift_l(prev_cr, cr0, 8), eqf(cur.y, K.lf, K) & __funnelshift_l(cr0, cr1, 8),
                              eqf(cur.z, K.lf, K) & __funnelshift_l(cr1, cr2, 8), eqf(cur.w, K.lf, K) & __funnelshift_l(cr2, cr3, 8));
        uint32_t s16 = mask16(sx0 & __funnelshift_l(prev_co, co0, 8), sx1 & __funnelshift_l(co0, co1, 8),
                              sx2 & __funnelshift_l(co1, co2, 8), sx3 & __funnelshift_l(co2, co3, 8));
        prev_cr = cr3; prev_co = co3;
        if (p < 1 || p + 16 > L) { const uint32_t vm = rng16(1 - p, L - p); c16 &= vm; s16 &= vm; }
        if (idx < 0) {                             // still on the first line: note its first two spaces
            uint32_t z = mask16(sx0, sx1, sx2, sx3);
            if (p < 0 || p + 16 > L) z &= rng16(-p, L - p);
            if (c16) z &= below(__ffs(c16) - 1);
            if (z && sp1 < 0) { sp1 = p + __ffs(z) - 1; z &= z - 1u; }
            if (z && sp1 >= 0 && sp2 < 0) sp2 = p + __ffs(z) - 1;
        }
        int lo = 0;
        while (c16) {
            const int k = __ffs(c16) - 1;          // the LF; the CR sits at k - 1
            c16 &= c16 - 1u;
            const int cr_pos = p + k - 1;
            const uint32_t seg = s16 & below(k) & ~below(lo);
            lo = k + 1;
            if (idx < 0) { idx = cr_pos; h0 = line_start = cr_pos + 2; continue; }
            if (seg) { if (!sep_cnt) sep_first = p + __ffs(seg) - 2; sep_cnt += (uint32_t)__popc(seg); }
            if (cr_pos == line_start && line_start > h0) { T = line_start - 2; break; }   // empty line: CRLFCRLF at T
            if (!hdr_err) {
                if (sep_cnt != 1u) hdr_err = sep_cnt ? (uint32_t)SHL_E_HDR_MULTI_SEP : (uint32_t)SHL_E_HDR_NO_SEP;
                else {
                    int vs = sep_first + 2, ve = cr_pos;
                    const uint32_t b0 = vs < ve ? ldb(m + vs) : 'x', b1 = ve > vs ? ldb(m + ve - 1) : 'x';
                    if (is_ws(b0) || is_ws(b1)) {
                        while (vs < ve && is_ws(ldb(m + vs))) vs++;
                        while (ve > vs && is_ws(ldb(m + ve - 1))) ve--;
                    }
                    const int klen = sep_first - line_start;
                    if (n_hdr < cap) {
                        shl_hdr_span rec;
                        rec.key_off = (uint16_t)line_start; rec.key_len = (uint16_t)klen;
                        rec.val_off = (uint16_t)vs; rec.val_len = (uint16_t)(ve - vs);
                        dst[(size_t)n_hdr * stride] = rec;
                    }
                    suspect |= (klen == 14) | (klen == 17);
                    n_hdr++;
                }
            }
            line_start = cr_pos + 2; sep_cnt = 0;
        }
        if (T >= 0) break;
        if (idx >= 0) {
            const uint32_t seg = s16 & ~below(lo);
            if (seg) { if (!sep_cnt) sep_first = p + __ffs(seg) - 2; sep_cnt += (uint32_t)__popc(seg); }
        }
        cur = nxt;
    }

    // ---- lanes are converged again from here on
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
    if (force_scalar || is_ws(ldb(m + h0)) || is_ws(ldb(m + T - 1))) {    // strip() would move the block's ends
        WellKnown w2 = wk; uint32_t n2 = 0;
        hdr_err = headers_scalar(m, h0, T + 4, dst, stride, cap, &n2, &w2);
        wk = w2; n_hdr = n2;
    } else if (suspect && !hdr_err && n_hdr <= cap) {
        for (uint32_t k = 0; k < n_hdr; k++) {
            const shl_hdr_span rec = dst[(size_t)k * stride];
            const int ko = rec.key_off, vs = rec.val_off, ve = rec.val_off + rec.val_len;
            if (rec.key_len == 14 && span_eq_lower(m, ko, ko + 14, "content-length")) { if (!wk.n_cl) { wk.cl_vs = vs; wk.cl_ve = ve; } wk.n_cl++; }
            if (rec.key_len == 17 && span_eq_lower(m, ko, ko + 17, "transfer-encoding")) { if (!wk.n_te) { wk.te_vs = vs; wk.te_ve = ve; } wk.n_te++; }
        }
    }
    r.n_headers = n_hdr;
    if (hdr_err) { r.status |= hdr_err; return; }
    finalize_headers(m, L, T + 4, get_or_head, wk, r.status, r.content_len);
}

// ----------------------------------------------------------------------------------- kernel
template <bool kPlace>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, 8) k_route(const RouteArgs a) {
    __shared__ shl_hdr_span stage[kWarpsPerBlock][kLaneCap * 32];
    __shared__ uint32_t s_codes[64];
    extern __shared__ uint32_t s_hist[];                       // kPlace: n_hosts bins (when they fit)

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const SwarK K = swar_consts();
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
