/* shl_oracle.c — CPU restatement of the reference's hot path.   *** TEST INFRASTRUCTURE ***
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this.  The product (libshellac_b200.so) never links, imports or calls it.
 *
 * What it restates (plain scalar C, one message at a time, no SIMD, no threads):
 *   - HttpParser.parse / _parse_first_line / _parse_headers / body framing / _parse_chunked
 *     (/root/reference/src/python/shellac/server/HttpParser.py:140-337) driven over a WHOLE
 *     message the way the reference's own test drives it (test/HttpParserTests.py:36-39):
 *         while data and not p.message_complete(): c = p.parse(data, len(data)); data = data[c:]
 *     with Python-2 `str` semantics (bytes; whitespace = {9..13, 32}; ASCII-only lower()).
 *   - cache key = request.url() verbatim (Server.py:327).
 *   - ketama placement as configured at Server.py:81-83.  The arithmetic is in libmemcached 0.44
 *     (third-party, NOT vendored, absent from this machine): hosts.c update_continuum,
 *     hash.c dispatch_host, libhashkit/md5.c.  Restated from its published algorithm.
 *
 * Parity pinning:
 *   parser    — PINNED: tests/test_oracle_golden.py checks every vector in
 *               tests/golden/parser_golden.json.gz, produced by running the unmodified reference
 *               (tests/golden/make_golden.py), incl. the literals of test/HttpParserTests.py.
 *   placement — PARITY UNPINNED against libmemcached itself (no reference test or fixture holds a
 *               placement result; pylibmc/libmemcached are absent).  Pinned only at the MD5 stage
 *               (RFC 1321 test suite + hashlib) and against a hashlib+bisect restatement.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <math.h>
#include "../include/shellac_b200.h"

/* ------------------------------------------------------------------ python-2 string helpers */
static int is_ws(uint8_t c) { return c == 32 || (c >= 9 && c <= 13); }          /* str.isspace, C locale */
static int is_digit(uint8_t c) { return c >= '0' && c <= '9'; }
static int is_alnum(uint8_t c) { return is_digit(c) || ((c | 32) >= 'a' && (c | 32) <= 'z'); }
static int hexval(uint8_t c) {
    if (is_digit(c)) return c - '0';
    c |= 32;
    if (c >= 'a' && c <= 'f') return c - 'a' + 10;
    return -1;
}
static uint8_t lower(uint8_t c) { return (c >= 'A' && c <= 'Z') ? (uint8_t)(c + 32) : c; }

/* float(s): 1 if s is a valid literal.  Grammar of PyOS_string_to_double after str.strip():
 *   [+-] ( digits [ '.' digits* ] | '.' digits ) [ (e|E) [+-] digits ]  |  [+-] (inf|infinity|nan)  */
static int py_float_ok(const uint8_t* s, int64_t n) {
    int64_t i = 0, j = n;
    while (i < j && is_ws(s[i])) i++;
    while (j > i && is_ws(s[j - 1])) j--;
    if (i == j) return 0;
    if (s[i] == '+' || s[i] == '-') i++;
    int64_t r = j - i;
    if (r == 3 && lower(s[i]) == 'n' && lower(s[i + 1]) == 'a' && lower(s[i + 2]) == 'n') return 1;
    if (r == 3 || r == 8) {
        static const char inf[] = "infinity";
        int ok = 1;
        for (int64_t k = 0; k < r; k++) if (lower(s[i + k]) != (uint8_t)inf[k]) { ok = 0; break; }
        if (ok) return 1;
    }
    int64_t nd = 0;
    while (i < j && is_digit(s[i])) { i++; nd++; }
    if (i < j && s[i] == '.') { i++; while (i < j && is_digit(s[i])) { i++; nd++; } }
    if (nd == 0) return 0;
    if (i < j && (s[i] == 'e' || s[i] == 'E')) {
        i++;
        if (i < j && (s[i] == '+' || s[i] == '-')) i++;
        int64_t ne = 0;
        while (i < j && is_digit(s[i])) { i++; ne++; }
        if (ne == 0) return 0;
    }
    return i == j;
}

/* int(s) base 10.  returns 0 invalid, 1 valid & fits int64 (*v), 2 valid but wider than int64. */
static int py_int10(const uint8_t* s, int64_t n, int64_t* v) {
    int64_t i = 0, j = n;
    while (i < j && is_ws(s[i])) i++;
    while (j > i && is_ws(s[j - 1])) j--;
    int neg = 0;
    if (i < j && (s[i] == '+' || s[i] == '-')) { neg = s[i] == '-'; i++; }
    if (i == j) return 0;
    uint64_t acc = 0; int wide = 0;
    for (; i < j; i++) {
        if (!is_digit(s[i])) return 0;
        uint64_t d = (uint64_t)(s[i] - '0');
        if (acc > (UINT64_MAX - d) / 10) wide = 1; else acc = acc * 10 + d;
    }
    if (wide) return 2;
    if (!neg && acc > (uint64_t)INT64_MAX) return 2;
    if (neg && acc > (uint64_t)INT64_MAX + 1) return 2;
    *v = neg ? (int64_t)(0 - acc) : (int64_t)acc;
    return 1;
}

/* ------------------------------------------------------------------ first line + headers */
static int64_t find_crlf(const uint8_t* m, int64_t from, int64_t L) {
    for (int64_t i = from; i + 1 < L; i++) if (m[i] == '\r' && m[i + 1] == '\n') return i;
    return -1;
}
static int64_t find_crlfcrlf(const uint8_t* m, int64_t from, int64_t L) {
    for (int64_t i = from; i + 3 < L; i++)
        if (m[i] == '\r' && m[i + 1] == '\n' && m[i + 2] == '\r' && m[i + 3] == '\n') return i;
    return -1;
}
static int span_eq(const uint8_t* m, int64_t a, int64_t b, const char* lit) {
    int64_t n = (int64_t)strlen(lit);
    return (b - a) == n && memcmp(m + a, lit, (size_t)n) == 0;
}
static int span_eq_lower(const uint8_t* m, int64_t a, int64_t b, const char* lit) {
    int64_t n = (int64_t)strlen(lit);
    if ((b - a) != n) return 0;
    for (int64_t k = 0; k < n; k++) if (lower(m[a + k]) != (uint8_t)lit[k]) return 0;
    return 1;
}

typedef struct {
    uint32_t status, first_line_len, headers_len, sp1, sp2, fl_end, slash, n_headers;
    int32_t status_code;
    int64_t content_len;
} msg_rec;

/* HttpParser._parse_first_line (HttpParser.py:304-321) on m[0:idx+2). returns outcome code or 0 */
static uint32_t first_line(const uint8_t* m, int64_t idx, msg_rec* r) {
    int64_t e = idx + 2;
    while (e > 0 && is_ws(m[e - 1])) e--;                          /* rstrip()                 */
    int64_t sp1 = -1, sp2 = -1;
    for (int64_t i = 0; i < e; i++) if (m[i] == ' ') { sp1 = i; break; }
    if (sp1 >= 0) for (int64_t i = sp1 + 1; i < e; i++) if (m[i] == ' ') { sp2 = i; break; }
    if (sp1 < 0 || sp2 < 0) return SHL_E_FIRSTLINE_PARTS;           /* split(' ', 2) -> 3 parts */
    r->sp1 = (uint32_t)sp1; r->sp2 = (uint32_t)sp2; r->fl_end = (uint32_t)e;
    int is_resp = sp1 >= 2 && m[0] == 'H' && m[1] == 'T';           /* a.startswith('HT')       */
    int64_t a = is_resp ? 0 : sp2 + 1, b = is_resp ? sp1 : e;       /* the part split('/') sees */
    int64_t slash = -1, ns = 0;
    for (int64_t i = a; i < b; i++) if (m[i] == '/') { if (!ns) slash = i; ns++; }
    if (is_resp) r->status &= ~SHL_F_IS_REQUEST;   /* set before split('/') can raise (HttpParser.py:310,317) */
    if (ns != 1) return SHL_E_VERSION_SPLIT;
    r->slash = (uint32_t)slash;
    if (!py_float_ok(m + slash + 1, b - slash - 1)) return SHL_E_VERSION_FLOAT;
    uint32_t vc = 2;
    if (span_eq(m, slash + 1, b, "1.1")) vc = 1; else if (span_eq(m, slash + 1, b, "1.0")) vc = 0;
    r->status = (r->status & ~SHL_F_VERSION_MASK) | (vc << SHL_F_VERSION_SHIFT);
    if (is_resp) {
        int64_t v = 0;
        int k = py_int10(m + sp1 + 1, sp2 - sp1 - 1, &v);
        if (k == 0) return SHL_E_STATUS_INT;
        if (k == 2 || v > INT32_MAX || v < INT32_MIN) r->status |= SHL_F_STATUS_WIDE;
        else r->status_code = (int32_t)v;
    }
    r->status |= SHL_F_FIRST_LINE_DONE;
    return 0;
}

/* HttpParser._parse_headers (HttpParser.py:323-337) on m[h0:h1) (h1 = one past CRLFCRLF).
 * Emits records; n_te / n_cl count transfer-encoding / content-length keys, te / cl hold the first. */
static uint32_t headers(const uint8_t* m, int64_t h0, int64_t h1, shl_hdr_span* hdr, uint64_t cap,
                        uint32_t* n_out, int* n_te, shl_hdr_span* te, int* n_cl, shl_hdr_span* cl) {
    int64_t s = h0, e = h1;
    while (s < e && is_ws(m[s])) s++;                              /* strip()                   */
    while (e > s && is_ws(m[e - 1])) e--;
    uint32_t n = 0; *n_te = 0; *n_cl = 0;
    int64_t ls = s;
    for (;;) {
        int64_t le = -1;                                           /* next CRLF inside [ls,e)   */
        for (int64_t i = ls; i + 1 < e; i++) if (m[i] == '\r' && m[i + 1] == '\n') { le = i; break; }
        int64_t lend = le < 0 ? e : le;
        int64_t sep = -1, nsep = 0;                                /* header.split(': ')        */
        for (int64_t i = ls; i + 1 < lend; i++)
            if (m[i] == ':' && m[i + 1] == ' ') { if (!nsep) sep = i; nsep++; i++; }
        if (nsep != 1) { *n_out = n; return nsep == 0 ? SHL_E_HDR_NO_SEP : SHL_E_HDR_MULTI_SEP; }
        int64_t vs = sep + 2, ve = lend;
        while (vs < ve && is_ws(m[vs])) vs++;                      /* value.strip()             */
        while (ve > vs && is_ws(m[ve - 1])) ve--;
        shl_hdr_span rec = { (uint16_t)ls, (uint16_t)(sep - ls), (uint16_t)vs, (uint16_t)(ve - vs) };
        if ((uint64_t)n >= cap) { *n_out = n; return SHL_E_LIB_RECORDS_FULL; }
        hdr[n++] = rec;
        if (span_eq_lower(m, ls, sep, "transfer-encoding")) { if (!*n_te) *te = rec; (*n_te)++; }
        if (span_eq_lower(m, ls, sep, "content-length")) { if (!*n_cl) *cl = rec; (*n_cl)++; }
        if (le < 0) break;
        ls = le + 2;
    }
    *n_out = n;
    return 0;
}

/* One whole message, header states only (parse() calls 1 and 2 + the Content-Length body call). */
static void parse_one(const uint8_t* m, int64_t L, msg_rec* r, shl_hdr_span* hdr, uint64_t cap) {
    memset(r, 0, sizeof *r);
    r->status = (3u << SHL_F_VERSION_SHIFT) | SHL_F_IS_REQUEST;      /* _is_request defaults to True (HttpParser.py:54) */
    if (L == 0) { r->status |= SHL_S_EMPTY; return; }
    int64_t idx = find_crlf(m, 0, L);
    if (idx < 0) { r->status |= SHL_S_INCOMPLETE_LINE; return; }
    r->first_line_len = (uint32_t)(idx + 2);
    uint32_t code = first_line(m, idx, r);
    if (code) { r->status |= code; return; }
    int64_t h0 = idx + 2;
    int64_t t = find_crlfcrlf(m, h0, L);
    if (t < 0) { r->status |= SHL_S_INCOMPLETE_HEADERS; return; }
    if (t + 4 > 65536) { r->status |= SHL_E_LIB_HDR_TOO_LARGE; return; }
    r->headers_len = (uint32_t)(t + 4 - h0);
    int n_te, n_cl; shl_hdr_span te = {0, 0, 0, 0}, cl = {0, 0, 0, 0};
    code = headers(m, h0, t + 4, hdr, cap, &r->n_headers, &n_te, &te, &n_cl, &cl);
    if (code) { r->status |= code; return; }
    r->status |= SHL_F_HEADERS_COMPLETE | SHL_F_ON_BODY;                       /* HttpParser.py:187-189 */
    int complete = 0;
    if ((r->status & SHL_F_IS_REQUEST) &&
        (span_eq(m, 0, r->sp1, "GET") || span_eq(m, 0, r->sp1, "HEAD"))) complete = 1;   /* :191 */
    int chunked = n_te == 1 && span_eq(m, te.val_off, te.val_off + te.val_len, "chunked"); /* :194 */
    if (chunked) r->status |= SHL_F_CHUNKED;
    else {
        if (n_cl > 1) { r->status |= SHL_E_CLEN_LIST | (complete ? SHL_F_MESSAGE_COMPLETE : 0); return; }
        if (n_cl == 1) {
            int64_t v = 0;
            int k = py_int10(m + cl.val_off, cl.val_len, &v);
            if (k == 0) { r->status |= SHL_E_CLEN_INT | (complete ? SHL_F_MESSAGE_COMPLETE : 0); return; }
            if (k == 2) { r->status |= SHL_F_CLEN_WIDE; v = (m[cl.val_off] == '-') ? INT64_MIN : INT64_MAX; }
            r->content_len = v;
        }
        if (r->content_len == 0 && !(r->status & SHL_F_CLEN_WIDE)) {           /* :198-200 */
            r->status &= ~SHL_F_ON_BODY; complete = 1;
        }
    }
    if (complete) { r->status |= SHL_F_MESSAGE_COMPLETE | SHL_S_OK; return; }
    int64_t avail = L - (t + 4);
    if (chunked) { r->status |= SHL_S_INCOMPLETE_BODY; return; }               /* the framer decides */
    if (avail > 0 && avail >= r->content_len && !(r->status & SHL_F_CLEN_WIDE && r->content_len > 0)) {
        r->status &= ~SHL_F_ON_BODY;                                           /* :213-223 */
        r->status |= SHL_F_MESSAGE_COMPLETE | SHL_S_OK;
    } else r->status |= SHL_S_INCOMPLETE_BODY;
}

int orc_parse(const uint8_t* arena, const uint64_t* off, const uint32_t* len, uint64_t n,
              const shl_parse_out* out) {
    uint64_t used = 0, kused = 0;
    for (uint64_t i = 0; i < n; i++) {
        msg_rec r;
        parse_one(arena + off[i], (int64_t)len[i], &r, out->hdr + used, out->hdr_cap - used);
        /* key = request.url() (Server.py:327); pylibmc refuses keys over 250 bytes before hashing (:335) */
        if ((r.status & SHL_F_IS_REQUEST) && (r.status & SHL_F_FIRST_LINE_DONE) && r.sp2 - r.sp1 - 1 > SHL_MAX_KEY_LEN)
            r.status |= SHL_F_KEY_TOO_LONG;
        if (out->lkeys) {                                    /* key.lower(), HttpParser.py:330 */
            out->lkey_base[i] = (uint32_t)kused;
            for (uint32_t k = 0; k < r.n_headers; k++) {
                const shl_hdr_span* h = out->hdr + used + k;
                for (uint32_t j = 0; j < h->key_len && kused < out->lkeys_cap; j++)
                    out->lkeys[kused++] = lower(arena[off[i] + h->key_off + j]);
            }
        }
        out->status[i] = r.status; out->first_line_len[i] = r.first_line_len;
        out->headers_len[i] = r.headers_len; out->sp1[i] = r.sp1; out->sp2[i] = r.sp2;
        out->fl_end[i] = r.fl_end; out->slash[i] = r.slash; out->status_code[i] = r.status_code;
        out->n_headers[i] = r.n_headers; out->hdr_base[i] = (uint32_t)used;
        out->content_len[i] = r.content_len;
        used += r.n_headers;
    }
    *out->hdr_used = used;
    if (out->lkeys_used) *out->lkeys_used = kused;
    return 0;
}

/* ------------------------------------------------------------------ chunked framing
 * CHUNK_HEADER_RX = (\r\n)?[a-z0-9]+(;[a-z0-9]+="?[a-z0-9\-_]+"?)?\r\n   IGNORECASE, anchored
 * (HttpParser.py:39-40).  Returns the match length at m[p:L) or -1.  [sz_a, sz_b) = the size run. */
static int64_t chunk_header_match(const uint8_t* m, int64_t p, int64_t L, int64_t* sz_a, int64_t* sz_b) {
    int64_t i = p;
    if (i + 1 < L && m[i] == '\r' && m[i + 1] == '\n') i += 2;
    int64_t a = i;
    while (i < L && is_alnum(m[i])) i++;
    if (i == a) return -1;
    *sz_a = a; *sz_b = i;
    if (i + 1 < L && m[i] == '\r' && m[i + 1] == '\n') return i + 2 - p;
    if (i < L && m[i] == ';') {
        int64_t j = i + 1, k = j;
        while (j < L && is_alnum(m[j])) j++;
        if (j == k || j >= L || m[j] != '=') return -1;
        j++;
        if (j < L && m[j] == '"') j++;
        k = j;
        while (j < L && (is_alnum(m[j]) || m[j] == '-' || m[j] == '_')) j++;
        if (j == k) return -1;
        if (j < L && m[j] == '"') j++;
        if (j + 1 < L && m[j] == '\r' && m[j + 1] == '\n') return j + 2 - p;
    }
    return -1;
}
/* int(size, 16) (HttpParser.py:234-241): optional 0x/0X, then hex digits. 0 = ValueError, 1 ok.
 * Saturates at 2^63 (Python would carry a long; any such chunk can never be satisfied). */
static int py_int16(const uint8_t* s, int64_t n, uint64_t* v) {
    int64_t i = 0;
    if (n >= 2 && s[0] == '0' && (s[1] | 32) == 'x') i = 2;
    if (i == n) return 0;
    uint64_t acc = 0;
    for (; i < n; i++) {
        int h = hexval(s[i]);
        if (h < 0) return 0;
        if (acc >> 59) acc = (uint64_t)1 << 63; else acc = acc * 16 + (uint64_t)h;
        if (acc > ((uint64_t)1 << 63)) acc = (uint64_t)1 << 63;
    }
    *v = acc;
    return 1;
}

int orc_frame(const uint8_t* arena, const uint64_t* off, const uint32_t* len, uint64_t n,
              const uint32_t* status, const uint32_t* first_line_len, const uint32_t* headers_len,
              const shl_frame_out* out) {
    uint64_t used = 0;
    for (uint64_t i = 0; i < n; i++) {
        out->fstatus[i] = 0xff; out->n_chunks[i] = 0; out->chunk_base[i] = (uint32_t)used;
        out->body_end[i] = 0; out->body_bytes[i] = 0; out->last_call[i] = 0;
        uint32_t st = status[i];
        if (!(st & SHL_F_CHUNKED) || !(st & SHL_F_HEADERS_COMPLETE) || (st & SHL_F_MESSAGE_COMPLETE) ||
            SHL_CODE(st) >= 16) continue;
        const uint8_t* m = arena + off[i];
        int64_t L = len[i], p = (int64_t)first_line_len[i] + headers_len[i];
        uint32_t nch = 0, fs = SHL_S_INCOMPLETE_BODY, last = 0; uint64_t bytes = 0;
        while (p < L) {
            int64_t a, b;
            int64_t idx = chunk_header_match(m, p, L, &a, &b);
            if (idx < 0) { last = (uint32_t)(L - p); p = L; break; }           /* :255-256 return length */
            uint64_t size;
            if (!py_int16(m + a, b - a, &size)) { fs = SHL_E_CHUNK_HEX; break; } /* ValueError */
            int64_t left = L - (p + idx);
            if (size == 0) {                                                    /* :268-276 */
                if (left >= 2 && m[p + idx] == '\r' && m[p + idx + 1] == '\n') {
                    last = (uint32_t)(idx + 2); p += idx + 2; fs = SHL_S_OK;
                } else {
                    p += idx;                                                   /* returns idx; then :278-287 */
                    last = (uint32_t)idx;
                    if (p < L) { last = (uint32_t)(L - p); p = L; }             /* next call returns length */
                }
                break;
            }
            if ((uint64_t)left < size) { last = (uint32_t)(idx + left); p = L; break; }   /* :291-294 */
            if (used + nch >= out->chunk_cap) { fs = SHL_E_LIB_RECORDS_FULL; break; }
            shl_chunk_span c = { (uint32_t)p, (uint32_t)idx, (uint32_t)(p + idx), (uint32_t)size };
            out->chunk[used + nch++] = c;
            bytes += size; last = (uint32_t)(idx + size);
            p += idx + (int64_t)size;                                           /* :295-301 */
        }
        out->fstatus[i] = fs; out->n_chunks[i] = nch; out->body_end[i] = (uint64_t)p;
        out->body_bytes[i] = bytes; out->last_call[i] = last;
        used += nch;
    }
    *out->chunk_used = used;
    return 0;
}

/* One CHUNK_HEADER_RX.match + _parse_chunk_size at the start of each window (HttpParser.py:253-261):
 * hdr_len = len(match.group()) or -1 (no match); size = int(size,16); ok = 0 when int() raises. */
int orc_chunk_header(const uint8_t* arena, const uint64_t* off, const uint32_t* len, uint64_t n,
                     int32_t* hdr_len, uint64_t* size, uint8_t* ok) {
    for (uint64_t i = 0; i < n; i++) {
        int64_t a = 0, b = 0;
        int64_t idx = chunk_header_match(arena + off[i], 0, (int64_t)len[i], &a, &b);
        hdr_len[i] = (int32_t)idx; size[i] = 0; ok[i] = 0;
        if (idx >= 0) ok[i] = (uint8_t)py_int16(arena + off[i] + a, b - a, &size[i]);
    }
    return 0;
}

/* ------------------------------------------------------------------ MD5 (RFC 1321) */
static const uint32_t K[64] = {
    0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501,
    0x698098d8, 0x8b44f7af, 0xffff5bb1, 0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821,
    0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa, 0xd62f105d, 0x02441453, 0xd8a1e681, 0xe7d3fbc8,
    0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8, 0x676f02d9, 0x8d2a4c8a,
    0xfffa3942, 0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70,
    0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05, 0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665,
    0xf4292244, 0x432aff97, 0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d, 0x85845dd1,
    0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1, 0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391 };
static const uint8_t S[64] = { 7,12,17,22,7,12,17,22,7,12,17,22,7,12,17,22, 5,9,14,20,5,9,14,20,5,9,14,20,5,9,14,20,
                               4,11,16,23,4,11,16,23,4,11,16,23,4,11,16,23, 6,10,15,21,6,10,15,21,6,10,15,21,6,10,15,21 };
static void md5_block(uint32_t h[4], const uint8_t blk[64]) {
    uint32_t w[16];
    for (int i = 0; i < 16; i++)
        w[i] = (uint32_t)blk[4 * i] | (uint32_t)blk[4 * i + 1] << 8 | (uint32_t)blk[4 * i + 2] << 16 | (uint32_t)blk[4 * i + 3] << 24;
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
    for (int i = 0; i < 64; i++) {
        uint32_t f; int g;
        if (i < 16) { f = (b & c) | (~b & d); g = i; }
        else if (i < 32) { f = (d & b) | (~d & c); g = (5 * i + 1) & 15; }
        else if (i < 48) { f = b ^ c ^ d; g = (3 * i + 5) & 15; }
        else { f = c ^ (b | ~d); g = (7 * i) & 15; }
        uint32_t t = a + f + K[i] + w[g];
        a = d; d = c; c = b;
        b = b + ((t << S[i]) | (t >> (32 - S[i])));
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d;
}
void orc_md5(const uint8_t* msg, uint64_t n, uint8_t out[16]) {
    uint32_t h[4] = { 0x67452301, 0xefcdab89, 0x98badcfe, 0x10325476 };
    uint64_t i = 0;
    for (; i + 64 <= n; i += 64) md5_block(h, msg + i);
    uint8_t tail[128]; uint64_t r = n - i;
    memset(tail, 0, sizeof tail);
    memcpy(tail, msg + i, (size_t)r);
    tail[r] = 0x80;
    uint64_t tl = (r < 56) ? 64 : 128, bits = n * 8;
    for (int k = 0; k < 8; k++) tail[tl - 8 + k] = (uint8_t)(bits >> (8 * k));
    md5_block(h, tail);
    if (tl == 128) md5_block(h, tail + 64);
    for (int k = 0; k < 4; k++) for (int b = 0; b < 4; b++) out[4 * k + b] = (uint8_t)(h[k] >> (8 * b));
}

/* ------------------------------------------------------------------ ketama ring
 * libmemcached 0.44 hosts.c update_continuum.
 *   mode 0 ('ketama': True): per server, i in 0..99: label "%s-%u" (port 11211) or "%s:%u-%u";
 *          value = LE32(md5(label)[0:4]).
 *   mode 1 (ketama_weighted): pts = floorf(pct*160/4*N + 1e-10)*4, pct = (float)w/(float)W; per
 *          label i in 0..pts/4-1, 4 points: LE32(md5[4a:4a+4]), a = 0..3.
 * Sort by value; ties keep insertion order (lower server index, lower point index first): the
 * convention DESIGN.md records for libmemcached's qsort with a value-only comparator.          */
typedef struct { uint32_t value, node; uint32_t seq; } ring_pt;
static int ring_cmp(const void* a, const void* b) {
    const ring_pt* x = a; const ring_pt* y = b;
    if (x->value != y->value) return x->value < y->value ? -1 : 1;
    return x->seq < y->seq ? -1 : (x->seq > y->seq);
}
int orc_ring_build(const char* const* hosts, const uint32_t* ports, const uint32_t* weights,
                   uint32_t n_hosts, int mode, uint32_t* values, uint32_t* nodes, uint32_t cap) {
    ring_pt* pts = malloc(sizeof(ring_pt) * ((size_t)n_hosts * 160 * 4 + 16));
    uint32_t np = 0;
    uint64_t W = 0;
    for (uint32_t s = 0; s < n_hosts; s++) W += weights ? weights[s] : 1;
    for (uint32_t s = 0; s < n_hosts; s++) {
        uint32_t reps = 100, per = 1;
        if (mode == SHL_RING_KETAMA_WEIGHTED) {
            float pct = (float)(weights ? weights[s] : 1) / (float)W;
            float t = pct * 160;            /* float * int -> float, as in the C original */
            t = t / 4;
            t = t * (float)n_hosts;
            uint32_t p = (uint32_t)floorf((float)((double)t + 0.0000000001)) * 4;
            reps = p / 4; per = 4;
        }
        for (uint32_t i = 0; i < reps; i++) {
            char label[320]; int ln;
            if (ports[s] == 11211) ln = snprintf(label, sizeof label, "%s-%u", hosts[s], i);
            else ln = snprintf(label, sizeof label, "%s:%u-%u", hosts[s], ports[s], i);
            uint8_t d[16];
            orc_md5((const uint8_t*)label, (uint64_t)ln, d);
            for (uint32_t a = 0; a < per; a++) {
                pts[np].value = (uint32_t)d[4 * a] | (uint32_t)d[4 * a + 1] << 8 | (uint32_t)d[4 * a + 2] << 16 | (uint32_t)d[4 * a + 3] << 24;
                pts[np].node = s; pts[np].seq = np; np++;
            }
        }
    }
    qsort(pts, np, sizeof *pts, ring_cmp);
    if (np > cap) { free(pts); return -(int)np; }
    for (uint32_t i = 0; i < np; i++) { values[i] = pts[i].value; nodes[i] = pts[i].node; }
    free(pts);
    return (int)np;
}

/* libmemcached hash.c dispatch_host: 1 server -> 0; else first point with value >= h, wrapping. */
int orc_place(const uint8_t* arena, const uint64_t* key_off, const uint32_t* key_len, uint64_t n,
              const uint32_t* values, const uint32_t* nodes, uint32_t n_points, uint32_t n_hosts,
              const shl_place_out* out) {
    for (uint64_t i = 0; i < n; i++) {
        uint8_t d[16];
        orc_md5(arena + key_off[i], key_len[i], d);
        uint32_t h = (uint32_t)d[0] | (uint32_t)d[1] << 8 | (uint32_t)d[2] << 16 | (uint32_t)d[3] << 24;
        uint32_t lo = 0, hi = n_points;
        while (lo < hi) { uint32_t mid = lo + (hi - lo) / 2; if (values[mid] < h) lo = mid + 1; else hi = mid; }
        if (lo == n_points) lo = 0;
        uint32_t node = n_hosts == 1 ? 0 : nodes[lo];
        if (out->md5) memcpy(out->md5 + 16 * i, d, 16);
        if (out->hash) out->hash[i] = h;
        if (out->node) out->node[i] = node;
        if (out->hist) out->hist[node]++;
    }
    return 0;
}

/* Body accumulation (HttpParser.py:213-214, 296-297, 340-347), identity content-coding: the bytes
 * _parse_body has written once the driver loop is over.  Bodies are laid out back to back, each
 * start rounded up to 16 bytes (the device allocates in the same granule, in arbitrary order). */
int orc_body(const uint8_t* arena, const uint64_t* off, const uint32_t* len, uint64_t n,
             const shl_parse_out* p, const shl_frame_out* f, const shl_body_out* out) {
    uint64_t used = 0;
    for (uint64_t i = 0; i < n; i++) {
        const uint8_t* m = arena + off[i];
        uint32_t st = p->status[i];
        uint64_t blen = 0;
        out->body_off[i] = used;
        if ((st & SHL_F_HEADERS_COMPLETE) && SHL_CODE(st) < 16) {
            int no_call = (st & SHL_F_IS_REQUEST) && (st & SHL_F_FIRST_LINE_DONE) &&
                          (span_eq(m, 0, p->sp1[i], "GET") || span_eq(m, 0, p->sp1[i], "HEAD"));
            if (!no_call) {
                if (st & SHL_F_CHUNKED) {
                    if (f->fstatus[i] != 0xff) {
                        for (uint32_t k = 0; k < f->n_chunks[i]; k++) {
                            shl_chunk_span c = f->chunk[f->chunk_base[i] + k];
                            if (used + blen + c.data_len <= out->body_cap) memcpy(out->body + used + blen, m + c.data_off, c.data_len);
                            blen += c.data_len;
                        }
                    }
                } else if ((st & SHL_F_MESSAGE_COMPLETE) && !(st & SHL_F_CLEN_WIDE)) {
                    int64_t body_off = (int64_t)p->first_line_len[i] + p->headers_len[i];
                    int64_t avail = (int64_t)len[i] - body_off, cl = p->content_len[i];
                    int64_t take = cl >= 0 ? (cl < avail ? cl : avail) : (avail + cl > 0 ? avail + cl : 0);
                    if (used + (uint64_t)take <= out->body_cap) memcpy(out->body + used, m + body_off, (size_t)take);
                    blen = (uint64_t)take;
                }
            }
        }
        out->body_len[i] = blen;
        used += (blen + 15) & ~(uint64_t)15;
    }
    *out->body_used = used;
    return 0;
}
