// shl_device.cuh — device helpers shared by the kernels of libshellac_b200.so (sm_100a).
//
// Python-2 `str` semantics restated for bytes on the device: whitespace = {9..13, 32}
// (str.strip / str.isspace in the C locale), ASCII-only lower(), int()/float() literal grammars.
// Reference call sites: HttpParser.py:307 (rstrip), :326 (strip), :329 (value.strip), :330
// (key.lower), :312/:319 (float), :313/:197 (int), :239-241 (int(..., 16)).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "../../include/shellac_b200.h"

#define SHL_FULL 0xffffffffu

namespace shl {

__device__ __forceinline__ bool is_ws(uint32_t c) { return c == 32u || (c - 9u) <= 4u; }
__device__ __forceinline__ bool is_digit(uint32_t c) { return (c - 48u) <= 9u; }
__device__ __forceinline__ bool is_alnum(uint32_t c) { return is_digit(c) || ((c | 32u) - 97u) <= 25u; }
__device__ __forceinline__ uint32_t lower(uint32_t c) { return (c - 65u) <= 25u ? c + 32u : c; }
__device__ __forceinline__ int hexval(uint32_t c) {
    if (is_digit(c)) return (int)c - 48;
    c |= 32u;
    return (c - 97u) <= 5u ? (int)c - 87 : -1;
}

__device__ __forceinline__ uint32_t ldb(const uint8_t* p) { return (uint32_t)__ldg(p); }

// unaligned little-endian 32-bit load (may touch up to 3 bytes past p+4, inside the arena padding)
__device__ __forceinline__ uint32_t ld32u(const uint8_t* p) {
    const uint32_t* a = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)3);
    const int sh = (int)(reinterpret_cast<uintptr_t>(p) & 3) * 8;
    const uint32_t lo = __ldg(a);
    const uint32_t hi = sh ? __ldg(a + 1) : 0u;
    return __funnelshift_r(lo, hi, sh);
}

// unaligned little-endian 64-bit load (may touch up to 7 bytes past p+8, inside the arena padding)
__device__ __forceinline__ uint64_t ld64u(const uint8_t* p) {
    const uint32_t* a = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)3);
    const int sh = (int)(reinterpret_cast<uintptr_t>(p) & 3) * 8;
    const uint32_t w0 = __ldg(a), w1 = __ldg(a + 1), w2 = sh ? __ldg(a + 2) : 0u;
    return (uint64_t)__funnelshift_r(w0, w1, sh) | ((uint64_t)__funnelshift_r(w1, w2, sh) << 32);
}

// ASCII lower() of four packed bytes (bytes >= 0x80 are left alone, like Python 2's str.lower in the C locale)
__device__ __forceinline__ uint32_t lower4(uint32_t x) {
    const uint32_t t = x & 0x7f7f7f7fu;
    const uint32_t ge_A = t + 0x3f3f3f3fu;                 // bit 7 of a byte set iff t >= 'A'
    const uint32_t gt_Z = t + 0x25252525u;                 // ... iff t > 'Z'
    return x | ((ge_A & ~gt_Z & ~x & 0x80808080u) >> 2);
}

// p[0:n] == lit (ci: lower(p[0:n]) == lit, lit lower-case).  ONE out-of-line copy, word at a time:
// the literal compares used to be unrolled byte loops inlined at every call site, a quarter of
// the route kernel's code, which the instruction cache paid for (profiles/r01/v8_notes).
__device__ __noinline__ bool span_eq_words(const uint8_t* p, const char* lit, int n, bool ci) {
    int k = 0;
    for (; k + 4 <= n; k += 4) {
        uint32_t x = ld32u(p + k);
        const uint32_t y = (uint32_t)(uint8_t)lit[k] | ((uint32_t)(uint8_t)lit[k + 1] << 8) |
                           ((uint32_t)(uint8_t)lit[k + 2] << 16) | ((uint32_t)(uint8_t)lit[k + 3] << 24);
        if (ci) x = lower4(x);
        if (x != y) return false;
    }
    for (; k < n; k++) {
        uint32_t c = ldb(p + k);
        if (ci) c = lower(c);
        if (c != (uint32_t)(uint8_t)lit[k]) return false;
    }
    return true;
}
// m[a:b] == lit  (exact)
template <int N>
__device__ __forceinline__ bool span_eq(const uint8_t* m, int64_t a, int64_t b, const char (&lit)[N]) {
    if (b - a != N - 1) return false;
    return span_eq_words(m + a, lit, N - 1, false);
}
// lower(m[a:b]) == lit  (lit is lower-case)
template <int N>
__device__ __forceinline__ bool span_eq_lower(const uint8_t* m, int64_t a, int64_t b, const char (&lit)[N]) {
    if (b - a != N - 1) return false;
    return span_eq_words(m + a, lit, N - 1, true);
}

// float(s) accepts?  [ws] [+-] (digits [. digits*] | . digits) [(e|E) [+-] digits] [ws]  |  inf / infinity / nan
__device__ __noinline__ bool py_float_ok(const uint8_t* s, int64_t n) {
    int64_t i = 0, j = n;
    while (i < j && is_ws(ldb(s + i))) i++;
    while (j > i && is_ws(ldb(s + j - 1))) j--;
    if (i == j) return false;
    uint32_t c = ldb(s + i);
    if (c == '+' || c == '-') i++;
    int64_t r = j - i;
    if (r == 3 && lower(ldb(s + i)) == 'n' && lower(ldb(s + i + 1)) == 'a' && lower(ldb(s + i + 2)) == 'n') return true;
    if (r == 3 || r == 8) {
        const char inf[9] = "infinity";
        bool ok = true;
        for (int64_t k = 0; k < r; k++) if (lower(ldb(s + i + k)) != (uint32_t)inf[k]) { ok = false; break; }
        if (ok) return true;
    }
    int64_t nd = 0;
    while (i < j && is_digit(ldb(s + i))) { i++; nd++; }
    if (i < j && ldb(s + i) == '.') { i++; while (i < j && is_digit(ldb(s + i))) { i++; nd++; } }
    if (nd == 0) return false;
    if (i < j && (ldb(s + i) | 32u) == 'e') {
        i++;
        if (i < j && (ldb(s + i) == '+' || ldb(s + i) == '-')) i++;
        int64_t ne = 0;
        while (i < j && is_digit(ldb(s + i))) { i++; ne++; }
        if (ne == 0) return false;
    }
    return i == j;
}

// int(s): 0 invalid, 1 valid & fits int64, 2 valid but wider than int64
__device__ __noinline__ int py_int10(const uint8_t* s, int64_t n, int64_t* v) {
    int64_t i = 0, j = n;
    while (i < j && is_ws(ldb(s + i))) i++;
    while (j > i && is_ws(ldb(s + j - 1))) j--;
    bool neg = false;
    if (i < j) { uint32_t c = ldb(s + i); if (c == '+' || c == '-') { neg = c == '-'; i++; } }
    if (i == j) return 0;
    unsigned long long acc = 0; bool wide = false;
    for (; i < j; i++) {
        uint32_t c = ldb(s + i);
        if (!is_digit(c)) return 0;
        unsigned long long d = c - 48u;
        if (acc > (0xffffffffffffffffull - d) / 10ull) wide = true; else acc = acc * 10ull + d;
    }
    if (wide) return 2;
    if (!neg && acc > 0x7fffffffffffffffull) return 2;
    if (neg && acc > 0x8000000000000000ull) return 2;
    *v = neg ? (int64_t)(0ull - acc) : (int64_t)acc;
    return 1;
}

// ---------------------------------------------------------------------------------------------
// MD5 (RFC 1321), one key per thread.  Key bytes are fetched as aligned 32-bit words and
// re-aligned with a funnel shift; reads may touch up to 7 bytes past the key inside the arena's
// padding, they are masked off.  Reference: libhashkit md5.c behind pylibmc (Server.py:335,432).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t rotl(uint32_t x, int s) { return __funnelshift_l(x, x, s); }

#define SHL_MD5_F(b, c, d) (((b) & (c)) | (~(b) & (d)))
#define SHL_MD5_G(b, c, d) (((d) & (b)) | (~(d) & (c)))
#define SHL_MD5_H(b, c, d) ((b) ^ (c) ^ (d))
#define SHL_MD5_I(b, c, d) ((c) ^ ((b) | ~(d)))
#define SHL_MD5_STEP(f, a, b, c, d, x, t, s) (a) += f((b), (c), (d)) + (x) + (t); (a) = rotl((a), (s)) + (b);

__device__ __forceinline__ void md5_block(uint32_t h[4], const uint32_t w[16]) {
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
    SHL_MD5_STEP(SHL_MD5_F, a, b, c, d, w[0], 0xd76aa478u, 7)  SHL_MD5_STEP(SHL_MD5_F, d, a, b, c, w[1], 0xe8c7b756u, 12)
    SHL_MD5_STEP(SHL_MD5_F, c, d, a, b, w[2], 0x242070dbu, 17) SHL_MD5_STEP(SHL_MD5_F, b, c, d, a, w[3], 0xc1bdceeeu, 22)
    SHL_MD5_STEP(SHL_MD5_F, a, b, c, d, w[4], 0xf57c0fafu, 7)  SHL_MD5_STEP(SHL_MD5_F, d, a, b, c, w[5], 0x4787c62au, 12)
    SHL_MD5_STEP(SHL_MD5_F, c, d, a, b, w[6], 0xa8304613u, 17) SHL_MD5_STEP(SHL_MD5_F, b, c, d, a, w[7], 0xfd469501u, 22)
    SHL_MD5_STEP(SHL_MD5_F, a, b, c, d, w[8], 0x698098d8u, 7)  SHL_MD5_STEP(SHL_MD5_F, d, a, b, c, w[9], 0x8b44f7afu, 12)
    SHL_MD5_STEP(SHL_MD5_F, c, d, a, b, w[10], 0xffff5bb1u, 17) SHL_MD5_STEP(SHL_MD5_F, b, c, d, a, w[11], 0x895cd7beu, 22)
    SHL_MD5_STEP(SHL_MD5_F, a, b, c, d, w[12], 0x6b901122u, 7) SHL_MD5_STEP(SHL_MD5_F, d, a, b, c, w[13], 0xfd987193u, 12)
    SHL_MD5_STEP(SHL_MD5_F, c, d, a, b, w[14], 0xa679438eu, 17) SHL_MD5_STEP(SHL_MD5_F, b, c, d, a, w[15], 0x49b40821u, 22)
    SHL_MD5_STEP(SHL_MD5_G, a, b, c, d, w[1], 0xf61e2562u, 5)  SHL_MD5_STEP(SHL_MD5_G, d, a, b, c, w[6], 0xc040b340u, 9)
    SHL_MD5_STEP(SHL_MD5_G, c, d, a, b, w[11], 0x265e5a51u, 14) SHL_MD5_STEP(SHL_MD5_G, b, c, d, a, w[0], 0xe9b6c7aau, 20)
    SHL_MD5_STEP(SHL_MD5_G, a, b, c, d, w[5], 0xd62f105du, 5)  SHL_MD5_STEP(SHL_MD5_G, d, a, b, c, w[10], 0x02441453u, 9)
    SHL_MD5_STEP(SHL_MD5_G, c, d, a, b, w[15], 0xd8a1e681u, 14) SHL_MD5_STEP(SHL_MD5_G, b, c, d, a, w[4], 0xe7d3fbc8u, 20)
    SHL_MD5_STEP(SHL_MD5_G, a, b, c, d, w[9], 0x21e1cde6u, 5)  SHL_MD5_STEP(SHL_MD5_G, d, a, b, c, w[14], 0xc33707d6u, 9)
    SHL_MD5_STEP(SHL_MD5_G, c, d, a, b, w[3], 0xf4d50d87u, 14) SHL_MD5_STEP(SHL_MD5_G, b, c, d, a, w[8], 0x455a14edu, 20)
    SHL_MD5_STEP(SHL_MD5_G, a, b, c, d, w[13], 0xa9e3e905u, 5) SHL_MD5_STEP(SHL_MD5_G, d, a, b, c, w[2], 0xfcefa3f8u, 9)
    SHL_MD5_STEP(SHL_MD5_G, c, d, a, b, w[7], 0x676f02d9u, 14) SHL_MD5_STEP(SHL_MD5_G, b, c, d, a, w[12], 0x8d2a4c8au, 20)
    SHL_MD5_STEP(SHL_MD5_H, a, b, c, d, w[5], 0xfffa3942u, 4)  SHL_MD5_STEP(SHL_MD5_H, d, a, b, c, w[8], 0x8771f681u, 11)
    SHL_MD5_STEP(SHL_MD5_H, c, d, a, b, w[11], 0x6d9d6122u, 16) SHL_MD5_STEP(SHL_MD5_H, b, c, d, a, w[14], 0xfde5380cu, 23)
    SHL_MD5_STEP(SHL_MD5_H, a, b, c, d, w[1], 0xa4beea44u, 4)  SHL_MD5_STEP(SHL_MD5_H, d, a, b, c, w[4], 0x4bdecfa9u, 11)
    SHL_MD5_STEP(SHL_MD5_H, c, d, a, b, w[7], 0xf6bb4b60u, 16) SHL_MD5_STEP(SHL_MD5_H, b, c, d, a, w[10], 0xbebfbc70u, 23)
    SHL_MD5_STEP(SHL_MD5_H, a, b, c, d, w[13], 0x289b7ec6u, 4) SHL_MD5_STEP(SHL_MD5_H, d, a, b, c, w[0], 0xeaa127fau, 11)
    SHL_MD5_STEP(SHL_MD5_H, c, d, a, b, w[3], 0xd4ef3085u, 16) SHL_MD5_STEP(SHL_MD5_H, b, c, d, a, w[6], 0x04881d05u, 23)
    SHL_MD5_STEP(SHL_MD5_H, a, b, c, d, w[9], 0xd9d4d039u, 4)  SHL_MD5_STEP(SHL_MD5_H, d, a, b, c, w[12], 0xe6db99e5u, 11)
    SHL_MD5_STEP(SHL_MD5_H, c, d, a, b, w[15], 0x1fa27cf8u, 16) SHL_MD5_STEP(SHL_MD5_H, b, c, d, a, w[2], 0xc4ac5665u, 23)
    SHL_MD5_STEP(SHL_MD5_I, a, b, c, d, w[0], 0xf4292244u, 6)  SHL_MD5_STEP(SHL_MD5_I, d, a, b, c, w[7], 0x432aff97u, 10)
    SHL_MD5_STEP(SHL_MD5_I, c, d, a, b, w[14], 0xab9423a7u, 15) SHL_MD5_STEP(SHL_MD5_I, b, c, d, a, w[5], 0xfc93a039u, 21)
    SHL_MD5_STEP(SHL_MD5_I, a, b, c, d, w[12], 0x655b59c3u, 6) SHL_MD5_STEP(SHL_MD5_I, d, a, b, c, w[3], 0x8f0ccc92u, 10)
    SHL_MD5_STEP(SHL_MD5_I, c, d, a, b, w[10], 0xffeff47du, 15) SHL_MD5_STEP(SHL_MD5_I, b, c, d, a, w[1], 0x85845dd1u, 21)
    SHL_MD5_STEP(SHL_MD5_I, a, b, c, d, w[8], 0x6fa87e4fu, 6)  SHL_MD5_STEP(SHL_MD5_I, d, a, b, c, w[15], 0xfe2ce6e0u, 10)
    SHL_MD5_STEP(SHL_MD5_I, c, d, a, b, w[6], 0xa3014314u, 15) SHL_MD5_STEP(SHL_MD5_I, b, c, d, a, w[13], 0x4e0811a1u, 21)
    SHL_MD5_STEP(SHL_MD5_I, a, b, c, d, w[4], 0xf7537e82u, 6)  SHL_MD5_STEP(SHL_MD5_I, d, a, b, c, w[11], 0xbd3af235u, 10)
    SHL_MD5_STEP(SHL_MD5_I, c, d, a, b, w[2], 0x2ad7d2bbu, 15) SHL_MD5_STEP(SHL_MD5_I, b, c, d, a, w[9], 0xeb86d391u, 21)
    h[0] += a; h[1] += b; h[2] += c; h[3] += d;
}

// 0x80 in some byte <=> x holds a zero byte (exact as a yes / no answer)
__device__ __forceinline__ uint32_t has_zero_byte(uint32_t x) { return (x - 0x01010101u) & ~x & 0x80808080u; }

// digest of key[0:n) into h[4] (little-endian words == digest bytes 0..15).  kSpaces: also report
// whether the key holds a ' ' (the route kernel takes the URL for granted between the first space
// of the line and " HTTP/1.x"; a space inside it means split(' ', 2) cut elsewhere, HttpParser.py:307).
template <bool kSpaces>
__device__ __forceinline__ uint32_t md5_key_t(const uint8_t* key, uint32_t n, uint32_t h[4]) {
    h[0] = 0x67452301u; h[1] = 0xefcdab89u; h[2] = 0x98badcfeu; h[3] = 0x10325476u;
    const uint32_t* ap = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(key) & ~(uintptr_t)3);
    const uint32_t o = (uint32_t)(reinterpret_cast<uintptr_t>(key) & 3);
    const int sh = (int)o * 8;
    const uint32_t nblk = (n + 8u) / 64u + 1u;
    // aligned word j holds key bytes [4j-o, 4j-o+4); only words that hold a key byte are touched.
    // Straight-line per word (lanes hold keys of different lengths): key word gi is data below
    // wn, the word with the 0x80 pad at wn, zero above.
    const uint32_t nw = (n + o + 3u) >> 2, wn = n >> 2;
    const uint32_t r8 = (n & 3u) * 8u, km = (1u << r8) - 1u, pd = 0x80u << r8;
    uint32_t spaces = 0u;
    uint32_t lo = (n > 0u) ? __ldg(ap) : 0u;
    for (uint32_t blk = 0; blk < nblk; blk++) {
        uint32_t w[16];
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const uint32_t gi = blk * 16u + (uint32_t)i;        // key word index
            const uint32_t hi = (gi + 1u < nw) ? __ldg(ap + gi + 1u) : 0u;
            uint32_t x = __funnelshift_r(lo, hi, sh);
            lo = hi;
            x = gi < wn ? x : (gi == wn ? (x & km) : 0u);
            if (kSpaces) spaces |= has_zero_byte(x ^ 0x20202020u);
            if (gi == wn) x |= pd;
            if (blk == nblk - 1u) { if (i == 14) x = n << 3; if (i == 15) x = n >> 29; }
            w[i] = x;
        }
        md5_block(h, w);
    }
    return spaces;
}
__device__ __forceinline__ void md5_key(const uint8_t* key, uint32_t n, uint32_t h[4]) { md5_key_t<false>(key, n, h); }

// does p[0:n) hold a ' '?  (the same question without the digest, for the parse-only kernel)
__device__ __forceinline__ bool span_has_space(const uint8_t* p, uint32_t n) {
    const uint32_t* ap = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)3);
    const uint32_t o = (uint32_t)(reinterpret_cast<uintptr_t>(p) & 3);
    const uint32_t nw = (n + o + 3u) >> 2;
    uint32_t spaces = 0u;
    for (uint32_t j = 0; j < nw; j++) {
        uint32_t x = __ldg(ap + j);
        if (j == 0u) x &= 0xffffffffu << (8u * o);
        if (j == nw - 1u) { const uint32_t e = (n + o) & 3u; if (e) x &= 0xffffffffu >> (8u * (4u - e)); }
        spaces |= has_zero_byte(x ^ 0x20202020u);
    }
    return n != 0u && spaces != 0u;
}

// lower_bound over the sorted continuum + wrap (libmemcached hash.c dispatch_host).
// `dir` (optional) is a 4097-entry directory on the top 12 hash bits: dir[b] = index of the first
// point whose value >= b << 20.  It bounds the search to one bucket, so the ~log2(n_points)
// dependent loads of a plain binary search become two independent directory loads plus the few
// steps a bucket needs (0-1 at 1 600 points, ~5 at 102 400).
constexpr int kRingDirBits = 12;
__device__ __forceinline__ uint32_t ring_lookup(const uint32_t* __restrict__ values, const uint32_t* __restrict__ nodes,
                                                const uint32_t* __restrict__ dir, uint32_t n_points, uint32_t n_hosts, uint32_t h) {
    if (n_hosts == 1u) return 0u;
    uint32_t lo = 0, hi = n_points;
    if (dir) { const uint32_t b = h >> (32 - kRingDirBits); lo = __ldg(dir + b); hi = __ldg(dir + b + 1); }
    while (lo < hi) {
        uint32_t mid = lo + ((hi - lo) >> 1);
        if (__ldg(values + mid) < h) lo = mid + 1; else hi = mid;
    }
    if (lo == n_points) lo = 0;
    return __ldg(nodes + lo);
}

}  // namespace shl
