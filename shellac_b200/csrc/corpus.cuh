// corpus.cuh — counter-based synthetic request corpora (BASELINE.json configs 2, 4, 5; recipe in
// SURVEY.md section 8d).  Message i is a pure function of (seed, i): the same code compiles for the
// host and for the device, so a 100 M-request corpus can be produced in HBM and any slice of it
// re-created on the CPU for the oracle.
//
// kind 0: GET request, 8..20 header lines drawn without replacement from 32 real header names
//         (mixed case), URL '/' + 8..96 chars of [A-Za-z0-9/_.-] (30 % with a ?k=v(&k=v){0..3}
//         query), HTTP/1.1 (5 % HTTP/1.0), value bytes 0x21..0x7e with single internal spaces,
//         never ": ", never empty; total length ~ N(512, 64) clipped to [256, 1024].
// kind 1: the URL alone (keys-only variant for the remap sweep, config 5).
#pragma once
#include <cstdint>

namespace shl {

#if defined(__CUDACC__)
#define SHL_HD __host__ __device__ __forceinline__
#else
#define SHL_HD inline
#endif

SHL_HD uint64_t mix64(uint64_t z) {
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}
struct Rng {
    uint64_t s;
    SHL_HD explicit Rng(uint64_t seed, uint64_t i) : s(mix64(seed ^ mix64(i))) {}
    SHL_HD uint32_t next() { s += 0x9e3779b97f4a7c15ull; return (uint32_t)(mix64(s) >> 32); }
    SHL_HD uint32_t below(uint32_t n) { return (uint32_t)(((uint64_t)next() * n) >> 32); }
};

// name table as one string + offsets (works in host and device code without static init)
SHL_HD const char* hdr_names() {
    return "Host\0User-Agent\0Accept\0Accept-Language\0Accept-Encoding\0Accept-Charset\0Connection\0"
           "Keep-Alive\0Cookie\0Referer\0Cache-Control\0Pragma\0DNT\0X-Forwarded-For\0X-Requested-With\0"
           "If-Modified-Since\0If-None-Match\0Via\0Origin\0Upgrade-Insecure-Requests\0TE\0Range\0"
           "X-Request-ID\0x-trace-id\0AUTHORIZATION\0From\0Max-Forwards\0Warning\0X-Real-IP\0"
           "accept-datetime\0X-Csrf-Token\0Forwarded\0";
}
SHL_HD int name_at(int k, const char*& p) {           // k-th name: pointer and length
    const char* s = hdr_names();
    for (int i = 0; i < k; i++) { while (*s) s++; s++; }
    p = s; int n = 0; while (s[n]) n++; return n;
}
SHL_HD uint8_t url_char(uint32_t r) {
    const char* set = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789/_.-";
    return (uint8_t)set[r % 66u];
}

struct Emit {
    uint8_t* out; uint32_t n;
    SHL_HD void put(uint8_t c) { if (out) out[n] = c; n++; }
    SHL_HD void puts(const char* s, int len) { for (int i = 0; i < len; i++) put((uint8_t)s[i]); }
};

SHL_HD void gen_url(Rng& g, Emit& e) {
    e.put('/');
    const uint32_t ul = 8u + g.below(89u);
    for (uint32_t i = 0; i < ul; i++) e.put(url_char(g.next()));
    if (g.below(100u) < 30u) {
        const uint32_t nq = 1u + g.below(4u);
        for (uint32_t q = 0; q < nq; q++) {
            e.put(q ? '&' : '?');
            e.put((uint8_t)('a' + g.below(26u))); e.put('=');
            const uint32_t vl = 1u + g.below(8u);
            for (uint32_t i = 0; i < vl; i++) e.put(url_char(g.next() % 62u));
        }
    }
}

// Writes message i at out (or only measures it when out == nullptr); returns its length.
SHL_HD uint32_t gen_message(uint64_t seed, uint64_t i, int kind, uint8_t* out) {
    Rng g(seed, i);
    Emit e; e.out = out; e.n = 0;
    if (kind == 1) { gen_url(g, e); return e.n; }
    // target total length ~ N(512, 64): Irwin-Hall of 4 uniforms has sd sqrt(4/12); scale to 64
    int t = 0;
    for (int k = 0; k < 4; k++) t += (int)g.below(1024u);
    int target = 512 + (int)(((long long)(t - 2046) * 64) / 591);
    target = target < 256 ? 256 : (target > 1024 ? 1024 : target);
    e.puts("GET ", 4);
    gen_url(g, e);
    e.puts(g.below(100u) < 5u ? " HTTP/1.0\r\n" : " HTTP/1.1\r\n", 11);
    const int H = 8 + (int)g.below(13u);
    const uint32_t start = g.below(32u), step = 1u + 2u * g.below(16u);   // odd step: a permutation of 32
    // bytes the header names + separators will take, so values can share what is left
    int fixed = (int)e.n + 2;
    for (int j = 0; j < H; j++) { const char* p; fixed += name_at((int)((start + (uint32_t)j * step) & 31u), p) + 4; }
    int room = target - fixed;
    if (room < H) room = H;
    // all structural draws first (value lengths), so measuring a message never touches its bytes
    int vls[20];
    int body = 0;
    for (int j = 0; j < H; j++) {
        int vl;
        if (j == H - 1) vl = room;
        else {
            const int share = room / (H - j);
            vl = share / 2 + (int)g.below((uint32_t)(share > 0 ? share : 1)) + 1;
            if (vl > room - (H - 1 - j)) vl = room - (H - 1 - j);
        }
        if (vl < 1) vl = 1;
        room -= vl;
        vls[j] = vl; body += vl;
    }
    if (!out) return (uint32_t)(fixed + body);
    for (int j = 0; j < H; j++) {
        const char* p;
        const int nl = name_at((int)((start + (uint32_t)j * step) & 31u), p);
        e.puts(p, nl); e.put(':'); e.put(' ');
        const int vl = vls[j];
        uint8_t prev = 'x';
        for (int k = 0; k < vl; k++) {
            uint8_t c = (uint8_t)(0x21u + g.below(94u));
            const bool interior = k > 0 && k < vl - 1;
            if (interior && prev != ' ' && prev != ':' && g.below(12u) == 0u) c = ' ';
            out[e.n++] = c;
            prev = c;
        }
        e.put('\r'); e.put('\n');
    }
    e.put('\r'); e.put('\n');
    return e.n;
}

}  // namespace shl
