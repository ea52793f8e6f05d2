// bm_emu.cpp — TEST INFRASTRUCTURE.  Runs the product's bitmap walker (shellac_b200/csrc/
// bitmap_walk.cuh, the same source the route kernel includes) on the CPU, over bitmaps built by a
// scalar loop, so that its framing logic can be checked against the oracle without a GPU
// (tests/test_bitmap_walk.py).  The classifier that produces the bitmaps on the device is fp16 SIMD
// and only runs there; this covers everything downstream of it.
#include <cstdint>
#include <cstring>
#include <vector>
#include "../shellac_b200/csrc/bitmap_walk.cuh"

extern "C" {

// Messages i0..i0+g-1 of (arena, off, len) are treated as ONE group lying in index order; the span
// starts at the 32-byte floor of the first message.  Per message: idx, T, n_c, n_hdr, err (0 / 20 /
// 21 as SHL_E_HDR_NO_SEP / _MULTI_SEP) and, for the first n_hdr lines, (line start, separator, end)
// relative to the message in lines[3 * (line_base[i] + k) ...].
int bm_emu_group(const uint8_t* arena, const uint64_t* off, const uint32_t* len, uint32_t g,
                 int32_t* idx, int32_t* T, uint32_t* n_c, uint32_t* n_hdr, uint32_t* err,
                 uint32_t* line_base, uint32_t* lines, uint32_t lines_cap, uint32_t* lines_used) {
    if (!g) return 0;
    const uint64_t a0 = off[0] & ~31ull;
    uint64_t end = 0;
    for (uint32_t i = 0; i < g; i++) {
        if (i && off[i] < off[i - 1] + len[i - 1]) return -1;           // not a span
        if (off[i] + len[i] > end) end = off[i] + len[i];
    }
    const uint32_t nbytes = (uint32_t)(end - a0);
    const uint32_t nw = (nbytes + 1) / 32 + 2;
    std::vector<shl::bm2> bm(nw + 3);
    for (auto& w : bm) w.x = w.y = 0u;
    for (uint32_t p = 0; p + 1 < nbytes; p++) {
        const uint8_t c0 = arena[a0 + p], c1 = arena[a0 + p + 1];
        const uint32_t q = p + 1;
        if (c0 == '\r' && c1 == '\n') bm[q >> 5].x |= 1u << (q & 31);
        if (c0 == ':' && c1 == ' ') bm[q >> 5].y |= 1u << (q & 31);
    }
    uint32_t used = *lines_used;
    for (uint32_t i = 0; i < g; i++) {
        const uint32_t s0 = (uint32_t)(off[i] - a0);
        shl::BmWalk w;
        shl::bm_walk(bm.data(), s0, (int)len[i], w);
        idx[i] = w.idx; T[i] = w.T; n_c[i] = w.n_c; n_hdr[i] = 0; err[i] = 0; line_base[i] = used;
        if (w.T < 0) continue;
        uint32_t nh = w.n_c - 1u;
        if (w.bad) {
            bool no_sep = false;
            nh = shl::bm_offender(bm.data(), s0, w, &no_sep);
            err[i] = no_sep ? 20u : 21u;
        }
        n_hdr[i] = nh;
        shl::BmCursor cc, cs;
        shl::bmc_init(cc, bm.data(), false, w.q_idx);
        shl::bmc_init(cs, bm.data(), true, w.q_idx);
        shl::EvCursor ec;
        shl::evc_init(ec, bm.data(), w.q_idx);
        uint32_t prev = w.q_idx;
        for (uint32_t k = 0; k < nh; k++) {
            const uint32_t sq = shl::bmc_next(cs, bm.data(), true), eq = shl::bmc_next(cc, bm.data(), false);
            if (shl::evc_pop(ec, bm.data()) != sq || shl::evc_pop(ec, bm.data()) != eq) return -3;   // both cursors must agree
            if (used >= lines_cap) return -2;
            lines[3 * used + 0] = prev + 1u - s0;
            lines[3 * used + 1] = sq - 1u - s0;
            lines[3 * used + 2] = eq - 1u - s0;
            used++;
            prev = eq;
        }
    }
    *lines_used = used;
    return 0;
}

}  // extern "C"
