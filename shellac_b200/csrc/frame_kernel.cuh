// frame_kernel.cuh — K2: chunked-body framing, and the single-header primitive.
//
// Replaces HttpParser._parse_chunked / _parse_chunk_size and CHUNK_HEADER_RX
// (/root/reference/src/python/shellac/server/HttpParser.py:39-40, 234-301) for a whole message
// (buf_len == 0 at every chunk header, so the stale-buf_len quirk of :296 cannot arise here).
// The walk is a pointer chase (each header says where the next one is), so it is latency-bound,
// not bandwidth-bound: bodies are skipped by arithmetic and never read.  One thread per message,
// two walks: count, allocate the message's chunk records with one atomic, emit.
#pragma once
#include "shl_device.cuh"

namespace shl {

// CHUNK_HEADER_RX = (\r\n)?[a-z0-9]+(;[a-z0-9]+="?[a-z0-9\-_]+"?)?\r\n , IGNORECASE, anchored.
// Returns the match length at m[p:L) or -1; [a,b) is the size run.
__device__ __forceinline__ int64_t chunk_header_match(const uint8_t* m, int64_t p, int64_t L, int64_t& a, int64_t& b) {
    int64_t i = p;
    if (i + 1 < L && ldb(m + i) == '\r' && ldb(m + i + 1) == '\n') i += 2;
    a = i;
    while (i < L && is_alnum(ldb(m + i))) i++;
    if (i == a) return -1;
    b = i;
    if (i + 1 < L && ldb(m + i) == '\r' && ldb(m + i + 1) == '\n') return i + 2 - p;
    if (i < L && ldb(m + i) == ';') {
        int64_t j = i + 1, k = j;
        while (j < L && is_alnum(ldb(m + j))) j++;
        if (j == k || j >= L || ldb(m + j) != '=') return -1;
        j++;
        if (j < L && ldb(m + j) == '"') j++;
        k = j;
        while (j < L) { const uint32_t c = ldb(m + j); if (!(is_alnum(c) || c == '-' || c == '_')) break; j++; }
        if (j == k) return -1;
        if (j < L && ldb(m + j) == '"') j++;
        if (j + 1 < L && ldb(m + j) == '\r' && ldb(m + j + 1) == '\n') return j + 2 - p;
    }
    return -1;
}

// int(size, 16) (HttpParser.py:234-241): optional 0x/0X then hex digits; false = ValueError.
// Saturates at 2^63: such a chunk can never be satisfied by a < 2 GiB message.
__device__ __forceinline__ bool py_int16(const uint8_t* s, int64_t n, unsigned long long& v) {
    int64_t i = 0;
    if (n >= 2 && ldb(s) == '0' && (ldb(s + 1) | 32u) == 'x') i = 2;
    if (i == n) return false;
    unsigned long long acc = 0;
    for (; i < n; i++) {
        const int h = hexval(ldb(s + i));
        if (h < 0) return false;
        if (acc >> 59) acc = 1ull << 63; else acc = acc * 16ull + (unsigned long long)h;
        if (acc > (1ull << 63)) acc = 1ull << 63;
    }
    v = acc;
    return true;
}

struct FrameArgs {
    const uint8_t* arena; const uint64_t* off; const uint32_t* len; uint64_t n;
    const uint32_t* status; const uint32_t* first_line_len; const uint32_t* headers_len;
    shl_frame_out out;
};

// The common chunk header, [CRLF] 1-8 hex digits CRLF, matched and converted from ONE 16-byte
// window held in registers (three independent aligned loads, so a chunk costs one memory round
// trip instead of a dozen dependent byte loads; the walk is a pointer chase and nothing else).
// true: *idx = match length, *size = int(digits, 16).  false: not this shape (extension,
// longer run, "0x" prefix, bad digit, no match, window past the message) - the caller then runs
// the byte-wise restatement of the regex, which decides.
__device__ __forceinline__ bool chunk_header_fast(const uint8_t* m, int64_t p, int64_t L, int64_t* idx, unsigned long long* size) {
    if (p + 16 > L) return false;
    const uint8_t* s = m + p;
    const unsigned long long* q = reinterpret_cast<const unsigned long long*>(reinterpret_cast<uintptr_t>(s) & ~(uintptr_t)7);
    const int sh = (int)(reinterpret_cast<uintptr_t>(s) & 7) * 8;
    const unsigned long long q0 = __ldg(q), q1 = __ldg(q + 1), q2 = __ldg(q + 2);
    unsigned long long lo = sh ? (q0 >> sh) | (q1 << (64 - sh)) : q0;      // bytes 0..7 of the window
    unsigned long long hi = sh ? (q1 >> sh) | (q2 << (64 - sh)) : q1;      // bytes 8..15
    int used = 0;
    if ((lo & 0xffffull) == 0x0a0dull) {                                    // optional leading CRLF
        lo = (lo >> 16) | (hi << 48); hi >>= 16; used = 2;
    }
    unsigned long long acc = 0; int nd = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const uint32_t c = (uint32_t)(lo >> (8 * k)) & 0xffu;
        const int h = hexval(c);
        if (nd == k) { if (h >= 0) { acc = acc * 16ull + (unsigned long long)h; nd = k + 1; } }
    }
    if (nd == 0) return false;
    // the two bytes after the run: CRLF, else something the exact matcher has to look at
    const unsigned long long tail = nd == 8 ? hi : (lo >> (8 * nd)) | (hi << (64 - 8 * nd));
    if ((tail & 0xffffull) != 0x0a0dull) return false;
    if (nd >= 2 && (lo & 0xffull) == '0' && ((lo >> 8) & 0xdfull) == 'X') return false;   // "0x..." cannot be: 'x' is not hex; kept for clarity
    *idx = used + nd + 2; *size = acc;
    return true;
}

struct Walk { uint32_t fs, nch, last; int64_t end; unsigned long long bytes; };

template <bool kEmit>
__device__ __forceinline__ Walk walk_chunks(const uint8_t* m, int64_t L, int64_t p, shl_chunk_span* dst, uint32_t cap) {
    Walk w; w.fs = SHL_S_INCOMPLETE_BODY; w.nch = 0; w.last = 0; w.bytes = 0;
    while (p < L) {
        int64_t idx;
        unsigned long long size;
        if (!chunk_header_fast(m, p, L, &idx, &size)) {
            int64_t a, b;
            idx = chunk_header_match(m, p, L, a, b);
            if (idx < 0) { w.last = (uint32_t)(L - p); p = L; break; }             // :255-256
            if (!py_int16(m + a, b - a, size)) { w.fs = SHL_E_CHUNK_HEX; break; }   // :239-241
        }
        const int64_t left = L - (p + idx);
        if (size == 0ull) {                                                          // :268-287
            if (left >= 2 && ldb(m + p + idx) == '\r' && ldb(m + p + idx + 1) == '\n') {
                w.last = (uint32_t)(idx + 2); p += idx + 2; w.fs = SHL_S_OK;
            } else {
                p += idx; w.last = (uint32_t)idx;
                if (p < L) { w.last = (uint32_t)(L - p); p = L; }
            }
            break;
        }
        if ((unsigned long long)left < size) { w.last = (uint32_t)(idx + left); p = L; break; }   // :291-294
        if (kEmit && w.nch < cap) {
            shl_chunk_span c;
            c.hdr_off = (uint32_t)p; c.hdr_len = (uint32_t)idx; c.data_off = (uint32_t)(p + idx); c.data_len = (uint32_t)size;
            dst[w.nch] = c;
        }
        w.nch++;
        w.bytes += size; w.last = (uint32_t)(idx + (int64_t)size);
        p += idx + (int64_t)size;                                                    // :295-301
    }
    w.end = p;
    return w;
}

__global__ void __launch_bounds__(128) k_frame(const FrameArgs a) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const uint32_t st = a.status[i];
    uint32_t fs = 0xffu, nch = 0, last = 0, base = 0; uint64_t end = 0, bytes = 0;
    if ((st & SHL_F_CHUNKED) && (st & SHL_F_HEADERS_COMPLETE) && !(st & SHL_F_MESSAGE_COMPLETE) && (st & 0xffu) < 16u) {
        const uint8_t* m = a.arena + a.off[i];
        const int64_t L = a.len[i], p0 = (int64_t)a.first_line_len[i] + a.headers_len[i];
        Walk w = walk_chunks<false>(m, L, p0, nullptr, 0);
        fs = w.fs; nch = w.nch; last = w.last; end = (uint64_t)w.end; bytes = w.bytes;
        if (nch) {
            const unsigned long long b = atomicAdd(reinterpret_cast<unsigned long long*>(a.out.chunk_used), (unsigned long long)nch);
            base = (uint32_t)b;
            if (b + nch <= a.out.chunk_cap) walk_chunks<true>(m, L, p0, a.out.chunk + b, nch);
            else { fs = SHL_E_LIB_RECORDS_FULL; nch = 0; }
        }
    }
    a.out.fstatus[i] = fs; a.out.n_chunks[i] = nch; a.out.chunk_base[i] = base;
    a.out.body_end[i] = end; a.out.body_bytes[i] = bytes; a.out.last_call[i] = last;
}

// One CHUNK_HEADER_RX.match + _parse_chunk_size at the start of each window (HttpParser.py:253-261),
// the primitive the incremental facade needs.
__global__ void __launch_bounds__(128) k_chunk_header(const uint8_t* arena, const uint64_t* off, const uint32_t* len,
                                                      uint64_t n, int32_t* hdr_len, uint64_t* size, uint8_t* ok) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int64_t a = 0, b = 0;
    const uint8_t* m = arena + off[i];
    const int64_t idx = chunk_header_match(m, 0, (int64_t)len[i], a, b);
    unsigned long long v = 0; bool good = false;
    if (idx >= 0) good = py_int16(m + a, b - a, v);
    hdr_len[i] = (int32_t)idx; size[i] = good ? v : 0ull; ok[i] = good ? 1 : 0;
}

// ----------------------------------------------------------------------------------- K3+K4
// key -> MD5 -> LE32 -> continuum lower-bound -> node (+ per-node histogram), one key per thread.
struct PlaceArgs {
    const uint8_t* arena; const uint64_t* key_off; const uint32_t* key_len; uint64_t n;
    const uint32_t* ring_values; const uint32_t* ring_nodes; const uint32_t* ring_dir; uint32_t n_points, n_hosts;
    shl_place_out out;
};

__global__ void __launch_bounds__(256) k_place(const PlaceArgs a) {
    extern __shared__ uint32_t s_hist[];
    const bool smem_hist = a.out.hist != nullptr && a.n_hosts <= 4096u;
    if (smem_hist) for (uint32_t i = threadIdx.x; i < a.n_hosts; i += blockDim.x) s_hist[i] = 0u;
    __syncthreads();
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < a.n) {
        uint32_t h[4];
        md5_key(a.arena + a.key_off[i], a.key_len[i], h);
        const uint32_t node = ring_lookup(a.ring_values, a.ring_nodes, a.ring_dir, a.n_points, a.n_hosts, h[0]);
        if (a.out.md5) {
            uint32_t* d = reinterpret_cast<uint32_t*>(a.out.md5) + 4 * i;
            d[0] = h[0]; d[1] = h[1]; d[2] = h[2]; d[3] = h[3];
        }
        if (a.out.hash) a.out.hash[i] = h[0];
        if (a.out.node) a.out.node[i] = node;
        if (a.out.hist) {
            if (smem_hist) atomicAdd(&s_hist[node], 1u);
            else atomicAdd(reinterpret_cast<unsigned long long*>(a.out.hist) + node, 1ull);
        }
    }
    __syncthreads();
    if (smem_hist)
        for (uint32_t i = threadIdx.x; i < a.n_hosts; i += blockDim.x)
            if (s_hist[i]) atomicAdd(reinterpret_cast<unsigned long long*>(a.out.hist) + i, (unsigned long long)s_hist[i]);
}

}  // namespace shl

// ----------------------------------------------------------------------------------- N1: body gather
// De-chunk / body extraction: what the reference accumulates with _parse_body as it walks a
// message (HttpParser.py:213-214 for Content-Length bodies, :296-297 per chunk, :340-347).  The
// framing kernels only emit spans; this one copies the bytes those spans select into one
// contiguous, 16-byte-aligned body per message, so `body().read()` can be served from a single
// slice.  Pure data movement: 2 x body_len algorithmic bytes, HBM-bound.  gzip inflate stays on
// the host (zlib), as in the reference.
namespace shl {

struct BodyArgs {
    const uint8_t* arena; const uint64_t* off; const uint32_t* len; uint64_t n;
    const uint32_t* status; const uint32_t* first_line_len; const uint32_t* headers_len; const int64_t* content_len;
    const uint32_t* sp1;                                            // method span end (GET / HEAD have no body call)
    const uint32_t* fstatus; const uint32_t* n_chunks; const uint32_t* chunk_base; const shl_chunk_span* chunk;
    const uint64_t* body_bytes;
    shl_body_out out;
};

// Plan: body length of message i and its slot in the body arena (one atomic per message).
__global__ void __launch_bounds__(128) k_body_plan(const BodyArgs a) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const uint32_t st = a.status[i];
    uint64_t blen = 0;
    const bool hdr_done = (st & SHL_F_HEADERS_COMPLETE) && (st & 0xffu) < 16u;
    if (hdr_done) {
        const uint8_t* m = a.arena + a.off[i];
        const uint32_t s1 = a.sp1[i];
        const bool no_body_call = (st & SHL_F_IS_REQUEST) && (st & SHL_F_FIRST_LINE_DONE) &&
                                  ((s1 == 3u && ldb(m) == 'G' && ldb(m + 1) == 'E' && ldb(m + 2) == 'T') ||
                                   (s1 == 4u && ldb(m) == 'H' && ldb(m + 1) == 'E' && ldb(m + 2) == 'A' && ldb(m + 3) == 'D'));
        if (!no_body_call) {
            if (st & SHL_F_CHUNKED) { if (a.fstatus[i] != 0xffu) blen = a.body_bytes[i]; }
            else if ((st & SHL_F_MESSAGE_COMPLETE) && !(st & SHL_F_CLEN_WIDE)) {       // partial data stays buffered (:209)
                const int64_t avail = (int64_t)a.len[i] - ((int64_t)a.first_line_len[i] + a.headers_len[i]);
                const int64_t cl = a.content_len[i];
                blen = (uint64_t)(cl >= 0 ? (cl < avail ? cl : avail) : (avail + cl > 0 ? avail + cl : 0));   // data[:cl]
            }
        }
    }
    uint64_t boff = 0;
    if (blen) {
        boff = atomicAdd(reinterpret_cast<unsigned long long*>(a.out.body_used), (unsigned long long)((blen + 15ull) & ~15ull));
        if (boff + blen > a.out.body_cap) blen = 0;                 // arena full: caller sees body_used > body_cap
    }
    a.out.body_off[i] = boff; a.out.body_len[i] = blen;
}

// Block-cooperative copy of n bytes, any alignment on either side: 16-byte aligned stores, two
// aligned 16-byte loads per store re-aligned with funnel shifts; scalar bytes at the two edges.
__device__ __forceinline__ void block_copy(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src, uint64_t n) {
    const uint64_t head0 = (16u - (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15)) & 15u;
    const uint64_t head = n < head0 ? n : head0;
    if (threadIdx.x < head) dst[threadIdx.x] = ldb(src + threadIdx.x);
    dst += head; src += head; n -= head;
    const uint64_t nv = n >> 4;
    const uint32_t s = (uint32_t)(reinterpret_cast<uintptr_t>(src) & 15);
    const uint4* sv = reinterpret_cast<const uint4*>(src - s);
    uint4* dv = reinterpret_cast<uint4*>(dst);
    if (s == 0) {
        for (uint64_t v = threadIdx.x; v < nv; v += blockDim.x) dv[v] = __ldg(sv + v);
    } else {
        const int wsel = (int)(s >> 2), sh = (int)(s & 3) * 8;
        for (uint64_t v = threadIdx.x; v < nv; v += blockDim.x) {
            const uint4 a0 = __ldg(sv + v), a1 = __ldg(sv + v + 1);
            const uint32_t w[9] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w, 0u};
            uint32_t o[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                uint32_t lo = 0, hi = 0;
#pragma unroll
                for (int j = 0; j < 4; j++) if (wsel == j) { lo = w[j + k]; hi = w[j + k + 1]; }
                o[k] = __funnelshift_r(lo, hi, sh);
            }
            dv[v] = make_uint4(o[0], o[1], o[2], o[3]);
        }
    }
    const uint64_t done = nv << 4, tail = n - done;
    if (threadIdx.x < tail) dst[done + threadIdx.x] = ldb(src + done + threadIdx.x);
}

// Blocks (i, s), s < kBodySlices, share message i: a body is cut into 32 KiB slices and block s
// copies slices s, s + kBodySlices, ...  (bodies span 4 KiB - 1 MiB: one block per message would
// leave the tail of the grid to a few blocks grinding through megabyte bodies alone).
constexpr int kBodySlices = 8;
constexpr uint64_t kBodySlice = 32768;

__global__ void __launch_bounds__(256) k_body_copy(const BodyArgs a) {
    const uint64_t i = blockIdx.x;
    const uint64_t blen = a.out.body_len[i];
    if ((uint64_t)blockIdx.y * kBodySlice >= blen) return;
    const uint8_t* m = a.arena + a.off[i];
    uint8_t* dst = a.out.body + a.out.body_off[i];
    const uint32_t st = a.status[i];
    if (st & SHL_F_CHUNKED) {
        const shl_chunk_span* c = a.chunk + a.chunk_base[i];
        const uint32_t nc = a.n_chunks[i];
        uint32_t k = 0;
        uint64_t at = 0;                                            // body offset of chunk k
        for (uint64_t lo = (uint64_t)blockIdx.y * kBodySlice; lo < blen; lo += kBodySlices * kBodySlice) {
            const uint64_t hi = min(blen, lo + kBodySlice);
            while (k < nc && at + c[k].data_len <= lo) { at += c[k].data_len; k++; }
            uint32_t kk = k; uint64_t aa = at;
            while (kk < nc && aa < hi) {                            // chunks overlapping [lo, hi)
                const uint64_t dl = c[kk].data_len;
                const uint64_t b0 = max(aa, lo), b1 = min(aa + dl, hi);
                block_copy(dst + b0, m + c[kk].data_off + (b0 - aa), b1 - b0);
                aa += dl; kk++;
            }
        }
    } else {
        const uint8_t* src = m + a.first_line_len[i] + a.headers_len[i];
        for (uint64_t lo = (uint64_t)blockIdx.y * kBodySlice; lo < blen; lo += kBodySlices * kBodySlice)
            block_copy(dst + lo, src + lo, min(blen, lo + kBodySlice) - lo);
    }
}

}  // namespace shl

// ----------------------------------------------------------------------------------- N3: group by node
// After placement the cache client talks to each memcached node separately (pylibmc does this
// behind Server.py:335 / :432).  These two kernels turn node[n] into a permutation grouped by
// node: start[k] .. start[k+1] index the requests of node k in order[] (unplaced ones, node ==
// 0xffffffff, are left out; order inside a node is arbitrary).
namespace shl {

// one block: exclusive scan of the per-node histogram into start[0..n_hosts], cursor[k] = start[k]
__global__ void __launch_bounds__(1024) k_group_scan(const uint64_t* __restrict__ hist, uint32_t n_hosts,
                                                     uint32_t* __restrict__ start, uint32_t* __restrict__ cursor) {
    __shared__ uint32_t warp_sum[32];
    __shared__ uint32_t carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < n_hosts; base += blockDim.x) {
        const uint32_t k = base + threadIdx.x;
        const uint32_t v = k < n_hosts ? (uint32_t)hist[k] : 0u;
        uint32_t inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(SHL_FULL, inc, d); if (lane >= d) inc += t; }
        if (lane == 31) warp_sum[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = warp_sum[lane], wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(SHL_FULL, wi, d); if (lane >= d) wi += t; }
            warp_sum[lane] = wi - w;                               // exclusive over warps
        }
        __syncthreads();
        const uint32_t excl = carry + warp_sum[warp] + inc - v;
        if (k < n_hosts) { start[k] = excl; cursor[k] = excl; }
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) start[n_hosts] = carry;
}

// thread per request; lanes of a warp that hit the same node share one atomic (__match_any_sync)
__global__ void __launch_bounds__(256) k_group_scatter(const uint32_t* __restrict__ node, uint64_t n, uint32_t n_hosts,
                                                       uint32_t* __restrict__ cursor, uint32_t* __restrict__ order) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const uint32_t nd = i < n ? node[i] : 0xffffffffu;
    const bool placed = nd < n_hosts;
    const uint32_t key = placed ? nd : 0xffffffffu;
    const uint32_t peers = __match_any_sync(SHL_FULL, key);
    if (!placed) return;
    const int leader = __ffs(peers) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(&cursor[nd], (uint32_t)__popc(peers));
    base = __shfl_sync(peers, base, leader);
    order[base + __popc(peers & ((1u << lane) - 1u))] = (uint32_t)i;
}

}  // namespace shl

// ----------------------------------------------------------------------------------- N3: memcached binary frames
// What leaves the proxy for each memcached node after placement: the reference's client is built
// with binary=True (pylibmc.Client(..., binary=True), Server.py:81-83) and issues get(key) at
// Server.py:335 and set(key, obj, time=ttl) at :432.  pylibmc / libmemcached are third-party and
// absent here; the frames follow the memcached binary protocol as published (24-byte header:
// magic 0x80, opcode, key length, extras length, data type, vbucket, total body length, opaque,
// CAS; big-endian fields) — restated in oracle/py_port.py (mc_get_frame / mc_set_header).
//   op 0  get : header + key                                   (opcode 0x00)
//   op 1  set : header + extras {flags u32 = 0, expiry u32 = ttl} + key; the VALUE (the re-serialised
//               response, HttpParser.__str__) is appended by the caller: total body counts it.
// Requests come in the node-grouped order of k_group_scatter, so node k's stream is one contiguous
// range of the output: frames[frame_off[start[k]] : frame_off[start[k + 1]]].
namespace shl {

constexpr int kMcScanItems = 2048;              // requests per block of the size scan (256 threads x 8)

struct McArgs {
    const uint8_t* arena; const uint64_t* off; const uint32_t* sp1; const uint32_t* sp2;
    const uint32_t* order; uint64_t n;          // n = requests that have a node (start[n_hosts])
    const uint32_t* value_len;                  // op 1: bytes of the value that will follow each frame, per REQUEST index (may be null: 0)
    uint32_t op, ttl;
    uint64_t* frame_off;                        // [n + 1]
    uint64_t* block_sums;                       // [ceil(n / kMcScanItems) + 1] scratch
    uint8_t* frames; uint64_t frames_cap; uint64_t* frames_used;
};

__device__ __forceinline__ uint32_t mc_frame_bytes(const McArgs& a, uint64_t j) {
    const uint32_t i = a.order[j];
    return 24u + (a.op ? 8u : 0u) + (a.sp2[i] - a.sp1[i] - 1u);
}

// pass 1: per-block exclusive scan of the frame sizes (local offsets) + the block's total
__global__ void __launch_bounds__(256) k_mc_sizes(const McArgs a) {
    __shared__ uint32_t warp_sum[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t base = (uint64_t)blockIdx.x * kMcScanItems + (uint64_t)threadIdx.x * 8u;
    uint32_t v[8], sum = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) { v[k] = base + k < a.n ? mc_frame_bytes(a, base + k) : 0u; sum += v[k]; }
    uint32_t inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t t = __shfl_up_sync(SHL_FULL, inc, d); if (lane >= d) inc += t; }
    if (lane == 31) warp_sum[warp] = inc;
    __syncthreads();
    uint32_t wbase = 0;
    for (int w = 0; w < warp; w++) wbase += warp_sum[w];
    uint32_t run = wbase + inc - sum;
#pragma unroll
    for (int k = 0; k < 8; k++) { if (base + k < a.n) a.frame_off[base + k] = run; run += v[k]; }
    if (threadIdx.x == 255) a.block_sums[blockIdx.x] = run;
}
// pass 2: one block turns the block totals into block bases (exclusive), total at [nblocks]
__global__ void __launch_bounds__(1024) k_mc_block_scan(uint64_t* __restrict__ sums, uint32_t nblocks, uint64_t* __restrict__ used,
                                                        uint64_t* __restrict__ frame_off, uint64_t n) {
    __shared__ unsigned long long part[1024];
    const uint32_t per = (nblocks + 1023u) / 1024u;
    const uint32_t lo = threadIdx.x * per, hi = min(nblocks, lo + per);
    unsigned long long s = 0;
    for (uint32_t b = lo; b < hi; b++) s += sums[b];
    part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) { unsigned long long run = 0; for (int t = 0; t < 1024; t++) { const unsigned long long x = part[t]; part[t] = run; run += x; } *used = run; frame_off[n] = run; }
    __syncthreads();
    unsigned long long run = part[threadIdx.x];
    for (uint32_t b = lo; b < hi; b++) { const unsigned long long x = sums[b]; sums[b] = run; run += x; }
}
// pass 3: one warp per request writes its frame (lanes 0..23 / 0..31 the header + extras, then the key)
__global__ void __launch_bounds__(256) k_mc_emit(const McArgs a) {
    const uint64_t j = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (j >= a.n) return;
    const uint32_t i = a.order[j];
    const uint32_t klen = a.sp2[i] - a.sp1[i] - 1u, ext = a.op ? 8u : 0u;
    const uint64_t at = a.frame_off[j] + a.block_sums[j / kMcScanItems];
    __syncwarp();
    if (lane == 0) a.frame_off[j] = at;
    const uint32_t hdr = 24u + ext;
    if (at + hdr + klen > a.frames_cap) return;
    const uint32_t body = ext + klen + (a.op && a.value_len ? a.value_len[i] : 0u);
    uint8_t* f = a.frames + at;
    if ((uint32_t)lane < hdr) {
        uint32_t b = 0;
        switch (lane) {
            case 0: b = 0x80u; break;                      // magic: request
            case 1: b = a.op ? 0x01u : 0x00u; break;       // opcode: Get / Set
            case 2: b = klen >> 8; break;                  // key length, big-endian
            case 3: b = klen & 0xffu; break;
            case 4: b = ext; break;                        // extras length
            case 8: b = body >> 24; break;                 // total body length
            case 9: b = (body >> 16) & 0xffu; break;
            case 10: b = (body >> 8) & 0xffu; break;
            case 11: b = body & 0xffu; break;
            case 12: b = i >> 24; break;                   // opaque = request index: pairs the reply with the request
            case 13: b = (i >> 16) & 0xffu; break;
            case 14: b = (i >> 8) & 0xffu; break;
            case 15: b = i & 0xffu; break;
            case 28: b = a.ttl >> 24; break;               // set extras: flags (0) then expiry
            case 29: b = (a.ttl >> 16) & 0xffu; break;
            case 30: b = (a.ttl >> 8) & 0xffu; break;
            case 31: b = a.ttl & 0xffu; break;
            default: b = 0; break;                         // data type, vbucket, CAS, flags
        }
        f[lane] = (uint8_t)b;
    }
    const uint8_t* key = a.arena + a.off[i] + a.sp1[i] + 1u;
    for (uint32_t k = (uint32_t)lane; k < klen; k += 32u) f[hdr + k] = (uint8_t)ldb(key + k);
}

}  // namespace shl
