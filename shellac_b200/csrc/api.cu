// api.cu — the C ABI of libshellac_b200.so (declared in include/shellac_b200.h).
//
// Host side of the boundary: a context that owns the device arenas, the SoA output block, the
// stream and the ketama ring; the H2D -> kernels -> D2H choreography of the *_batch entry points;
// and the ring builder (libmemcached hosts.c update_continuum, restated).  No PyTorch, no CPU
// parsing or hashing of request data: if CUDA is unavailable every entry point fails.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <dlfcn.h>
#include <nccl.h>      // types and prototypes only: the library is bound at run time (see NcclApi)

#include "parse_kernel.cuh"
#include "frame_kernel.cuh"
#include "corpus.cuh"

namespace {

std::string g_last_error;

struct DevParse {             // device SoA block (one allocation)
    shl_parse_out o{};
    void* block = nullptr;
};

}  // namespace

struct shl_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[16]{};
    uint64_t max_msgs = 0, max_bytes = 0, max_records = 0;
    uint8_t* d_arena = nullptr; uint64_t* d_off = nullptr; uint32_t* d_len = nullptr;
    uint64_t* d_koff = nullptr; uint32_t* d_klen = nullptr;
    shl_parse_out dp{}; shl_frame_out df{}; shl_place_out dq{};
    shl_parse_out cur_dp{};               // where the LAST *_batch parse left its per-message fields (dp, or the small-batch block)
    const uint32_t* cur_node = nullptr; const uint64_t* cur_hist = nullptr;   // ... and the last placement its nodes / histogram
    uint64_t* d_counters = nullptr;       // [0]=hdr_used [1]=chunk_used [2..66)=code_hist
    uint64_t* h_counters = nullptr;       // pinned mirror
    int32_t* d_ch_len = nullptr; uint64_t* d_ch_size = nullptr; uint8_t* d_ch_ok = nullptr;
    // ring
    uint32_t* d_ring_values = nullptr; uint32_t* d_ring_nodes = nullptr; uint32_t* d_ring_dir = nullptr;
    std::vector<uint32_t> ring_values, ring_nodes;
    uint32_t n_points = 0, n_hosts = 0;
    uint64_t* d_hist = nullptr; uint32_t hist_cap = 0;
    uint32_t* d_gstart = nullptr; uint32_t* d_gcursor = nullptr; uint32_t* d_gorder = nullptr; uint32_t gcap = 0;   // grouping (lazy)
    uint8_t* d_body = nullptr; uint64_t* d_body_off = nullptr; uint64_t* d_body_len = nullptr;   // body gather (lazy)
    // counters shared between GPUs: [placement histogram n_hosts | outcome codes 64], two buffers so
    // that the all-reduce of batch k (comm stream) overlaps the kernel of batch k + 1 (ctx stream)
    static constexpr int kStatBufs = 4;   // a buffer is reused four batches later: the ranks of a job need not run in lock-step
    uint64_t* d_stats[kStatBufs] = {}; uint64_t* h_stats = nullptr;   // h_stats: pinned, kStatBufs x stats_len
    uint32_t stats_len = 0; int stat_k = 0, stat_last = -1;
    cudaStream_t comm_stream = nullptr;
    cudaEvent_t ev_stats_ready[kStatBufs]{}, ev_stats_done[kStatBufs]{};
    ncclComm_t comm = nullptr; int comm_rank = 0, comm_world = 1;
    uint64_t* d_scratch = nullptr; uint64_t* h_scratch = nullptr;             // 64 words for shl_comm_allreduce_u64
    uint64_t cur_n = 0;                   // messages of the last *_batch parse (frame_batch checks it)
    shl_msg_rec* d_crec = nullptr;        // compact records (lazy)
    uint64_t* d_mc_off = nullptr; uint64_t* d_mc_sums = nullptr; uint8_t* d_mc_frames = nullptr; uint32_t* d_mc_vlen = nullptr;   // memcached frames (lazy)
    uint64_t mc_frames_cap = 0;
    uint8_t* d_lkeys = nullptr; uint32_t* d_lkey_base = nullptr;   // lower-cased header names (lazy; d_counters[2] counts the bytes)
    int force_scalar = 0;
    int sm_count = 148, grid_waves = 16;
    uint64_t launches = 0;
    // pipelined host-buffer path: chunks round-robin over these streams so H2D, kernel and D2H overlap
    static constexpr int kPipe = 4, kMaxChunks = 64;
    int chunk_msgs = 65536, max_chunks = 8;     // SHL_CHUNK_MSGS / SHL_MAX_CHUNKS: messages per chunk, chunks per batch
    cudaStream_t pipe[kPipe]{};
    cudaEvent_t ev_fork = nullptr, ev_join[kPipe]{}, ev_chunk[kMaxChunks]{};
    uint64_t* d_chunk_used = nullptr;     // [kMaxChunks]
    uint64_t* h_chunk_used = nullptr;     // pinned mirror
    int pipeline = 1;                     // SHL_PIPELINE=0 disables the chunked path
    // small-batch path (one reactor tick): every result of a call lands in ONE device block that
    // comes back with ONE copy into a pinned mirror, and is handed out from there on the host
    static constexpr uint64_t kSmallMsgs = 4096;             // beyond this the host-side hand-out costs more than the copies it saves
    uint8_t* d_small = nullptr; uint8_t* h_small = nullptr; uint64_t small_bytes = 0;
    int small_path = 1;                   // SHL_SMALL=0 disables it
    std::string err;
};

namespace {

int fail(shl_ctx* c, int code, const std::string& msg) {
    if (c) c->err = msg;
    g_last_error = msg;
    return code;
}
#define CU(call)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess)                                                                \
            return fail(ctx, SHL_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

// NCCL is bound with dlopen when the first communicator is made, not with a DT_NEEDED entry: a
// process that loads this library and then imports PyTorch would otherwise be handed the system's
// libnccl.so.2 where torch expects the newer one it bundles under the same soname (and fails with
// an undefined symbol).  If NCCL is already in the process (torch's), dlopen returns that copy.
struct NcclApi {
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    void* handle = nullptr;
    std::string why;
    bool load() {
        if (handle) return true;
        const char* name = getenv("SHL_NCCL_LIB");
        handle = dlopen(name && name[0] ? name : "libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
        if (!handle) { why = std::string("dlopen(libnccl.so.2): ") + dlerror(); return false; }
#define SHL_NCCL_SYM(f) f = reinterpret_cast<decltype(f)>(dlsym(handle, "nccl" #f)); if (!f) { why = "libnccl.so.2 lacks nccl" #f; handle = nullptr; return false; }
        SHL_NCCL_SYM(GetUniqueId) SHL_NCCL_SYM(CommInitRank) SHL_NCCL_SYM(AllReduce) SHL_NCCL_SYM(CommDestroy) SHL_NCCL_SYM(GetErrorString)
#undef SHL_NCCL_SYM
        return true;
    }
};
NcclApi g_nccl;

// Makes `device` current for the scope and puts the caller's device back on the way out.
struct DeviceGuard {
    int prev = -1; bool switched = false;
    cudaError_t enter(int device) {
        cudaError_t e = cudaGetDevice(&prev);
        if (e != cudaSuccess) return e;
        if (prev != device) { e = cudaSetDevice(device); switched = e == cudaSuccess; }
        return e;
    }
    ~DeviceGuard() { if (switched) cudaSetDevice(prev); }
};

template <typename T>
cudaError_t dalloc(T** p, uint64_t n) { return cudaMalloc(reinterpret_cast<void**>(p), std::max<uint64_t>(n, 1) * sizeof(T)); }

// ---------------------------------------------------------------------------- host MD5 (ring labels only)
struct Md5 {
    static uint32_t rol(uint32_t x, int s) { return (x << s) | (x >> (32 - s)); }
    static void block(uint32_t h[4], const uint8_t* p) {
        static const uint32_t K[64] = {
            0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501,
            0x698098d8, 0x8b44f7af, 0xffff5bb1, 0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821,
            0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa, 0xd62f105d, 0x02441453, 0xd8a1e681, 0xe7d3fbc8,
            0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8, 0x676f02d9, 0x8d2a4c8a,
            0xfffa3942, 0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70,
            0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05, 0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665,
            0xf4292244, 0x432aff97, 0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d, 0x85845dd1,
            0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1, 0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391};
        static const int S[16] = {7, 12, 17, 22, 5, 9, 14, 20, 4, 11, 16, 23, 6, 10, 15, 21};
        uint32_t w[16];
        for (int i = 0; i < 16; i++) memcpy(&w[i], p + 4 * i, 4);
        uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
        for (int i = 0; i < 64; i++) {
            uint32_t f; int g;
            switch (i >> 4) {
                case 0: f = (b & c) | (~b & d); g = i; break;
                case 1: f = (d & b) | (~d & c); g = (5 * i + 1) & 15; break;
                case 2: f = b ^ c ^ d; g = (3 * i + 5) & 15; break;
                default: f = c ^ (b | ~d); g = (7 * i) & 15; break;
            }
            const uint32_t t = a + f + K[i] + w[g];
            a = d; d = c; c = b;
            b += rol(t, S[(i >> 4) * 4 + (i & 3)]);
        }
        h[0] += a; h[1] += b; h[2] += c; h[3] += d;
    }
    static void digest(const std::string& s, uint32_t h[4]) {
        h[0] = 0x67452301; h[1] = 0xefcdab89; h[2] = 0x98badcfe; h[3] = 0x10325476;
        std::string m = s;
        m.push_back((char)0x80);
        while (m.size() % 64 != 56) m.push_back('\0');
        const uint64_t bits = (uint64_t)s.size() * 8;
        for (int k = 0; k < 8; k++) m.push_back((char)(bits >> (8 * k)));
        for (size_t i = 0; i < m.size(); i += 64) block(h, reinterpret_cast<const uint8_t*>(m.data()) + i);
    }
};

shl_parse_out slice_parse(void* block, uint64_t n, shl_hdr_span* hdr, uint64_t cap, uint64_t* used) {
    // 12 arrays, each n elements, 8-byte aligned strides
    uint8_t* p = static_cast<uint8_t*>(block);
    const uint64_t s4 = (n * 4 + 15) & ~15ull, s8 = (n * 8 + 15) & ~15ull;
    shl_parse_out o{};
    o.status = (uint32_t*)p; p += s4; o.first_line_len = (uint32_t*)p; p += s4; o.headers_len = (uint32_t*)p; p += s4;
    o.sp1 = (uint32_t*)p; p += s4; o.sp2 = (uint32_t*)p; p += s4; o.fl_end = (uint32_t*)p; p += s4;
    o.slash = (uint32_t*)p; p += s4; o.status_code = (int32_t*)p; p += s4; o.n_headers = (uint32_t*)p; p += s4;
    o.hdr_base = (uint32_t*)p; p += s4; o.content_len = (int64_t*)p; p += s8;
    o.hdr = hdr; o.hdr_cap = cap; o.hdr_used = used;
    return o;
}
uint64_t parse_block_bytes(uint64_t n) { return 10 * ((n * 4 + 15) & ~15ull) + ((n * 8 + 15) & ~15ull); }

}  // namespace

extern "C" {

void shl_ctx_destroy(shl_ctx* ctx);

int shl_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

int shl_device_mem_info(int device, uint64_t* free_bytes, uint64_t* total_bytes) {
    shl_ctx* ctx = nullptr;
    DeviceGuard guard;
    CU(guard.enter(device));
    size_t f = 0, t = 0;
    CU(cudaMemGetInfo(&f, &t));
    if (free_bytes) *free_bytes = f;
    if (total_bytes) *total_bytes = t;
    return SHL_OK;
}

int shl_device_pci_bus_id(int device, char* buf, int len) {
    if (!buf || len < 13) return -1;
    return cudaDeviceGetPCIBusId(buf, len, device) == cudaSuccess ? 0 : -2;
}

const char* shl_last_error(shl_ctx* ctx) { return ctx ? ctx->err.c_str() : g_last_error.c_str(); }

static int stats_alloc(shl_ctx* ctx, uint32_t n_hosts) {
    const uint32_t len = n_hosts + SHL_N_CODES;
    for (int k = 0; k < shl_ctx::kStatBufs; k++) { cudaFree(ctx->d_stats[k]); ctx->d_stats[k] = nullptr; }
    if (ctx->h_stats) { cudaFreeHost(ctx->h_stats); ctx->h_stats = nullptr; }
    for (int k = 0; k < shl_ctx::kStatBufs; k++) { CU(dalloc(&ctx->d_stats[k], len)); CU(cudaMemset(ctx->d_stats[k], 0, (size_t)len * 8)); }
    CU(cudaHostAlloc(reinterpret_cast<void**>(&ctx->h_stats), shl_ctx::kStatBufs * (size_t)len * 8, cudaHostAllocDefault));
    memset(ctx->h_stats, 0, shl_ctx::kStatBufs * (size_t)len * 8);
    ctx->stats_len = len; ctx->stat_k = 0; ctx->stat_last = -1;
    return SHL_OK;
}
static uint64_t* cur_codes(shl_ctx* ctx) { return ctx->d_stats[ctx->stat_k] + (ctx->stats_len - SHL_N_CODES); }

static int ctx_setup(shl_ctx* ctx, int device) {
    CU(cudaSetDevice(device));
    CU(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    CU(cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device));
    { const char* gw = getenv("SHL_GRID_WAVES"); if (gw && atoi(gw) > 0) ctx->grid_waves = atoi(gw); }
    for (auto& e : ctx->ev) CU(cudaEventCreate(&e));
    const uint64_t n = ctx->max_msgs;
    CU(dalloc(&ctx->d_arena, ctx->max_bytes + 64));
    CU(dalloc(&ctx->d_off, n)); CU(dalloc(&ctx->d_len, n));
    CU(dalloc(&ctx->d_koff, n)); CU(dalloc(&ctx->d_klen, n));
    CU(dalloc(&ctx->d_counters, 66));
    CU(cudaHostAlloc(reinterpret_cast<void**>(&ctx->h_counters), 66 * sizeof(uint64_t), cudaHostAllocDefault));
    if (int rc = stats_alloc(ctx, 0)) return rc;
    CU(cudaStreamCreateWithFlags(&ctx->comm_stream, cudaStreamNonBlocking));
    for (auto& e : ctx->ev_stats_ready) CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto& e : ctx->ev_stats_done) CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CU(dalloc(&ctx->d_scratch, 64));
    CU(cudaHostAlloc(reinterpret_cast<void**>(&ctx->h_scratch), 64 * sizeof(uint64_t), cudaHostAllocDefault));
    void* blk = nullptr;
    CU(cudaMalloc(&blk, parse_block_bytes(n)));
    shl_hdr_span* hdr = nullptr; CU(dalloc(&hdr, ctx->max_records));
    ctx->dp = slice_parse(blk, n, hdr, ctx->max_records, ctx->d_counters + 0);
    ctx->cur_dp = ctx->dp;
    shl_frame_out& f = ctx->df;
    CU(dalloc(&f.fstatus, n)); CU(dalloc(&f.n_chunks, n)); CU(dalloc(&f.chunk_base, n));
    CU(dalloc(&f.body_end, n)); CU(dalloc(&f.body_bytes, n)); CU(dalloc(&f.last_call, n));
    CU(dalloc(&f.chunk, ctx->max_records)); f.chunk_cap = ctx->max_records; f.chunk_used = ctx->d_counters + 1;
    shl_place_out& q = ctx->dq;
    CU(dalloc(&q.md5, 16 * n)); CU(dalloc(&q.hash, n)); CU(dalloc(&q.node, n));
    CU(dalloc(&ctx->d_ch_len, n)); CU(dalloc(&ctx->d_ch_size, n)); CU(dalloc(&ctx->d_ch_ok, n));
    { const char* cv = getenv("SHL_CARVEOUT");      // percent of the SM's 228 KiB given to shared memory (default: the driver's choice)
      if (cv && atoi(cv) > 0) {
          CU(cudaFuncSetAttribute(shl::k_route<true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, atoi(cv)));
          CU(cudaFuncSetAttribute(shl::k_route<false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, atoi(cv)));
      } }
    CU(cudaFuncSetAttribute(shl::k_route<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384));
    CU(cudaFuncSetAttribute(shl::k_route<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384));
    CU(cudaFuncSetAttribute(shl::k_route<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384));
    CU(cudaFuncSetAttribute(shl::k_route<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384));
    for (auto& st : ctx->pipe) CU(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
    for (auto& e : ctx->ev_join) CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto& e : ctx->ev_chunk) CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CU(dalloc(&ctx->d_chunk_used, shl_ctx::kMaxChunks));
    CU(cudaHostAlloc(reinterpret_cast<void**>(&ctx->h_chunk_used), shl_ctx::kMaxChunks * sizeof(uint64_t), cudaHostAllocDefault));
    { const char* pl = getenv("SHL_PIPELINE"); if (pl && pl[0] == '0') ctx->pipeline = 0; }
    { const char* cm = getenv("SHL_CHUNK_MSGS"); if (cm && atoi(cm) >= 1024) ctx->chunk_msgs = atoi(cm); }
    { const char* mc = getenv("SHL_MAX_CHUNKS"); if (mc && atoi(mc) >= 2) ctx->max_chunks = atoi(mc); }
    { const char* sm = getenv("SHL_SMALL"); if (sm && sm[0] == '0') ctx->small_path = 0; }
    const char* fs = getenv("SHL_FORCE_SCALAR");
    ctx->force_scalar = (fs && fs[0] == '1') ? 1 : 0;
    return SHL_OK;
}

int shl_ctx_create(int device, uint64_t max_msgs, uint64_t max_bytes, uint64_t max_records, shl_ctx** out) {
    if (!out) return fail(nullptr, SHL_ERR_ARG, "shl_ctx_create: out is null");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(nullptr, SHL_ERR_NO_DEVICE, "no CUDA device: libshellac_b200 has no CPU fallback");
    if (device < 0 || device >= ndev) return fail(nullptr, SHL_ERR_ARG, "shl_ctx_create: bad device index");
    shl_ctx* ctx = new shl_ctx();
    ctx->device = device;
    ctx->max_msgs = std::max<uint64_t>(max_msgs, 1);
    ctx->max_bytes = std::max<uint64_t>(max_bytes, 64);
    ctx->max_records = max_records ? max_records : std::max<uint64_t>(ctx->max_bytes / 16, 1024);
    if (const int rc = ctx_setup(ctx, device)) {           // frees whatever was allocated before the failure
        const std::string why = ctx->err;
        shl_ctx_destroy(ctx);
        return fail(nullptr, rc, why);
    }
    *out = ctx;
    return SHL_OK;
}

void shl_ctx_destroy(shl_ctx* ctx) {
    if (!ctx) return;
    DeviceGuard guard; guard.enter(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    cudaFree(ctx->d_arena); cudaFree(ctx->d_off); cudaFree(ctx->d_len); cudaFree(ctx->d_koff); cudaFree(ctx->d_klen);
    cudaFree(ctx->dp.status); cudaFree(ctx->dp.hdr);
    cudaFree(ctx->df.fstatus); cudaFree(ctx->df.n_chunks); cudaFree(ctx->df.chunk_base); cudaFree(ctx->df.body_end);
    cudaFree(ctx->df.body_bytes); cudaFree(ctx->df.last_call); cudaFree(ctx->df.chunk);
    cudaFree(ctx->dq.md5); cudaFree(ctx->dq.hash); cudaFree(ctx->dq.node);
    cudaFree(ctx->d_ch_len); cudaFree(ctx->d_ch_size); cudaFree(ctx->d_ch_ok);
    if (ctx->h_counters) cudaFreeHost(ctx->h_counters);
    cudaFree(ctx->d_counters);
    if (ctx->comm) g_nccl.CommDestroy(ctx->comm);
    for (int k = 0; k < shl_ctx::kStatBufs; k++) cudaFree(ctx->d_stats[k]);
    if (ctx->h_stats) cudaFreeHost(ctx->h_stats);
    cudaFree(ctx->d_scratch); if (ctx->h_scratch) cudaFreeHost(ctx->h_scratch);
    if (ctx->comm_stream) { cudaStreamSynchronize(ctx->comm_stream); cudaStreamDestroy(ctx->comm_stream); }
    for (auto& e : ctx->ev_stats_ready) if (e) cudaEventDestroy(e);
    for (auto& e : ctx->ev_stats_done) if (e) cudaEventDestroy(e);
    cudaFree(ctx->d_ring_values); cudaFree(ctx->d_ring_nodes); cudaFree(ctx->d_ring_dir); cudaFree(ctx->d_hist);
    cudaFree(ctx->d_body); cudaFree(ctx->d_body_off); cudaFree(ctx->d_body_len);
    cudaFree(ctx->d_lkeys); cudaFree(ctx->d_lkey_base); cudaFree(ctx->d_crec);
    cudaFree(ctx->d_mc_off); cudaFree(ctx->d_mc_sums); cudaFree(ctx->d_mc_frames); cudaFree(ctx->d_mc_vlen);
    cudaFree(ctx->d_gstart); cudaFree(ctx->d_gcursor); cudaFree(ctx->d_gorder);
    for (auto& e : ctx->ev) if (e) cudaEventDestroy(e);
    for (auto& st : ctx->pipe) if (st) { cudaStreamSynchronize(st); cudaStreamDestroy(st); }
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    for (auto& e : ctx->ev_join) if (e) cudaEventDestroy(e);
    for (auto& e : ctx->ev_chunk) if (e) cudaEventDestroy(e);
    cudaFree(ctx->d_chunk_used); if (ctx->h_chunk_used) cudaFreeHost(ctx->h_chunk_used);
    cudaFree(ctx->d_small); if (ctx->h_small) cudaFreeHost(ctx->h_small);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int shl_sync(shl_ctx* ctx) {
    if (!ctx) return SHL_ERR_ARG;
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaGetLastError());
    return SHL_OK;
}

int shl_pinned_alloc(void** ptr, uint64_t bytes) {
    shl_ctx* ctx = nullptr;
    if (!ptr) return SHL_ERR_ARG;
    CU(cudaHostAlloc(ptr, std::max<uint64_t>(bytes, 1), cudaHostAllocDefault));
    return SHL_OK;
}
int shl_pinned_free(void* ptr) { shl_ctx* ctx = nullptr; CU(cudaFreeHost(ptr)); return SHL_OK; }
int shl_device_alloc(shl_ctx* ctx, void** dptr, uint64_t bytes) {
    if (!ctx || !dptr) return SHL_ERR_ARG;
    CU(cudaSetDevice(ctx->device));
    CU(cudaMalloc(dptr, std::max<uint64_t>(bytes, 1)));
    return SHL_OK;
}
int shl_device_free(shl_ctx* ctx, void* dptr) { if (!ctx) return SHL_ERR_ARG; CU(cudaSetDevice(ctx->device)); CU(cudaFree(dptr)); return SHL_OK; }
int shl_memcpy_h2d(shl_ctx* ctx, void* dst, const void* src, uint64_t bytes) {
    if (!ctx) return SHL_ERR_ARG;
    CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return SHL_OK;
}
int shl_memcpy_d2h(shl_ctx* ctx, void* dst, const void* src, uint64_t bytes) {
    if (!ctx) return SHL_ERR_ARG;
    CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return SHL_OK;
}
int shl_memcpy_d2d(shl_ctx* ctx, void* dst, const void* src, uint64_t bytes) {
    if (!ctx) return SHL_ERR_ARG;
    CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, ctx->stream));
    return SHL_OK;
}
int shl_memset_dev(shl_ctx* ctx, void* dst, int value, uint64_t bytes) {
    if (!ctx) return SHL_ERR_ARG;
    CU(cudaMemsetAsync(dst, value, bytes, ctx->stream));
    return SHL_OK;
}
int shl_event_record(shl_ctx* ctx, int slot) {
    if (!ctx || slot < 0 || slot >= 16) return SHL_ERR_ARG;
    CU(cudaEventRecord(ctx->ev[slot], ctx->stream));
    return SHL_OK;
}
int shl_event_elapsed_ms(shl_ctx* ctx, int a, int b, float* ms) {
    if (!ctx || !ms || a < 0 || a >= 16 || b < 0 || b >= 16) return SHL_ERR_ARG;
    CU(cudaEventSynchronize(ctx->ev[b]));
    CU(cudaEventElapsedTime(ms, ctx->ev[a], ctx->ev[b]));
    return SHL_OK;
}
void* shl_stream(shl_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
uint64_t shl_launch_count(shl_ctx* ctx) { return ctx ? ctx->launches : 0; }

// ------------------------------------------------------------------------------------ ring
// libmemcached 0.44 hosts.c update_continuum, restated (PARITY UNPINNED: see DESIGN.md).
int shl_ring_build(shl_ctx* ctx, const char* const* hosts, const uint32_t* ports, const uint32_t* weights,
                   uint32_t n_hosts, int mode) {
    if (!ctx || !hosts || !ports || n_hosts == 0) return fail(ctx, SHL_ERR_ARG, "shl_ring_build: bad argument");
    struct Pt { uint32_t value, node; };
    std::vector<Pt> pts;
    pts.reserve((size_t)n_hosts * 160);
    uint64_t W = 0;
    for (uint32_t s = 0; s < n_hosts; s++) W += weights ? weights[s] : 1;
    for (uint32_t s = 0; s < n_hosts; s++) {
        uint32_t reps = 100, per = 1;                                   // MEMCACHED_POINTS_PER_SERVER
        if (mode == SHL_RING_KETAMA_WEIGHTED) {                         // MEMCACHED_POINTS_PER_SERVER_KETAMA = 160
            const float pct = (float)(weights ? weights[s] : 1) / (float)W;
            float t = pct * 160;
            t = t / 4;
            t = t * (float)n_hosts;
            reps = (uint32_t)floorf((float)((double)t + 0.0000000001));
            per = 4;
        }
        for (uint32_t i = 0; i < reps; i++) {
            char label[400];
            int ln;
            if (ports[s] == 11211) ln = snprintf(label, sizeof label, "%s-%u", hosts[s], i);
            else ln = snprintf(label, sizeof label, "%s:%u-%u", hosts[s], ports[s], i);
            uint32_t h[4];
            Md5::digest(std::string(label, (size_t)ln), h);
            for (uint32_t k = 0; k < per; k++) pts.push_back({h[k], s});
        }
    }
    // value-only comparator; equal values keep insertion order (lower server index first)
    std::stable_sort(pts.begin(), pts.end(), [](const Pt& a, const Pt& b) { return a.value < b.value; });
    ctx->ring_values.resize(pts.size()); ctx->ring_nodes.resize(pts.size());
    for (size_t i = 0; i < pts.size(); i++) { ctx->ring_values[i] = pts[i].value; ctx->ring_nodes[i] = pts[i].node; }
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    cudaFree(ctx->d_ring_values); cudaFree(ctx->d_ring_nodes);
    ctx->d_ring_values = ctx->d_ring_nodes = nullptr;
    CU(dalloc(&ctx->d_ring_values, pts.size())); CU(dalloc(&ctx->d_ring_nodes, pts.size()));
    CU(cudaMemcpy(ctx->d_ring_values, ctx->ring_values.data(), pts.size() * 4, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(ctx->d_ring_nodes, ctx->ring_nodes.data(), pts.size() * 4, cudaMemcpyHostToDevice));
    {   // directory on the top kRingDirBits hash bits: dir[b] = first point with value >= b << (32 - bits)
        const uint32_t nb = 1u << shl::kRingDirBits;
        std::vector<uint32_t> dir(nb + 1);
        size_t j = 0;
        for (uint32_t b = 0; b < nb; b++) {
            const uint32_t base = b << (32 - shl::kRingDirBits);
            while (j < pts.size() && ctx->ring_values[j] < base) j++;
            dir[b] = (uint32_t)j;
        }
        dir[nb] = (uint32_t)pts.size();
        if (!ctx->d_ring_dir) CU(dalloc(&ctx->d_ring_dir, nb + 1));
        CU(cudaMemcpy(ctx->d_ring_dir, dir.data(), (nb + 1) * 4, cudaMemcpyHostToDevice));
    }
    if (ctx->stats_len != n_hosts + SHL_N_CODES) {
        CU(cudaStreamSynchronize(ctx->comm_stream));
        if (int rc = stats_alloc(ctx, n_hosts)) return rc;
    }
    ctx->n_points = (uint32_t)pts.size(); ctx->n_hosts = n_hosts;
    if (ctx->hist_cap < n_hosts) {
        cudaFree(ctx->d_hist); ctx->d_hist = nullptr;
        CU(dalloc(&ctx->d_hist, n_hosts));
        ctx->hist_cap = n_hosts;
    }
    return SHL_OK;
}
int shl_ring_size(shl_ctx* ctx, uint32_t* n_points, uint32_t* n_hosts) {
    if (!ctx) return SHL_ERR_ARG;
    if (n_points) *n_points = ctx->n_points;
    if (n_hosts) *n_hosts = ctx->n_hosts;
    return SHL_OK;
}
int shl_ring_export(shl_ctx* ctx, uint32_t* values, uint32_t* nodes, uint32_t cap) {
    if (!ctx || cap < ctx->n_points) return fail(ctx, SHL_ERR_ARG, "shl_ring_export: capacity too small");
    if (values) memcpy(values, ctx->ring_values.data(), (size_t)ctx->n_points * 4);
    if (nodes) memcpy(nodes, ctx->ring_nodes.data(), (size_t)ctx->n_points * 4);
    return SHL_OK;
}

// ------------------------------------------------------------------------------------ device entry points
static int check_arena(shl_ctx* ctx, const void* arena) {
    if ((reinterpret_cast<uintptr_t>(arena) & 31) != 0)
        return fail(ctx, SHL_ERR_ARG, "arena base must be 32-byte aligned (and padded by 64 readable bytes)");
    return SHL_OK;
}

static int launch_route(shl_ctx* ctx, const uint8_t* d_arena, const uint64_t* d_off, const uint32_t* d_len, uint64_t n,
                        const shl_parse_out* d_out, const shl_place_out* d_q, cudaStream_t stream = nullptr, uint32_t bias = 0,
                        shl_msg_rec* d_crec = nullptr) {
    if (!stream && ctx) stream = ctx->stream;
    if (!ctx || !d_arena || !d_off || !d_len || !d_out) return fail(ctx, SHL_ERR_ARG, "null pointer");
    CU(cudaSetDevice(ctx->device));
    if (int rc = check_arena(ctx, d_arena)) return rc;
    if (d_q && !ctx->d_ring_values) return fail(ctx, SHL_ERR_NO_RING, "shl_ring_build() has not been called");
    if (n == 0) return SHL_OK;
    shl::RouteArgs a{};
    a.arena = d_arena; a.off = d_off; a.len = d_len; a.n = n; a.out = *d_out;
    a.code_hist = cur_codes(ctx); a.force_scalar = ctx->force_scalar;
    a.k = shl::swar_host_consts();
    a.hdr_base_bias = bias;
    a.crec = d_crec;
    // a warp takes 32 messages when there are enough of them to fill the machine; a reactor-sized
    // batch is spread over more warps (down to one message each) so that the kernel's critical
    // path is one message's worth of rows, not thirty-two
    uint32_t gsz = shl::kGroup;
    while (gsz > 1 && n / gsz < 148ull * 8ull) gsz >>= 1;
    a.group = gsz;
    const uint64_t groups = (n + gsz - 1) / gsz;
    // no more blocks than the device holds at once (times SHL_GRID_WAVES): each warp then loops over groups
    const uint64_t resident = (uint64_t)ctx->sm_count * SHL_OPT_MINBLOCKS * (uint64_t)ctx->grid_waves;
    const unsigned blocks = (unsigned)std::min<uint64_t>((groups + shl::kWarpsPerBlock - 1) / shl::kWarpsPerBlock, resident);
    if (d_q) {
        a.ring_values = ctx->d_ring_values; a.ring_nodes = ctx->d_ring_nodes; a.ring_dir = ctx->d_ring_dir;
        a.n_points = ctx->n_points; a.n_hosts = ctx->n_hosts; a.place = *d_q;
        const size_t sh = (d_q->hist && ctx->n_hosts <= 4096u) ? ctx->n_hosts * 4 : 0;
        if (gsz == shl::kGroup) shl::k_route<true, true><<<blocks, shl::kWarpsPerBlock * 32, sh, stream>>>(a);
        else shl::k_route<true, false><<<blocks, shl::kWarpsPerBlock * 32, sh, stream>>>(a);
    } else {
        if (gsz == shl::kGroup) shl::k_route<false, true><<<blocks, shl::kWarpsPerBlock * 32, 0, stream>>>(a);
        else shl::k_route<false, false><<<blocks, shl::kWarpsPerBlock * 32, 0, stream>>>(a);
    }
    ctx->launches++;
    CU(cudaGetLastError());
    return SHL_OK;
}

int shl_parse_dev(shl_ctx* ctx, const uint8_t* d_arena, const uint64_t* d_off, const uint32_t* d_len, uint64_t n,
                  const shl_parse_out* d_out) {
    return launch_route(ctx, d_arena, d_off, d_len, n, d_out, nullptr);
}
int shl_route_dev(shl_ctx* ctx, const uint8_t* d_arena, const uint64_t* d_off, const uint32_t* d_len, uint64_t n,
                  const shl_parse_out* d_pout, const shl_place_out* d_qout) {
    if (!d_qout) return fail(ctx, SHL_ERR_ARG, "shl_route_dev: qout is null");
    return launch_route(ctx, d_arena, d_off, d_len, n, d_pout, d_qout);
}
int shl_frame_dev(shl_ctx* ctx, const uint8_t* d_arena, const uint64_t* d_off, const uint32_t* d_len, uint64_t n,
                  const uint32_t* d_status, const uint32_t* d_fl, const uint32_t* d_hl, const shl_frame_out* d_out) {
    if (!ctx || !d_arena || !d_off || !d_len || !d_status || !d_fl || !d_hl || !d_out) return fail(ctx, SHL_ERR_ARG, "null pointer");
    if (n == 0) return SHL_OK;
    CU(cudaSetDevice(ctx->device));
    shl::FrameArgs a{};
    a.arena = d_arena; a.off = d_off; a.len = d_len; a.n = n; a.status = d_status; a.first_line_len = d_fl; a.headers_len = d_hl;
    a.out = *d_out;
    shl::k_frame<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(a);
    ctx->launches++;
    CU(cudaGetLastError());
    return SHL_OK;
}
int shl_place_dev(shl_ctx* ctx, const uint8_t* d_arena, const uint64_t* d_key_off, const uint32_t* d_key_len, uint64_t n,
                  const shl_place_out* d_out) {
    if (!ctx || !d_arena || !d_key_off || !d_key_len || !d_out) return fail(ctx, SHL_ERR_ARG, "null pointer");
    if (!ctx->d_ring_values) return fail(ctx, SHL_ERR_NO_RING, "shl_ring_build() has not been called");
    if (n == 0) return SHL_OK;
    CU(cudaSetDevice(ctx->device));
    shl::PlaceArgs a{};
    a.arena = d_arena; a.key_off = d_key_off; a.key_len = d_key_len; a.n = n;
    a.ring_values = ctx->d_ring_values; a.ring_nodes = ctx->d_ring_nodes; a.ring_dir = ctx->d_ring_dir;
    a.n_points = ctx->n_points; a.n_hosts = ctx->n_hosts;
    a.out = *d_out;
    const size_t sh = (d_out->hist && ctx->n_hosts <= 4096u) ? ctx->n_hosts * 4 : 0;
    shl::k_place<<<(unsigned)((n + 255) / 256), 256, sh, ctx->stream>>>(a);
    ctx->launches++;
    CU(cudaGetLastError());
    return SHL_OK;
}

// ------------------------------------------------------------------------------------ host-buffer entry points
static int stage_in(shl_ctx* ctx, const uint8_t* arena, uint64_t arena_bytes, const uint64_t* off, const uint32_t* len,
                    uint64_t n, uint64_t* d_off, uint32_t* d_len) {
    if (!ctx || !off || !len) return fail(ctx, SHL_ERR_ARG, "null pointer");
    if (n > ctx->max_msgs || arena_bytes > ctx->max_bytes)
        return fail(ctx, SHL_ERR_ARG, "batch larger than the capacity given to shl_ctx_create");
    CU(cudaSetDevice(ctx->device));
    if (arena)   // NULL: reuse the bytes the previous *_batch call uploaded
        CU(cudaMemcpyAsync(ctx->d_arena, arena, arena_bytes, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(d_off, off, n * 8, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(d_len, len, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    return SHL_OK;
}
#define D2H(dst, src, bytes) do { if (dst) CU(cudaMemcpyAsync((dst), (src), (bytes), cudaMemcpyDeviceToHost, ctx->stream)); } while (0)

static int parse_out_d2h(shl_ctx* ctx, const shl_parse_out* out, uint64_t n) {
    const shl_parse_out& d = ctx->dp;
    D2H(out->status, d.status, n * 4); D2H(out->first_line_len, d.first_line_len, n * 4);
    D2H(out->headers_len, d.headers_len, n * 4); D2H(out->sp1, d.sp1, n * 4); D2H(out->sp2, d.sp2, n * 4);
    D2H(out->fl_end, d.fl_end, n * 4); D2H(out->slash, d.slash, n * 4); D2H(out->status_code, d.status_code, n * 4);
    D2H(out->n_headers, d.n_headers, n * 4); D2H(out->hdr_base, d.hdr_base, n * 4); D2H(out->content_len, d.content_len, n * 8);
    if (out->hdr || out->hdr_used) {
        // the record count is only known after the kernel: fetch it, then copy exactly that many
        CU(cudaMemcpyAsync(ctx->h_counters, ctx->d_counters, 8, cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
        const uint64_t used = ctx->h_counters[0];
        if (out->hdr_used) *out->hdr_used = used;
        const uint64_t copy = std::min<uint64_t>(std::min<uint64_t>(used, d.hdr_cap), out->hdr_cap);
        if (out->hdr && copy) CU(cudaMemcpyAsync(out->hdr, d.hdr, copy * sizeof(shl_hdr_span), cudaMemcpyDeviceToHost, ctx->stream));
        if (used > d.hdr_cap || (out->hdr && used > out->hdr_cap))
            return fail(ctx, SHL_ERR_CAPACITY, "header record arena overflowed: raise max_records / hdr_cap");
    }
    return SHL_OK;
}
static int place_out_d2h(shl_ctx* ctx, const shl_place_out* out, uint64_t n) {
    const shl_place_out& d = ctx->dq;
    D2H(out->md5, d.md5, n * 16); D2H(out->hash, d.hash, n * 4); D2H(out->node, d.node, n * 4);
    if (out->hist) {
        // `hist` is ADDED to: bring the device bins back through the pinned mirror
        std::vector<uint64_t> tmp(ctx->n_hosts);
        CU(cudaMemcpyAsync(tmp.data(), ctx->d_hist, (size_t)ctx->n_hosts * 8, cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
        for (uint32_t i = 0; i < ctx->n_hosts; i++) out->hist[i] += tmp[i];
    }
    return SHL_OK;
}

// ---- small batches: one block out, one copy back ---------------------------------------------------
// A reactor tick hands over a few hundred messages.  Bringing their results back field by field
// costs some twenty cudaMemcpyAsync calls and three stream syncs (~115 us a call whatever the
// size); here the kernel writes every field into one device block laid out for this n, one D2H
// brings the block to a pinned mirror, and the fields are copied out to the caller's arrays on
// the host.  Returns 1 when the call does not qualify (the general path then runs).
struct SmallLayout {
    uint64_t content_len, md5, u32[10], hash, node, used, hist, rec, rec_cap, total;
};
static SmallLayout small_layout(uint64_t n, uint32_t n_hosts, uint64_t arena_bytes) {
    SmallLayout l{}; uint64_t o = 0;
    l.md5 = o; o += 16 * n;                                  // 16-byte stores: keep it first (the block is 256-byte aligned)
    l.content_len = o; o += 8 * n;
    for (int k = 0; k < 10; k++) { l.u32[k] = o; o += 4 * n; }
    l.hash = o; o += 4 * n;
    l.node = o; o += 4 * n;
    o = (o + 7) & ~7ull;
    l.used = o; o += 8;
    l.hist = o; o += 8ull * n_hosts;
    l.rec = o; l.rec_cap = arena_bytes / 5 + 64; o += 8 * l.rec_cap;   // a header line is at least "k: " CRLF: the records always fit
    l.total = o;
    return l;
}
static int route_small(shl_ctx* ctx, const uint8_t* arena, uint64_t arena_bytes, const uint64_t* off, const uint32_t* len,
                       uint64_t n, const shl_parse_out* pout, const shl_place_out* qout) {
    if (!ctx->small_path || !arena || n == 0 || n > shl_ctx::kSmallMsgs) return 1;
    const uint32_t nh = qout ? ctx->n_hosts : 0u;
    const SmallLayout l = small_layout(n, nh, arena_bytes);
    if (!ctx->d_small) {
        const SmallLayout big = small_layout(std::min<uint64_t>(shl_ctx::kSmallMsgs, ctx->max_msgs), 4096,
                                             std::min<uint64_t>(ctx->max_bytes, 4ull << 20));   // bigger ticks and rings beyond 4 096 hosts take the general path
        CU(cudaMalloc(reinterpret_cast<void**>(&ctx->d_small), big.total));
        CU(cudaHostAlloc(reinterpret_cast<void**>(&ctx->h_small), big.total, cudaHostAllocDefault));
        ctx->small_bytes = big.total;
    }
    if (l.total > ctx->small_bytes) return 1;
    if (int rc = stage_in(ctx, arena, arena_bytes, off, len, n, ctx->d_off, ctx->d_len)) return rc;
    uint8_t* d = ctx->d_small;
    CU(cudaMemsetAsync(d + l.used, 0, 8 + 8ull * nh, ctx->stream));
    shl_parse_out o{};
    o.content_len = reinterpret_cast<int64_t*>(d + l.content_len);
    o.status = reinterpret_cast<uint32_t*>(d + l.u32[0]); o.first_line_len = reinterpret_cast<uint32_t*>(d + l.u32[1]);
    o.headers_len = reinterpret_cast<uint32_t*>(d + l.u32[2]); o.sp1 = reinterpret_cast<uint32_t*>(d + l.u32[3]);
    o.sp2 = reinterpret_cast<uint32_t*>(d + l.u32[4]); o.fl_end = reinterpret_cast<uint32_t*>(d + l.u32[5]);
    o.slash = reinterpret_cast<uint32_t*>(d + l.u32[6]); o.status_code = reinterpret_cast<int32_t*>(d + l.u32[7]);
    o.n_headers = reinterpret_cast<uint32_t*>(d + l.u32[8]); o.hdr_base = reinterpret_cast<uint32_t*>(d + l.u32[9]);
    o.hdr = reinterpret_cast<shl_hdr_span*>(d + l.rec); o.hdr_cap = l.rec_cap; o.hdr_used = reinterpret_cast<uint64_t*>(d + l.used);
    shl_place_out q{};
    if (qout) {
        q.md5 = d + l.md5; q.hash = reinterpret_cast<uint32_t*>(d + l.hash); q.node = reinterpret_cast<uint32_t*>(d + l.node);
        q.hist = qout->hist ? reinterpret_cast<uint64_t*>(d + l.hist) : nullptr;
    }
    ctx->cur_dp = o; ctx->cur_n = n;
    if (qout) { ctx->cur_node = q.node; ctx->cur_hist = q.hist; }
    if (int rc = launch_route(ctx, ctx->d_arena, ctx->d_off, ctx->d_len, n, &o, qout ? &q : nullptr)) return rc;
    // one copy brings the fields and as many records as a batch normally has (24 per message);
    // the rare batch with more fetches the rest with a second copy
    const uint64_t first_recs = std::min<uint64_t>(l.rec_cap, 24 * n + 64);
    CU(cudaMemcpyAsync(ctx->h_small, d, l.rec + 8 * first_recs, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    const uint8_t* h = ctx->h_small;
    const uint64_t used = *reinterpret_cast<const uint64_t*>(h + l.used);
    if (used > l.rec_cap) return fail(ctx, SHL_ERR_CAPACITY, "small-batch record block overflowed (cannot happen: a header line is >= 5 bytes)");
    if (used > first_recs) {
        CU(cudaMemcpyAsync(ctx->h_small + l.rec + 8 * first_recs, d + l.rec + 8 * first_recs, 8 * (used - first_recs), cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
    }
    if (pout->hdr && used > pout->hdr_cap)
        return fail(ctx, SHL_ERR_CAPACITY, "header record arena overflowed: raise max_records / hdr_cap");
#define OUT(dst, src_off, bytes) do { if (dst) std::memcpy((dst), h + (src_off), (bytes)); } while (0)
    OUT(pout->content_len, l.content_len, 8 * n);
    OUT(pout->status, l.u32[0], 4 * n); OUT(pout->first_line_len, l.u32[1], 4 * n); OUT(pout->headers_len, l.u32[2], 4 * n);
    OUT(pout->sp1, l.u32[3], 4 * n); OUT(pout->sp2, l.u32[4], 4 * n); OUT(pout->fl_end, l.u32[5], 4 * n);
    OUT(pout->slash, l.u32[6], 4 * n); OUT(pout->status_code, l.u32[7], 4 * n); OUT(pout->n_headers, l.u32[8], 4 * n);
    OUT(pout->hdr_base, l.u32[9], 4 * n);
    OUT(pout->hdr, l.rec, 8 * used);
    if (pout->hdr_used) *pout->hdr_used = used;
    if (qout) {
        OUT(qout->md5, l.md5, 16 * n); OUT(qout->hash, l.hash, 4 * n); OUT(qout->node, l.node, 4 * n);
        if (qout->hist) {                                   // `hist` is ADDED to
            const uint64_t* bins = reinterpret_cast<const uint64_t*>(h + l.hist);
            for (uint32_t i = 0; i < nh; i++) qout->hist[i] += bins[i];
        }
    }
#undef OUT
    return SHL_OK;
}

// ---- chunked, multi-stream version of parse/route for large batches -----------------------------
// The batch is cut into message ranges; chunk c runs H2D(arena slice, index slice) -> kernel ->
// D2H(result slices) on stream c % kPipe, so the copy engines of both directions and the SMs are
// busy at the same time (PCIe is full duplex; one stream would serialise the three stages).
// Each chunk owns a slice of the record arena and its own allocation counter; hdr_base values are
// biased so they index the caller's arena directly.  Returns 1 when the batch is not chunkable
// (offsets not locally dense) so the caller falls back to the single-stream path.
static int route_pipelined(shl_ctx* ctx, const uint8_t* arena, uint64_t arena_bytes, const uint64_t* off, const uint32_t* len,
                           uint64_t n, const shl_parse_out* pout, const shl_place_out* qout, shl_msg_rec* crec = nullptr) {
    constexpr int kPipe = shl_ctx::kPipe, kMaxChunks = shl_ctx::kMaxChunks;
    if (!ctx->pipeline || !arena || n < 65536) return 1;
    ctx->cur_dp = ctx->dp; ctx->cur_n = n;
    if (qout) { ctx->cur_node = ctx->dq.node; ctx->cur_hist = qout->hist ? ctx->d_hist : nullptr; }
    const int nch = (int)std::min<uint64_t>((uint64_t)std::min(kMaxChunks, ctx->max_chunks), std::max<uint64_t>(2, n / (uint64_t)ctx->chunk_msgs));
    const uint64_t per = ((n + nch - 1) / nch + 31) & ~31ull;
    // cheap density probe (the exact byte range of a chunk is computed right before it is queued,
    // so the host scan of the index overlaps the copies of the previous chunks): the arena is
    // chunkable when sampled offsets are non-decreasing
    {
        const uint64_t stride = std::max<uint64_t>(1, n / 2048);
        uint64_t prev = 0;
        for (uint64_t i = 0; i < n; i += stride) { if (off[i] < prev) return 1; prev = off[i]; }
    }
    const uint64_t host_cap = pout->hdr ? pout->hdr_cap : ctx->dp.hdr_cap;
    const uint64_t cap_c = std::min<uint64_t>(ctx->dp.hdr_cap, host_cap) / nch;
    if (cap_c == 0 || cap_c * nch > 0xffffffffull) return 1;
    CU(cudaMemsetAsync(ctx->d_chunk_used, 0, kMaxChunks * 8, ctx->stream));
    shl_place_out dq = ctx->dq;
    if (crec) {               // compact output: the record carries the node; digests only if asked for
        dq.hash = nullptr; dq.node = nullptr; dq.md5 = qout->md5 ? ctx->dq.md5 : nullptr;
        ctx->cur_node = nullptr;
    }
    if (qout) {
        dq.hist = qout->hist ? ctx->d_hist : nullptr;
        if (dq.hist) CU(cudaMemsetAsync(ctx->d_hist, 0, (size_t)ctx->n_hosts * 8, ctx->stream));
    }
    CU(cudaEventRecord(ctx->ev_fork, ctx->stream));
    for (int s = 0; s < kPipe; s++) CU(cudaStreamWaitEvent(ctx->pipe[s], ctx->ev_fork, 0));
    const shl_parse_out& d = ctx->dp;
#define D2S(dst, src, cnt, esz) do { if (dst) CU(cudaMemcpyAsync((uint8_t*)(dst) + i0 * (esz), (const uint8_t*)(src) + i0 * (esz), (cnt) * (esz), cudaMemcpyDeviceToHost, st)); } while (0)
    for (int c = 0; c < nch; c++) {
        const uint64_t i0 = std::min<uint64_t>(n, c * per), i1 = std::min<uint64_t>(n, i0 + per), m = i1 - i0;
        if (!m) continue;
        cudaStream_t st = ctx->pipe[c % kPipe];
        uint64_t lo = ~0ull, hi = 0;
        for (uint64_t i = i0; i < i1; i++) { lo = std::min(lo, off[i]); hi = std::max(hi, off[i] + len[i]); }
        if (hi > arena_bytes) return fail(ctx, SHL_ERR_ARG, "a message lies outside the arena");
        lo &= ~31ull;
        CU(cudaMemcpyAsync(ctx->d_arena + lo, arena + lo, hi - lo, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(ctx->d_off + i0, off + i0, m * 8, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(ctx->d_len + i0, len + i0, m * 4, cudaMemcpyHostToDevice, st));
        shl_parse_out o = d;
        o.status += i0; o.first_line_len += i0; o.headers_len += i0; o.sp1 += i0; o.sp2 += i0; o.fl_end += i0; o.slash += i0;
        o.status_code += i0; o.n_headers += i0; o.hdr_base += i0; o.content_len += i0;
        o.hdr = d.hdr + (uint64_t)c * cap_c; o.hdr_cap = cap_c; o.hdr_used = ctx->d_chunk_used + c;
        shl_place_out q = dq;
        if (qout) { if (q.md5) q.md5 += 16 * i0; if (q.hash) q.hash += i0; if (q.node) q.node += i0; }
        if (int rc = launch_route(ctx, ctx->d_arena, ctx->d_off + i0, ctx->d_len + i0, m, &o, qout ? &q : nullptr, st,
                                  (uint32_t)((uint64_t)c * cap_c), crec ? ctx->d_crec + i0 : nullptr)) return rc;
        if (crec) CU(cudaMemcpyAsync(crec + i0, ctx->d_crec + i0, m * sizeof(shl_msg_rec), cudaMemcpyDeviceToHost, st));
        D2S(pout->status, d.status, m, 4); D2S(pout->first_line_len, d.first_line_len, m, 4); D2S(pout->headers_len, d.headers_len, m, 4);
        D2S(pout->sp1, d.sp1, m, 4); D2S(pout->sp2, d.sp2, m, 4); D2S(pout->fl_end, d.fl_end, m, 4); D2S(pout->slash, d.slash, m, 4);
        D2S(pout->status_code, d.status_code, m, 4); D2S(pout->n_headers, d.n_headers, m, 4); D2S(pout->hdr_base, d.hdr_base, m, 4);
        D2S(pout->content_len, d.content_len, m, 8);
        if (qout) { D2S(qout->md5, dq.md5, m, 16); D2S(qout->hash, dq.hash, m, 4); D2S(qout->node, dq.node, m, 4); }
        CU(cudaMemcpyAsync(ctx->h_chunk_used + c, ctx->d_chunk_used + c, 8, cudaMemcpyDeviceToHost, st));
        CU(cudaEventRecord(ctx->ev_chunk[c], st));
    }
#undef D2S
    // the record count of a chunk is only known once its kernel ran: fetch exactly that many
    uint64_t used_total = 0; bool overflow = false;
    for (int c = 0; c < nch; c++) {
        const uint64_t i0 = std::min<uint64_t>(n, c * per), i1 = std::min<uint64_t>(n, i0 + per);
        if (i0 == i1) continue;
        CU(cudaEventSynchronize(ctx->ev_chunk[c]));
        const uint64_t used = ctx->h_chunk_used[c];
        used_total += std::min(used, cap_c);
        if (used > cap_c) overflow = true;
        const uint64_t copy = std::min(used, cap_c);
        if (pout->hdr && copy)
            CU(cudaMemcpyAsync(pout->hdr + (uint64_t)c * cap_c, d.hdr + (uint64_t)c * cap_c, copy * sizeof(shl_hdr_span),
                               cudaMemcpyDeviceToHost, ctx->pipe[c % kPipe]));
    }
    if (pout->hdr_used) *pout->hdr_used = used_total;
    for (int s = 0; s < kPipe; s++) {
        CU(cudaEventRecord(ctx->ev_join[s], ctx->pipe[s]));
        CU(cudaStreamWaitEvent(ctx->stream, ctx->ev_join[s], 0));
    }
    if (qout) { shl_place_out only_hist{}; only_hist.hist = qout->hist; if (int rc = place_out_d2h(ctx, &only_hist, 0)) return rc; }
    if (overflow) return fail(ctx, SHL_ERR_CAPACITY, "header record arena overflowed: raise max_records / hdr_cap");
    return SHL_OK;
}

// The caller asked for the lower-cased header names: give the device parse block the arena for
// them (allocated on first use: header names cannot outgrow the message bytes).
static int with_lkeys(shl_ctx* ctx, const shl_parse_out* host, shl_parse_out* dev) {
    *dev = ctx->dp;
    if (!host->lkeys) return SHL_OK;
    if (!host->lkeys_used || !host->lkey_base) return fail(ctx, SHL_ERR_ARG, "lkeys needs lkeys_used and lkey_base too");
    if (!ctx->d_lkeys) { CU(dalloc(&ctx->d_lkeys, ctx->max_bytes + 64)); CU(dalloc(&ctx->d_lkey_base, ctx->max_msgs)); }
    dev->lkeys = ctx->d_lkeys; dev->lkeys_cap = std::min<uint64_t>(ctx->max_bytes, host->lkeys_cap);
    dev->lkeys_used = ctx->d_counters + 2; dev->lkey_base = ctx->d_lkey_base;
    CU(cudaMemsetAsync(ctx->d_counters + 2, 0, 8, ctx->stream));
    return SHL_OK;
}
static int lkeys_d2h(shl_ctx* ctx, const shl_parse_out* host, uint64_t n) {
    if (!host->lkeys) return SHL_OK;
    CU(cudaMemcpyAsync(host->lkey_base, ctx->d_lkey_base, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(ctx->h_counters + 2, ctx->d_counters + 2, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    const uint64_t used = ctx->h_counters[2];
    *host->lkeys_used = used;
    if (used > std::min<uint64_t>(ctx->max_bytes, host->lkeys_cap)) return fail(ctx, SHL_ERR_CAPACITY, "lower-cased key arena overflowed: raise lkeys_cap");
    if (used) CU(cudaMemcpyAsync(host->lkeys, ctx->d_lkeys, used, cudaMemcpyDeviceToHost, ctx->stream));
    return SHL_OK;
}

int shl_parse_batch(shl_ctx* ctx, const uint8_t* arena, uint64_t arena_bytes, const uint64_t* off, const uint32_t* len,
                    uint64_t n, const shl_parse_out* out) {
    if (!out) return fail(ctx, SHL_ERR_ARG, "out is null");
    if (ctx && off && len && n <= ctx->max_msgs && arena_bytes <= ctx->max_bytes && !out->lkeys) {
        CU(cudaSetDevice(ctx->device));
        int rc = route_small(ctx, arena, arena_bytes, off, len, n, out, nullptr);
        if (rc <= 0) return rc;
        rc = route_pipelined(ctx, arena, arena_bytes, off, len, n, out, nullptr);
        if (rc <= 0) return rc;
    }
    if (int rc = stage_in(ctx, arena, arena_bytes, off, len, n, ctx->d_off, ctx->d_len)) return rc;
    CU(cudaMemsetAsync(ctx->d_counters, 0, 8, ctx->stream));
    ctx->cur_dp = ctx->dp; ctx->cur_n = n;
    shl_parse_out dp;
    if (int rc = with_lkeys(ctx, out, &dp)) return rc;
    if (int rc = shl_parse_dev(ctx, ctx->d_arena, ctx->d_off, ctx->d_len, n, &dp)) return rc;
    if (int rc = parse_out_d2h(ctx, out, n)) return rc;
    return lkeys_d2h(ctx, out, n);
}

int shl_route_batch(shl_ctx* ctx, const uint8_t* arena, uint64_t arena_bytes, const uint64_t* off, const uint32_t* len,
                    uint64_t n, const shl_parse_out* pout, const shl_place_out* qout) {
    if (!pout || !qout) return fail(ctx, SHL_ERR_ARG, "out is null");
    if (ctx && !ctx->d_ring_values) return fail(ctx, SHL_ERR_NO_RING, "shl_ring_build() has not been called");
    if (ctx && off && len && n <= ctx->max_msgs && arena_bytes <= ctx->max_bytes && !pout->lkeys) {
        CU(cudaSetDevice(ctx->device));
        int rc = route_small(ctx, arena, arena_bytes, off, len, n, pout, qout);
        if (rc <= 0) return rc;
        rc = route_pipelined(ctx, arena, arena_bytes, off, len, n, pout, qout);
        if (rc <= 0) return rc;
    }
    if (int rc = stage_in(ctx, arena, arena_bytes, off, len, n, ctx->d_off, ctx->d_len)) return rc;
    CU(cudaMemsetAsync(ctx->d_counters, 0, 8, ctx->stream));
    shl_place_out dq = ctx->dq;
    dq.hist = qout->hist ? ctx->d_hist : nullptr;
    if (dq.hist) CU(cudaMemsetAsync(ctx->d_hist, 0, (size_t)ctx->n_hosts * 8, ctx->stream));
    ctx->cur_dp = ctx->dp; ctx->cur_n = n;
    ctx->cur_node = ctx->dq.node; ctx->cur_hist = dq.hist;
    shl_parse_out dp;
    if (int rc = with_lkeys(ctx, pout, &dp)) return rc;
    if (int rc = shl_route_dev(ctx, ctx->d_arena, ctx->d_off, ctx->d_len, n, &dp, &dq)) return rc;
    if (int rc = parse_out_d2h(ctx, pout, n)) return rc;
    if (int rc = lkeys_d2h(ctx, pout, n)) return rc;
    return place_out_d2h(ctx, qout, n);
}

// ---- compact output: one 32-byte record per message
int shl_route_compact_dev(shl_ctx* ctx, const uint8_t* d_arena, const uint64_t* d_off, const uint32_t* d_len, uint64_t n,
                          shl_msg_rec* d_rec, shl_hdr_span* d_hdr, uint64_t hdr_cap, uint64_t* d_hdr_used, uint8_t* d_md5, uint64_t* d_hist) {
    if (!ctx || !d_rec || !d_hdr || !d_hdr_used) return fail(ctx, SHL_ERR_ARG, "shl_route_compact_dev: null pointer");
    if (ctx->n_hosts > 65534u) return fail(ctx, SHL_ERR_ARG, "compact records hold the node in 16 bits: rings of at most 65 534 nodes");
    shl_parse_out o{};
    o.hdr = d_hdr; o.hdr_cap = hdr_cap; o.hdr_used = d_hdr_used;
    shl_place_out q{};
    q.md5 = d_md5; q.hist = d_hist;
    return launch_route(ctx, d_arena, d_off, d_len, n, &o, &q, nullptr, 0, d_rec);
}

int shl_route_compact_batch(shl_ctx* ctx, const uint8_t* arena, uint64_t arena_bytes, const uint64_t* off, const uint32_t* len,
                            uint64_t n, shl_msg_rec* rec, shl_hdr_span* hdr, uint64_t hdr_cap, uint64_t* hdr_used,
                            uint8_t* md5, uint64_t* hist) {
    if (!ctx || !rec || !off || !len) return fail(ctx, SHL_ERR_ARG, "shl_route_compact_batch: null pointer");
    if (!ctx->d_ring_values) return fail(ctx, SHL_ERR_NO_RING, "shl_ring_build() has not been called");
    if (ctx->n_hosts > 65534u) return fail(ctx, SHL_ERR_ARG, "compact records hold the node in 16 bits: rings of at most 65 534 nodes");
    if (n > ctx->max_msgs || arena_bytes > ctx->max_bytes) return fail(ctx, SHL_ERR_ARG, "batch larger than the capacity given to shl_ctx_create");
    CU(cudaSetDevice(ctx->device));
    if (!ctx->d_crec) CU(dalloc(&ctx->d_crec, ctx->max_msgs));
    shl_parse_out pout{};
    pout.hdr = hdr; pout.hdr_cap = hdr_cap; pout.hdr_used = hdr_used;
    shl_place_out qout{};
    qout.md5 = md5; qout.hist = hist;
    int rc = route_pipelined(ctx, arena, arena_bytes, off, len, n, &pout, &qout, rec);
    if (rc <= 0) return rc;
    if (int rc2 = stage_in(ctx, arena, arena_bytes, off, len, n, ctx->d_off, ctx->d_len)) return rc2;
    CU(cudaMemsetAsync(ctx->d_counters, 0, 8, ctx->stream));
    shl_place_out dq = ctx->dq;
    dq.hist = hist ? ctx->d_hist : nullptr;
    dq.md5 = md5 ? ctx->dq.md5 : nullptr; dq.hash = nullptr; dq.node = nullptr;
    if (dq.hist) CU(cudaMemsetAsync(ctx->d_hist, 0, (size_t)ctx->n_hosts * 8, ctx->stream));
    ctx->cur_node = nullptr; ctx->cur_hist = dq.hist;
    if (int rc2 = launch_route(ctx, ctx->d_arena, ctx->d_off, ctx->d_len, n, &ctx->dp, &dq, nullptr, 0, ctx->d_crec)) return rc2;
    CU(cudaMemcpyAsync(rec, ctx->d_crec, n * sizeof(shl_msg_rec), cudaMemcpyDeviceToHost, ctx->stream));
    if (int rc2 = parse_out_d2h(ctx, &pout, n)) return rc2;
    return place_out_d2h(ctx, &qout, n);
}

int shl_place_batch(shl_ctx* ctx, const uint8_t* arena, uint64_t arena_bytes, const uint64_t* key_off, const uint32_t* key_len,
                    uint64_t n, const shl_place_out* out) {
    if (!out) return fail(ctx, SHL_ERR_ARG, "out is null");
    if (ctx && !ctx->d_ring_values) return fail(ctx, SHL_ERR_NO_RING, "shl_ring_build() has not been called");
    if (int rc = stage_in(ctx, arena, arena_bytes, key_off, key_len, n, ctx->d_koff, ctx->d_klen)) return rc;
    shl_place_out dq = ctx->dq;
    dq.hist = out->hist ? ctx->d_hist : nullptr;
    if (dq.hist) CU(cudaMemsetAsync(ctx->d_hist, 0, (size_t)ctx->n_hosts * 8, ctx->stream));
    ctx->cur_node = ctx->dq.node; ctx->cur_hist = dq.hist;
    if (int rc = shl_place_dev(ctx, ctx->d_arena, ctx->d_koff, ctx->d_klen, n, &dq)) return rc;
    return place_out_d2h(ctx, out, n);
}

int shl_frame_batch(shl_ctx* ctx, const uint8_t* arena, uint64_t arena_bytes, const uint64_t* off, const uint32_t* len,
                    uint64_t n, const uint32_t* status, const uint32_t* first_line_len, const uint32_t* headers_len,
                    const shl_frame_out* out) {
    if (!out || !status || !first_line_len || !headers_len) return fail(ctx, SHL_ERR_ARG, "null pointer");
    if (int rc = stage_in(ctx, arena, arena_bytes, off, len, n, ctx->d_off, ctx->d_len)) return rc;
    // the parse fields travel with the call so it can be used on its own (they are tiny next to the arena)
    if (n != ctx->cur_n) { ctx->cur_dp = ctx->dp; ctx->cur_n = n; }   // not the continuation of the last parse: the small-batch block is laid out for ITS n
    const shl_parse_out& cur = ctx->cur_dp;
    CU(cudaMemcpyAsync(cur.status, status, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(cur.first_line_len, first_line_len, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(cur.headers_len, headers_len, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemsetAsync(ctx->d_counters + 1, 0, 8, ctx->stream));
    if (int rc = shl_frame_dev(ctx, ctx->d_arena, ctx->d_off, ctx->d_len, n, cur.status, cur.first_line_len,
                               cur.headers_len, &ctx->df)) return rc;
    const shl_frame_out& d = ctx->df;
    D2H(out->fstatus, d.fstatus, n * 4); D2H(out->n_chunks, d.n_chunks, n * 4); D2H(out->chunk_base, d.chunk_base, n * 4);
    D2H(out->body_end, d.body_end, n * 8); D2H(out->body_bytes, d.body_bytes, n * 8); D2H(out->last_call, d.last_call, n * 4);
    if (out->chunk || out->chunk_used) {
        CU(cudaMemcpyAsync(ctx->h_counters + 1, ctx->d_counters + 1, 8, cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
        const uint64_t used = ctx->h_counters[1];
        if (out->chunk_used) *out->chunk_used = used;
        const uint64_t copy = std::min<uint64_t>(std::min<uint64_t>(used, d.chunk_cap), out->chunk_cap);
        if (out->chunk && copy) CU(cudaMemcpyAsync(out->chunk, d.chunk, copy * sizeof(shl_chunk_span), cudaMemcpyDeviceToHost, ctx->stream));
        if (used > d.chunk_cap || (out->chunk && used > out->chunk_cap))
            return fail(ctx, SHL_ERR_CAPACITY, "chunk record arena overflowed: raise max_records / chunk_cap");
    }
    return SHL_OK;
}

int shl_body_dev(shl_ctx* ctx, const uint8_t* d_arena, const uint64_t* d_off, const uint32_t* d_len, uint64_t n,
                 const shl_parse_out* p, const shl_frame_out* f, const shl_body_out* d_out) {
    if (!ctx || !d_arena || !d_off || !d_len || !p || !f || !d_out) return fail(ctx, SHL_ERR_ARG, "null pointer");
    if (n == 0) return SHL_OK;
    CU(cudaSetDevice(ctx->device));
    shl::BodyArgs a{};
    a.arena = d_arena; a.off = d_off; a.len = d_len; a.n = n;
    a.status = p->status; a.first_line_len = p->first_line_len; a.headers_len = p->headers_len; a.content_len = p->content_len;
    a.sp1 = p->sp1;
    a.fstatus = f->fstatus; a.n_chunks = f->n_chunks; a.chunk_base = f->chunk_base; a.chunk = f->chunk; a.body_bytes = f->body_bytes;
    a.out = *d_out;
    shl::k_body_plan<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(a);
    shl::k_body_copy<<<dim3((unsigned)n, shl::kBodySlices), 256, 0, ctx->stream>>>(a);
    ctx->launches += 2;
    CU(cudaGetLastError());
    return SHL_OK;
}

int shl_body_batch(shl_ctx* ctx, uint64_t n, const shl_body_out* out) {
    if (!ctx || !out) return fail(ctx, SHL_ERR_ARG, "null pointer");
    if (n > ctx->max_msgs) return fail(ctx, SHL_ERR_ARG, "batch larger than the capacity given to shl_ctx_create");
    CU(cudaSetDevice(ctx->device));
    if (!ctx->d_body) {
        CU(cudaMalloc(reinterpret_cast<void**>(&ctx->d_body), ctx->max_bytes + 16 * ctx->max_msgs + 64));
        CU(dalloc(&ctx->d_body_off, ctx->max_msgs)); CU(dalloc(&ctx->d_body_len, ctx->max_msgs));
    }
    shl_body_out d{};
    d.body_off = ctx->d_body_off; d.body_len = ctx->d_body_len; d.body = ctx->d_body;
    d.body_cap = std::min<uint64_t>(ctx->max_bytes + 16 * ctx->max_msgs, out->body ? out->body_cap : ~0ull);
    d.body_used = ctx->d_counters + 1;                  // the frame pass is done with its counter
    CU(cudaMemsetAsync(ctx->d_counters + 1, 0, 8, ctx->stream));
    if (int rc = shl_body_dev(ctx, ctx->d_arena, ctx->d_off, ctx->d_len, n, &ctx->cur_dp, &ctx->df, &d)) return rc;
    D2H(out->body_off, d.body_off, n * 8); D2H(out->body_len, d.body_len, n * 8);
    CU(cudaMemcpyAsync(ctx->h_counters + 1, ctx->d_counters + 1, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    const uint64_t used = ctx->h_counters[1];
    if (out->body_used) *out->body_used = used;
    if (used > d.body_cap) return fail(ctx, SHL_ERR_CAPACITY, "body arena overflowed: raise body_cap");
    if (out->body && used) CU(cudaMemcpyAsync(out->body, d.body, used, cudaMemcpyDeviceToHost, ctx->stream));
    return SHL_OK;
}

int shl_group_dev(shl_ctx* ctx, const uint32_t* d_node, uint64_t n, uint32_t n_hosts, const uint64_t* d_hist,
                  uint32_t* d_start, uint32_t* d_cursor, uint32_t* d_order) {
    if (!ctx || !d_node || !d_hist || !d_start || !d_cursor || !d_order || n_hosts == 0) return fail(ctx, SHL_ERR_ARG, "null pointer");
    CU(cudaSetDevice(ctx->device));
    shl::k_group_scan<<<1, 1024, 0, ctx->stream>>>(d_hist, n_hosts, d_start, d_cursor);
    if (n) shl::k_group_scatter<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(d_node, n, n_hosts, d_cursor, d_order);
    ctx->launches += 2;
    CU(cudaGetLastError());
    return SHL_OK;
}

int shl_group_batch(shl_ctx* ctx, uint64_t n, uint32_t* start, uint32_t* order) {
    if (!ctx || !start || !order) return fail(ctx, SHL_ERR_ARG, "null pointer");
    if (!ctx->d_ring_values) return fail(ctx, SHL_ERR_NO_RING, "shl_ring_build() has not been called");
    if (n > ctx->max_msgs) return fail(ctx, SHL_ERR_ARG, "batch larger than the capacity given to shl_ctx_create");
    CU(cudaSetDevice(ctx->device));
    if (!ctx->d_gorder) CU(dalloc(&ctx->d_gorder, ctx->max_msgs));
    if (ctx->gcap < ctx->n_hosts + 1) {
        cudaFree(ctx->d_gstart); cudaFree(ctx->d_gcursor); ctx->d_gstart = ctx->d_gcursor = nullptr;
        CU(dalloc(&ctx->d_gstart, ctx->n_hosts + 1)); CU(dalloc(&ctx->d_gcursor, ctx->n_hosts + 1));
        ctx->gcap = ctx->n_hosts + 1;
    }
    if (!ctx->cur_node || !ctx->cur_hist) return fail(ctx, SHL_ERR_ARG, "shl_group_batch needs a preceding route / place call with a histogram");
    if (int rc = shl_group_dev(ctx, ctx->cur_node, n, ctx->n_hosts, ctx->cur_hist, ctx->d_gstart, ctx->d_gcursor, ctx->d_gorder)) return rc;
    CU(cudaMemcpyAsync(start, ctx->d_gstart, (size_t)(ctx->n_hosts + 1) * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(order, ctx->d_gorder, n * 4, cudaMemcpyDeviceToHost, ctx->stream));   // only start[n_hosts] entries are meaningful
    return SHL_OK;
}

// ---- memcached binary-protocol frames per node (what pylibmc sends behind Server.py:335 / :432)
int shl_mc_frames_dev(shl_ctx* ctx, const uint8_t* d_arena, const uint64_t* d_off, const uint32_t* d_sp1, const uint32_t* d_sp2,
                      const uint32_t* d_order, uint64_t n_placed, int op, uint32_t ttl, const uint32_t* d_value_len,
                      const shl_mc_out* d_out) {
    if (!ctx || !d_arena || !d_off || !d_sp1 || !d_sp2 || !d_order || !d_out || !d_out->frame_off || !d_out->frames || !d_out->frames_used)
        return fail(ctx, SHL_ERR_ARG, "shl_mc_frames_dev: null pointer");
    if (op != SHL_MC_GET && op != SHL_MC_SET) return fail(ctx, SHL_ERR_ARG, "shl_mc_frames_dev: op is SHL_MC_GET or SHL_MC_SET");
    CU(cudaSetDevice(ctx->device));
    if (n_placed > ctx->max_msgs) return fail(ctx, SHL_ERR_ARG, "shl_mc_frames_dev: more requests than the ctx was created for");
    if (!ctx->d_mc_sums) CU(dalloc(&ctx->d_mc_sums, ctx->max_msgs / shl::kMcScanItems + 2));
    shl::McArgs a{};
    a.arena = d_arena; a.off = d_off; a.sp1 = d_sp1; a.sp2 = d_sp2; a.order = d_order; a.n = n_placed;
    a.value_len = d_value_len; a.op = (uint32_t)op; a.ttl = ttl;
    a.frame_off = d_out->frame_off; a.block_sums = ctx->d_mc_sums;
    a.frames = d_out->frames; a.frames_cap = d_out->frames_cap; a.frames_used = d_out->frames_used;
    const unsigned nb = (unsigned)((n_placed + shl::kMcScanItems - 1) / shl::kMcScanItems);
    if (nb) shl::k_mc_sizes<<<nb, 256, 0, ctx->stream>>>(a);
    shl::k_mc_block_scan<<<1, 1024, 0, ctx->stream>>>(ctx->d_mc_sums, nb, a.frames_used, a.frame_off, n_placed);
    if (n_placed) shl::k_mc_emit<<<(unsigned)((n_placed * 32 + 255) / 256), 256, 0, ctx->stream>>>(a);
    ctx->launches += nb ? 3 : 1;
    CU(cudaGetLastError());
    return SHL_OK;
}

int shl_mc_frames_batch(shl_ctx* ctx, uint64_t n, uint64_t n_placed, int op, uint32_t ttl, const uint32_t* value_len,
                        const shl_mc_out* out) {
    if (!ctx || !out || !out->frame_off) return fail(ctx, SHL_ERR_ARG, "shl_mc_frames_batch: null pointer");
    if (!ctx->d_gorder || !ctx->cur_node) return fail(ctx, SHL_ERR_ARG, "shl_mc_frames_batch needs a preceding shl_route_batch + shl_group_batch");
    if (n > ctx->max_msgs || n_placed > n) return fail(ctx, SHL_ERR_ARG, "shl_mc_frames_batch: bad counts");
    CU(cudaSetDevice(ctx->device));
    const uint64_t cap = ctx->max_bytes + 32 * ctx->max_msgs;
    if (!ctx->d_mc_frames) {
        CU(dalloc(&ctx->d_mc_off, ctx->max_msgs + 1)); CU(dalloc(&ctx->d_mc_frames, cap)); CU(dalloc(&ctx->d_mc_vlen, ctx->max_msgs));
        ctx->mc_frames_cap = cap;
    }
    if (value_len) CU(cudaMemcpyAsync(ctx->d_mc_vlen, value_len, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    shl_mc_out d{};
    d.frame_off = ctx->d_mc_off; d.frames = ctx->d_mc_frames; d.frames_cap = ctx->mc_frames_cap; d.frames_used = ctx->d_counters + 3;
    if (int rc = shl_mc_frames_dev(ctx, ctx->d_arena, ctx->d_off, ctx->cur_dp.sp1, ctx->cur_dp.sp2, ctx->d_gorder, n_placed, op, ttl,
                                   value_len ? ctx->d_mc_vlen : nullptr, &d)) return rc;
    CU(cudaMemcpyAsync(out->frame_off, ctx->d_mc_off, (n_placed + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaMemcpyAsync(ctx->h_counters + 3, ctx->d_counters + 3, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    const uint64_t used = ctx->h_counters[3];
    if (out->frames_used) *out->frames_used = used;
    if (used > ctx->mc_frames_cap || (out->frames && used > out->frames_cap))
        return fail(ctx, SHL_ERR_CAPACITY, "memcached frame arena overflowed: raise frames_cap");
    if (out->frames && used) CU(cudaMemcpyAsync(out->frames, ctx->d_mc_frames, used, cudaMemcpyDeviceToHost, ctx->stream));
    return SHL_OK;
}

int shl_chunk_header_batch(shl_ctx* ctx, const uint8_t* arena, uint64_t arena_bytes, const uint64_t* off, const uint32_t* len,
                           uint64_t n, int32_t* hdr_len, uint64_t* size, uint8_t* ok) {
    if (!hdr_len || !size || !ok) return fail(ctx, SHL_ERR_ARG, "null pointer");
    if (int rc = stage_in(ctx, arena, arena_bytes, off, len, n, ctx->d_off, ctx->d_len)) return rc;
    if (n) {
        shl::k_chunk_header<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(ctx->d_arena, ctx->d_off, ctx->d_len, n,
                                                                                 ctx->d_ch_len, ctx->d_ch_size, ctx->d_ch_ok);
        ctx->launches++;
        CU(cudaGetLastError());
    }
    D2H(hdr_len, ctx->d_ch_len, n * 4); D2H(size, ctx->d_ch_size, n * 8); D2H(ok, ctx->d_ch_ok, n);
    return SHL_OK;
}

int shl_code_hist(shl_ctx* ctx, uint64_t* out64, int reset) {
    if (!ctx || !out64) return SHL_ERR_ARG;
    CU(cudaSetDevice(ctx->device));
    CU(cudaMemcpyAsync(ctx->h_counters + 2, cur_codes(ctx), 64 * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (reset) CU(cudaMemsetAsync(cur_codes(ctx), 0, 64 * 8, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    memcpy(out64, ctx->h_counters + 2, 64 * 8);
    return SHL_OK;
}
void* shl_code_hist_dev(shl_ctx* ctx) { return ctx ? (void*)cur_codes(ctx) : nullptr; }
void* shl_node_hist_dev(shl_ctx* ctx) { return ctx ? (void*)ctx->d_hist : nullptr; }
// Test hook of the -DSHL_BOUNDS_CHECK build: out-of-bounds indices counted by the kernels so far
// (-1 in a build without the checks).
int shl_debug_oob_count(void) {
#ifdef SHL_BOUNDS_CHECK
    unsigned int v = 0;
    if (cudaMemcpyFromSymbol(&v, shl::g_shl_oob, sizeof v) != cudaSuccess) return -2;
    return (int)v;
#else
    return -1;
#endif
}
int shl_set_force_scalar(shl_ctx* ctx, int on) { if (!ctx) return SHL_ERR_ARG; ctx->force_scalar = on ? 1 : 0; return SHL_OK; }

// ------------------------------------------------------------------------------------ counters across GPUs
// The path has no exchange step: requests are independent and every GPU holds the whole ring.
// What the ranks of a job share are COUNTERS, `uint64[n_hosts + SHL_N_CODES]` (per-node placement
// histogram, per-outcome counts).  The library owns them (two device buffers), the NCCL
// communicator (one rank per process, one process per GPU) and a side stream: the all-reduce of
// batch k is ordered against the ctx stream by events only, so the kernel of batch k + 1 runs
// while it is in flight.  No PyTorch anywhere.
void* shl_stats_dev(shl_ctx* ctx) { return ctx ? (void*)ctx->d_stats[ctx->stat_k] : nullptr; }
uint32_t shl_stats_len(shl_ctx* ctx) { return ctx ? ctx->stats_len : 0u; }

int shl_comm_unique_id(uint8_t* id, uint32_t cap) {
    shl_ctx* ctx = nullptr;
    if (!id || cap < sizeof(ncclUniqueId)) return fail(nullptr, SHL_ERR_ARG, "shl_comm_unique_id: buffer smaller than SHL_COMM_ID_BYTES");
    if (!g_nccl.load()) return fail(ctx, SHL_ERR_CUDA, g_nccl.why);
    ncclUniqueId u;
    const ncclResult_t r = g_nccl.GetUniqueId(&u);
    if (r != ncclSuccess) return fail(ctx, SHL_ERR_CUDA, std::string("ncclGetUniqueId: ") + g_nccl.GetErrorString(r));
    memcpy(id, &u, sizeof u);
    return SHL_OK;
}
static_assert(sizeof(ncclUniqueId) == SHL_COMM_ID_BYTES, "SHL_COMM_ID_BYTES follows ncclUniqueId");

int shl_comm_init(shl_ctx* ctx, const uint8_t* id, int rank, int world) {
    if (!ctx || !id || world < 1 || rank < 0 || rank >= world) return fail(ctx, SHL_ERR_ARG, "shl_comm_init: bad argument");
    if (ctx->comm) return fail(ctx, SHL_ERR_ARG, "shl_comm_init: this ctx already has a communicator");
    if (!g_nccl.load()) return fail(ctx, SHL_ERR_CUDA, g_nccl.why);
    CU(cudaSetDevice(ctx->device));
    ncclUniqueId u;
    memcpy(&u, id, sizeof u);
    const ncclResult_t r = g_nccl.CommInitRank(&ctx->comm, world, u, rank);
    if (r != ncclSuccess) { ctx->comm = nullptr; return fail(ctx, SHL_ERR_CUDA, std::string("ncclCommInitRank: ") + g_nccl.GetErrorString(r)); }
    ctx->comm_rank = rank; ctx->comm_world = world;
    return SHL_OK;
}
int shl_comm_size(shl_ctx* ctx) { return ctx && ctx->comm ? ctx->comm_world : 1; }

int shl_stats_allreduce(shl_ctx* ctx) {
    if (!ctx) return SHL_ERR_ARG;
    CU(cudaSetDevice(ctx->device));
    const int k = ctx->stat_k, k2 = (k + 1) % shl_ctx::kStatBufs;
    const size_t bytes = (size_t)ctx->stats_len * 8;
    CU(cudaEventRecord(ctx->ev_stats_ready[k], ctx->stream));
    CU(cudaStreamWaitEvent(ctx->comm_stream, ctx->ev_stats_ready[k], 0));
    if (ctx->comm) {
        const ncclResult_t r = g_nccl.AllReduce(ctx->d_stats[k], ctx->d_stats[k], ctx->stats_len, ncclUint64, ncclSum, ctx->comm, ctx->comm_stream);
        if (r != ncclSuccess) return fail(ctx, SHL_ERR_CUDA, std::string("ncclAllReduce: ") + g_nccl.GetErrorString(r));
    }
    CU(cudaMemcpyAsync(ctx->h_stats + (size_t)k * ctx->stats_len, ctx->d_stats[k], bytes, cudaMemcpyDeviceToHost, ctx->comm_stream));
    CU(cudaEventRecord(ctx->ev_stats_done[k], ctx->comm_stream));
    ctx->stat_last = k;
    // the next buffer takes the next batches; its previous all-reduce (kStatBufs calls ago) must be over
    CU(cudaStreamWaitEvent(ctx->stream, ctx->ev_stats_done[k2], 0));
    CU(cudaMemsetAsync(ctx->d_stats[k2], 0, bytes, ctx->stream));
    ctx->stat_k = k2;
    return SHL_OK;
}
int shl_stats_join(shl_ctx* ctx) {
    if (!ctx) return SHL_ERR_ARG;
    for (int k = 0; k < shl_ctx::kStatBufs; k++) CU(cudaStreamWaitEvent(ctx->stream, ctx->ev_stats_done[k], 0));
    return SHL_OK;
}
int shl_stats_fetch(shl_ctx* ctx, uint64_t* out, uint32_t n) {
    if (!ctx || !out || n > ctx->stats_len) return fail(ctx, SHL_ERR_ARG, "shl_stats_fetch: bad argument");
    if (ctx->stat_last < 0) return fail(ctx, SHL_ERR_ARG, "shl_stats_fetch: no shl_stats_allreduce yet");
    CU(cudaEventSynchronize(ctx->ev_stats_done[ctx->stat_last]));
    memcpy(out, ctx->h_stats + (size_t)ctx->stat_last * ctx->stats_len, (size_t)n * 8);
    return SHL_OK;
}
int shl_comm_allreduce_u64(shl_ctx* ctx, uint64_t* vals, uint32_t n, int op) {
    if (!ctx || !vals || n == 0 || n > 64 || op < 0 || op > 1) return fail(ctx, SHL_ERR_ARG, "shl_comm_allreduce_u64: bad argument");
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));                // doubles as the barrier's "my work is done"
    CU(cudaStreamSynchronize(ctx->comm_stream));
    if (!ctx->comm) return SHL_OK;
    memcpy(ctx->h_scratch, vals, (size_t)n * 8);
    CU(cudaMemcpyAsync(ctx->d_scratch, ctx->h_scratch, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->comm_stream));
    const ncclResult_t r = g_nccl.AllReduce(ctx->d_scratch, ctx->d_scratch, n, ncclUint64, op ? ncclMax : ncclSum, ctx->comm, ctx->comm_stream);
    if (r != ncclSuccess) return fail(ctx, SHL_ERR_CUDA, std::string("ncclAllReduce: ") + g_nccl.GetErrorString(r));
    CU(cudaMemcpyAsync(ctx->h_scratch, ctx->d_scratch, (size_t)n * 8, cudaMemcpyDeviceToHost, ctx->comm_stream));
    CU(cudaStreamSynchronize(ctx->comm_stream));
    memcpy(vals, ctx->h_scratch, (size_t)n * 8);
    return SHL_OK;
}

// ------------------------------------------------------------------------------------ corpora
int shl_corpus_lengths_host(uint64_t seed, uint64_t first, uint64_t n, int kind, uint32_t* len) {
    if (!len) return SHL_ERR_ARG;
    for (uint64_t i = 0; i < n; i++) len[i] = shl::gen_message(seed, first + i, kind, nullptr);
    return SHL_OK;
}
int shl_corpus_fill_host(uint64_t seed, uint64_t first, uint64_t n, int kind, const uint64_t* off, uint8_t* arena) {
    if (!off || !arena) return SHL_ERR_ARG;
    for (uint64_t i = 0; i < n; i++) shl::gen_message(seed, first + i, kind, arena + off[i]);
    return SHL_OK;
}
}  // extern "C"

namespace shl {
__global__ void k_corpus_lengths(uint64_t seed, uint64_t first, uint64_t n, int kind, uint32_t* len) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) len[i] = gen_message(seed, first + i, kind, nullptr);
}
__global__ void k_corpus_fill(uint64_t seed, uint64_t first, uint64_t n, int kind, const uint64_t* off, uint8_t* arena) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) gen_message(seed, first + i, kind, arena + off[i]);
}
}  // namespace shl

extern "C" {
int shl_corpus_lengths_dev(shl_ctx* ctx, uint64_t seed, uint64_t first, uint64_t n, int kind, uint32_t* d_len) {
    if (!ctx || !d_len) return SHL_ERR_ARG;
    CU(cudaSetDevice(ctx->device));
    if (n) { shl::k_corpus_lengths<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(seed, first, n, kind, d_len); ctx->launches++; }
    CU(cudaGetLastError());
    return SHL_OK;
}
int shl_corpus_fill_dev(shl_ctx* ctx, uint64_t seed, uint64_t first, uint64_t n, int kind, const uint64_t* d_off, uint8_t* d_arena) {
    if (!ctx || !d_off || !d_arena) return SHL_ERR_ARG;
    CU(cudaSetDevice(ctx->device));
    if (n) { shl::k_corpus_fill<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(seed, first, n, kind, d_off, d_arena); ctx->launches++; }
    CU(cudaGetLastError());
    return SHL_OK;
}
}  // extern "C"

// =================================================================================================
// shl_place_static — placement for callers that must not allocate.
//
// Every other entry point works out of buffers that shl_ctx_create / shl_ring_build obtained with
// cudaMalloc.  Some callers cannot afford that: the proxy's single reactor thread (an allocation
// takes the driver's device-wide lock and can stall every batch in flight on that GPU), a small
// "which node owns this key?" probe in a process that never built a ctx, or a deployment where the
// HBM is already fully handed out to other tenants and an allocation may simply fail at run time.
// For them the kernel's working memory (staged keys, ring, results, histogram bins) lives in ONE
// file-scope device array of fixed size that exists from the moment the library is loaded on a
// GPU until the process exits.  Nothing here calls cudaMalloc, cudaFree or any allocating routine.
//
// The array is shared by every call and keeps whatever the previous call left in it; the bins that
// must start at zero are cleared on every call.  Two calls whose GPU work overlaps on one device
// would trample each other's keys and results: serialise calls per device.
// =================================================================================================
namespace {

constexpr uint64_t kStArena = 0;                                                 // SHL_STATIC_MAX_KEY_BYTES + pad
constexpr uint64_t kStOff = kStArena + SHL_STATIC_MAX_KEY_BYTES + 256;
constexpr uint64_t kStLen = kStOff + (uint64_t)SHL_STATIC_MAX_KEYS * 8;
constexpr uint64_t kStVals = kStLen + (uint64_t)SHL_STATIC_MAX_KEYS * 4;
constexpr uint64_t kStNodes = kStVals + (uint64_t)SHL_STATIC_MAX_POINTS * 4;
constexpr uint64_t kStMd5 = kStNodes + (uint64_t)SHL_STATIC_MAX_POINTS * 4;
constexpr uint64_t kStHash = kStMd5 + (uint64_t)SHL_STATIC_MAX_KEYS * 16;
constexpr uint64_t kStNode = kStHash + (uint64_t)SHL_STATIC_MAX_KEYS * 4;
constexpr uint64_t kStHist = kStNode + (uint64_t)SHL_STATIC_MAX_KEYS * 4;
constexpr uint64_t kStTotal = kStHist + (uint64_t)SHL_STATIC_MAX_HOSTS * 8;
static_assert(kStOff % 256 == 0 && kStLen % 256 == 0 && kStVals % 256 == 0 && kStNodes % 256 == 0 &&
              kStMd5 % 256 == 0 && kStHash % 256 == 0 && kStNode % 256 == 0 && kStHist % 256 == 0, "carve-outs stay aligned");

// ~536 KiB per GPU that loads the library.  In HOST code this name is only a stand-in for the
// symbol, not a device address: it must never be handed to a kernel or a memcpy directly.
static __device__ __align__(256) uint8_t g_place_scratch[kStTotal];

constexpr int kMaxDevices = 64;
uint8_t* g_place_scratch_dev[kMaxDevices] = {};      // device address of the array, per device, once resolved

}  // namespace

extern "C" int shl_place_static(int device, const uint8_t* arena, uint64_t arena_bytes, const uint64_t* key_off,
                                const uint32_t* key_len, uint64_t n, const uint32_t* ring_values,
                                const uint32_t* ring_nodes, uint32_t n_points, uint32_t n_hosts,
                                uint8_t* md5, uint32_t* hash, uint32_t* node, uint64_t* hist) {
    shl_ctx* ctx = nullptr;      // for the CU() macro: errors land in the process-wide last-error text
    if (!arena || !key_off || !key_len || !ring_values || !ring_nodes)
        return fail(nullptr, SHL_ERR_ARG, "shl_place_static: null input pointer");
    char why[200];
    if (n > SHL_STATIC_MAX_KEYS) {
        snprintf(why, sizeof why, "shl_place_static: %llu keys exceed SHL_STATIC_MAX_KEYS = %d", (unsigned long long)n, SHL_STATIC_MAX_KEYS);
        return fail(nullptr, SHL_ERR_ARG, why);
    }
    if (arena_bytes > SHL_STATIC_MAX_KEY_BYTES) {
        snprintf(why, sizeof why, "shl_place_static: %llu key bytes exceed SHL_STATIC_MAX_KEY_BYTES = %d", (unsigned long long)arena_bytes, SHL_STATIC_MAX_KEY_BYTES);
        return fail(nullptr, SHL_ERR_ARG, why);
    }
    if (n_points == 0 || n_points > SHL_STATIC_MAX_POINTS || n_hosts == 0 || n_hosts > SHL_STATIC_MAX_HOSTS) {
        snprintf(why, sizeof why, "shl_place_static: ring of %u points / %u hosts outside (0, %d] points / (0, %d] hosts",
                 n_points, n_hosts, SHL_STATIC_MAX_POINTS, SHL_STATIC_MAX_HOSTS);
        return fail(nullptr, SHL_ERR_ARG, why);
    }
    for (uint64_t i = 0; i < n; i++)
        if (key_off[i] > arena_bytes || key_len[i] > arena_bytes - key_off[i])
            return fail(nullptr, SHL_ERR_ARG, "shl_place_static: a key lies outside the arena");
    for (uint32_t i = 0; i < n_points; i++)
        if (ring_nodes[i] >= n_hosts) return fail(nullptr, SHL_ERR_ARG, "shl_place_static: ring node index >= n_hosts");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(nullptr, SHL_ERR_NO_DEVICE, "no CUDA device: libshellac_b200 has no CPU fallback");
    if (device < 0 || device >= ndev || device >= kMaxDevices) return fail(nullptr, SHL_ERR_ARG, "shl_place_static: bad device index");
    DeviceGuard guard;                                   // the caller's current device is put back on every way out
    CU(guard.enter(device));
    // Set-up, once per device and with that device current: ask the runtime where THIS device's
    // copy of the array lives.  The first call on a device does it, so make that first call
    // before any stream capture begins.
    uint8_t* base = g_place_scratch_dev[device];
    if (!base) {
        CU(cudaGetSymbolAddress(reinterpret_cast<void**>(&base), g_place_scratch));
        g_place_scratch_dev[device] = base;
    }
    if (n == 0) return SHL_OK;
    const cudaStream_t st = 0;                           // the device's default stream
    CU(cudaMemcpyAsync(base + kStArena, arena, arena_bytes, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(base + kStOff, key_off, n * 8, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(base + kStLen, key_len, n * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(base + kStVals, ring_values, (size_t)n_points * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(base + kStNodes, ring_nodes, (size_t)n_points * 4, cudaMemcpyHostToDevice, st));
    CU(cudaMemsetAsync(base + kStHist, 0, (size_t)n_hosts * 8, st));     // the bins are the one thing that must start at zero
    shl::PlaceArgs a{};
    a.arena = base + kStArena;
    a.key_off = reinterpret_cast<const uint64_t*>(base + kStOff);
    a.key_len = reinterpret_cast<const uint32_t*>(base + kStLen);
    a.n = n;
    a.ring_values = reinterpret_cast<const uint32_t*>(base + kStVals);
    a.ring_nodes = reinterpret_cast<const uint32_t*>(base + kStNodes);
    a.n_points = n_points; a.n_hosts = n_hosts;
    a.out.md5 = base + kStMd5;
    a.out.hash = reinterpret_cast<uint32_t*>(base + kStHash);
    a.out.node = reinterpret_cast<uint32_t*>(base + kStNode);
    a.out.hist = reinterpret_cast<uint64_t*>(base + kStHist);
    shl::k_place<<<(unsigned)((n + 255) / 256), 256, (size_t)n_hosts * 4, st>>>(a);
    CU(cudaGetLastError());
    uint64_t bins[SHL_STATIC_MAX_HOSTS];                 // on the stack (the success path allocates no host memory either;
                                                         //  the error paths build a std::string for shl_last_error)
    if (md5) CU(cudaMemcpyAsync(md5, base + kStMd5, n * 16, cudaMemcpyDeviceToHost, st));
    if (hash) CU(cudaMemcpyAsync(hash, base + kStHash, n * 4, cudaMemcpyDeviceToHost, st));
    if (node) CU(cudaMemcpyAsync(node, base + kStNode, n * 4, cudaMemcpyDeviceToHost, st));
    if (hist) CU(cudaMemcpyAsync(bins, base + kStHist, (size_t)n_hosts * 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (hist) for (uint32_t i = 0; i < n_hosts; i++) hist[i] += bins[i];
    return SHL_OK;
}
