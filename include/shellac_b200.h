/* shellac_b200.h — C ABI of libshellac_b200.so
 *
 * The B200-native replacement for the one data-parallel hot path of kmacrow/Shellac:
 * batched HTTP/1.1 message tokenising + body framing (reference: HttpParser.py:140-337),
 * cache key = request URL verbatim (reference: Server.py:327) and ketama/MD5 placement of each
 * key on a memcached continuum (reference: pylibmc.Client(..., behaviors={'ketama': True}),
 * Server.py:81-83, called at Server.py:335 and :432; arithmetic = libmemcached 0.44, not vendored).
 *
 * Plain C: pointers and sizes only.  Bind it with ctypes the way the
 * reference binds libmemcached through pylibmc.  Every entry point returns 0 on success and a
 * negative SHL_ERR_* on a library/CUDA failure (text via shl_last_error).  Per-message parse
 * outcomes are DATA (the `status` word), not errors: the Python facade raises the matching
 * exception type at the call at which the reference would have raised.
 *
 * Threading: one ctx per GPU; a ctx is not thread-safe (same contract as HttpParser.py:21).
 * All work is queued on the ctx's stream and is asynchronous until shl_sync().
 */
#ifndef SHELLAC_B200_H
#define SHELLAC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------
 * Library return codes
 * ---------------------------------------------------------------------------------------- */
#define SHL_OK              0
#define SHL_ERR_CUDA       -1   /* a CUDA runtime call failed; see shl_last_error()            */
#define SHL_ERR_ARG        -2   /* bad argument (null pointer, size beyond the ctx capacity)   */
#define SHL_ERR_NO_RING    -3   /* placement requested before shl_ring_build()                 */
#define SHL_ERR_NO_DEVICE  -4   /* no CUDA device: there is no CPU fallback, by design         */
#define SHL_ERR_CAPACITY   -5   /* header/chunk record arena overflowed (raise max_records)    */

/* ------------------------------------------------------------------------------------------
 * Per-message status word  (u32):  bits 0..7 = outcome code, bits 8.. = flags
 *
 * The outcome codes follow the exception taxonomy of the reference parser
 * (HttpParser.py:304-337, 194-200, 234-241); "INCOMPLETE_*" are the states in which the
 * reference's parse() returns `length` and waits for more bytes (HttpParser.py:156-157,
 * 177-178, 208-211, 255-256, 286-287, 291-294).
 * ---------------------------------------------------------------------------------------- */
enum {
    SHL_S_OK                 = 0,  /* message_complete() is True                                   */
    SHL_S_EMPTY              = 1,  /* length == 0: parse() returns None       (HttpParser.py:143)  */
    SHL_S_INCOMPLETE_LINE    = 2,  /* no CRLF yet                             (HttpParser.py:155)  */
    SHL_S_INCOMPLETE_HEADERS = 3,  /* no CRLFCRLF after the first line        (HttpParser.py:176)  */
    SHL_S_INCOMPLETE_BODY    = 4,  /* headers done, body/chunks not all here  (HttpParser.py:208,291) */
    SHL_E_FIRSTLINE_PARTS    = 16, /* ValueError: rstrip().split(' ',2) != 3 parts (HttpParser.py:307) */
    SHL_E_VERSION_SPLIT      = 17, /* ValueError: split('/') != 2 parts        (HttpParser.py:311,318) */
    SHL_E_VERSION_FLOAT      = 18, /* ValueError: float(version)               (HttpParser.py:312,319) */
    SHL_E_STATUS_INT         = 19, /* ValueError: int(status)                  (HttpParser.py:313) */
    SHL_E_HDR_NO_SEP         = 20, /* ValueError: header line without ": "     (HttpParser.py:328) */
    SHL_E_HDR_MULTI_SEP      = 21, /* ValueError: header line with >= 2 ": "   (HttpParser.py:328) */
    SHL_E_CLEN_INT           = 22, /* ValueError: int(content-length)          (HttpParser.py:197) */
    SHL_E_CLEN_LIST          = 23, /* TypeError : duplicate content-length -> int(list) (HttpParser.py:197) */
    SHL_E_CHUNK_HEX          = 24, /* ValueError: int(chunk-size, 16)          (HttpParser.py:239,241) */
    SHL_E_LIB_HDR_TOO_LARGE  = 32, /* library limit: header block ends beyond byte 65535 of the message */
    SHL_E_LIB_RECORDS_FULL   = 33  /* library limit: the header/chunk record arena is full        */
};
#define SHL_CODE(st)             ((st) & 0xffu)
#define SHL_F_IS_REQUEST         (1u << 8)   /* HttpParser.is_request()                            */
#define SHL_F_HEADERS_COMPLETE   (1u << 9)   /* HttpParser.headers_complete()                      */
#define SHL_F_MESSAGE_COMPLETE   (1u << 10)  /* HttpParser.message_complete()                      */
#define SHL_F_ON_BODY            (1u << 11)  /* HttpParser._on_body                                */
#define SHL_F_CHUNKED            (1u << 12)  /* HttpParser._chunked                                */
#define SHL_F_VERSION_SHIFT      13          /* 2 bits: 0 = "1.0", 1 = "1.1", 2 = other valid float text, 3 = not parsed */
#define SHL_F_VERSION_MASK       (3u << SHL_F_VERSION_SHIFT)
#define SHL_F_STATUS_WIDE        (1u << 15)  /* int(status) valid but does not fit status_code: host converts the span */
#define SHL_F_CLEN_WIDE          (1u << 16)  /* int(content-length) valid but beyond int64: host converts the span     */
#define SHL_F_FIRST_LINE_DONE    (1u << 17)  /* _parse_first_line returned (fields below are valid)                   */
#define SHL_F_KEY_TOO_LONG       (1u << 18)  /* request whose cache key (the URL, Server.py:327) is longer than
                                              * SHL_MAX_KEY_LEN: pylibmc raises ValueError before hashing, at
                                              * Server.py:335; such a request is parsed but NOT placed
                                              * (node = 0xffffffff, digest zero, not counted in the histogram)        */
#define SHL_MAX_KEY_LEN          250         /* MEMCACHED_MAX_KEY - 1 (libmemcached constants.h; restated)            */

/* One header line, offsets relative to the first byte of the message.
 * key   = line[:sep]            (NOT stripped; the reference lower()s it, HttpParser.py:330)
 * value = line[sep+2:].strip()  (py2 whitespace set {9..13, 32},     HttpParser.py:329)        */
typedef struct shl_hdr_span {
    uint16_t key_off, key_len, val_off, val_len;
} shl_hdr_span;

/* One chunk of a chunked body, offsets relative to the first byte of the message.
 * [hdr_off, hdr_off+hdr_len) is what CHUNK_HEADER_RX matched (HttpParser.py:39-40,253-260),
 * i.e. it includes the previous chunk's trailing CRLF; data follows immediately.
 * One reference parse() call consumes hdr_len + data_len bytes (HttpParser.py:263,299).         */
typedef struct shl_chunk_span {
    uint32_t hdr_off, hdr_len, data_off, data_len;
} shl_chunk_span;

/* Struct-of-arrays parse result.  Every pointer addresses `n` elements (host memory for the
 * *_batch entry points, device memory for the *_dev ones).  Offsets are relative to the message.
 *   method = [0, sp1)   url = [sp1+1, sp2)   c = [sp2+1, fl_end)      (HttpParser.py:307)
 *   request : version text = [slash+1, fl_end)                           (HttpParser.py:318-319)
 *   response: version text = [slash+1, sp1), reason = c, status text = [sp1+1, sp2)  (:311-314)  */
typedef struct shl_parse_out {
    uint32_t* status;          /* code | flags                                                     */
    uint32_t* first_line_len;  /* nb_parsed of the first parse() call = idx+2 (HttpParser.py:163)   */
    uint32_t* headers_len;     /* nb_parsed of the second parse() call    (HttpParser.py:184)       */
    uint32_t* sp1;
    uint32_t* sp2;
    uint32_t* fl_end;          /* end of the rstrip()ped first line                                */
    uint32_t* slash;           /* position of the '/' that split() used                            */
    int32_t*  status_code;     /* int(b) for responses, 0 for requests                             */
    uint32_t* n_headers;       /* header lines tokenised (lines before the offending one on error) */
    uint32_t* hdr_base;        /* index of this message's first record in hdr[]                    */
    int64_t*  content_len;     /* HttpParser._content_len after the header state (0 if chunked)    */
    shl_hdr_span* hdr;         /* record arena, `hdr_cap` entries                                  */
    uint64_t  hdr_cap;
    uint64_t* hdr_used;        /* one element: records written.  The records of a message are contiguous
                                * from hdr_base[i]; between messages the arena may have GAPS (the host-buffer
                                * calls cut batches of >= 65 536 messages into chunks that each own a slice
                                * of it): records in use end at max(hdr_base[i] + n_headers[i]), not at hdr_used */
    /* Optional (all four NULL / 0 = not wanted): the header NAMES lower-cased as the reference stores
     * them (key.lower(), HttpParser.py:330; ASCII only, like Python 2's str.lower()).  Message i's
     * names lie back to back, in record order, from lkeys[lkey_base[i]]: name k starts
     * sum(key_len[0..k-1]) bytes in and is key_len[k] long.                                        */
    uint8_t*  lkeys;           /* byte arena, `lkeys_cap` bytes                                    */
    uint64_t  lkeys_cap;
    uint64_t* lkeys_used;      /* one element: bytes handed out                                    */
    uint32_t* lkey_base;       /* [n]                                                              */
} shl_parse_out;

/* Chunked-body framing result (struct-of-arrays, n elements each). */
typedef struct shl_frame_out {
    uint32_t* fstatus;         /* SHL_S_OK / SHL_S_INCOMPLETE_BODY / SHL_E_CHUNK_HEX / 0xff = not chunked */
    uint32_t* n_chunks;        /* data-carrying chunks (the terminating 0-chunk is not counted)     */
    uint32_t* chunk_base;      /* index of the first record in chunk[]                              */
    uint64_t* body_end;        /* message-relative offset one past the last byte parse() consumed   */
    uint64_t* body_bytes;      /* sum of data_len = length of the de-chunked body                   */
    uint32_t* last_call;       /* nb_parsed of the final parse() call (HttpParser.py:270,276,280)   */
    shl_chunk_span* chunk;     /* record arena, `chunk_cap` entries                                 */
    uint64_t  chunk_cap;
    uint64_t* chunk_used;
} shl_frame_out;

/* Body gather result: message i's body is body[body_off[i] : body_off[i] + body_len[i]]
 * (each body starts 16-byte aligned; body_used[0] = bytes of the arena handed out).            */
typedef struct shl_body_out {
    uint64_t* body_off;        /* [n]                                                              */
    uint64_t* body_len;        /* [n] bytes _parse_body has written for this message               */
    uint8_t*  body;            /* byte arena, `body_cap` bytes                                     */
    uint64_t  body_cap;
    uint64_t* body_used;       /* one element                                                      */
} shl_body_out;

/* Compact per-message record (32 bytes, one memory sector): everything shl_parse_out + the node of
 * shl_place_out say about a message, for callers that bring results back over PCIe (72 bytes per
 * request become 32; the MD5 is optional and the hash, its first four bytes, is dropped).
 * Offsets are relative to the message and fit 16 bits by the library's 64 KiB header limit; a first
 * line that does not (no CRLF within 64 KiB) is SHL_E_LIB_HDR_TOO_LARGE here.  `tail`: requests
 * carry their node (0xffff = not placed; rings of at most 65 534 nodes), responses their status
 * code (SHL_F_STATUS_WIDE when it does not fit 16 bits: the host converts the span).            */
typedef struct shl_msg_rec {
    uint32_t status;
    uint32_t hdr_base;
    int64_t  content_len;
    uint16_t first_line_len, headers_len, sp1, sp2, fl_end, slash;
    uint16_t n_headers;
    uint16_t tail;
} shl_msg_rec;

/* Placement result (struct-of-arrays, n elements each). */
typedef struct shl_place_out {
    uint8_t*  md5;             /* 16 bytes per key: RFC 1321 digest (libhashkit md5.c)              */
    uint32_t* hash;            /* LE32(md5[0:4])            (libhashkit hashkit_md5)               */
    uint32_t* node;            /* server index in ring order (libmemcached hash.c dispatch_host)   */
    uint64_t* hist;            /* n_nodes counters, ADDED to (caller zeroes); may be NULL          */
} shl_place_out;

/* Ring modes (libmemcached hosts.c update_continuum).                                           */
#define SHL_RING_KETAMA           0  /* 'ketama': True -> 100 points/server, 1 MD5 per point: the reference's setting (Server.py:83) */
#define SHL_RING_KETAMA_WEIGHTED  1  /* ketama_weighted -> 160 points/server, 4 points per MD5 (BASELINE config 4)                  */

typedef struct shl_ctx shl_ctx;

/* ------------------------------------------------------------------------------------------
 * Context / memory.  Mirrors pylibmc.Client construction (Server.py:81-83): one object owns the
 * device buffers, the stream and the ring.
 * ---------------------------------------------------------------------------------------- */
int  shl_device_count(void);
/* "0000:1b:00.0"-style PCI address of a device (len >= 13), so that the host side can place its
 * pinned arenas on the GPU's NUMA node (shellac_b200/hostmem.py).  No reference counterpart. */
int  shl_device_pci_bus_id(int device, char* buf, int len);
/* Free / total device memory (cudaMemGetInfo), e.g. to size a ctx for the 180 GB of a B200 or to
 * check that a call allocated nothing.  No reference counterpart. */
int  shl_device_mem_info(int device, uint64_t* free_bytes, uint64_t* total_bytes);
int  shl_ctx_create(int device, uint64_t max_msgs, uint64_t max_bytes, uint64_t max_records, shl_ctx** out);
void shl_ctx_destroy(shl_ctx* ctx);
const char* shl_last_error(shl_ctx* ctx);           /* ctx may be NULL: last error of ctx creation */
int  shl_sync(shl_ctx* ctx);
int  shl_pinned_alloc(void** ptr, uint64_t bytes);  /* page-locked host memory for arenas/outputs  */
int  shl_pinned_free(void* ptr);
int  shl_device_alloc(shl_ctx* ctx, void** dptr, uint64_t bytes);
int  shl_device_free(shl_ctx* ctx, void* dptr);
int  shl_memcpy_h2d(shl_ctx* ctx, void* dst, const void* src, uint64_t bytes);   /* async on the ctx stream */
int  shl_memcpy_d2h(shl_ctx* ctx, void* dst, const void* src, uint64_t bytes);
int  shl_memcpy_d2d(shl_ctx* ctx, void* dst, const void* src, uint64_t bytes);
int  shl_memset_dev(shl_ctx* ctx, void* dst, int value, uint64_t bytes);
/* Device-side timing on the ctx stream (CUDA events). slot in [0,16). */
int  shl_event_record(shl_ctx* ctx, int slot);
int  shl_event_elapsed_ms(shl_ctx* ctx, int slot_start, int slot_stop, float* ms);
/* Stream handle (cudaStream_t) so a caller can order other work (e.g. an NCCL all-reduce) after ours. */
void* shl_stream(shl_ctx* ctx);

/* ------------------------------------------------------------------------------------------
 * Ring.  Replaces libmemcached's update_continuum (hosts.c) behind pylibmc.Client(servers, ...)
 * (Server.py:81-83).  hosts[i]/ports[i] are the '%s:%d' pieces in CLI order (Server.py:81,493-505);
 * weights may be NULL (all 1).  Host index = list position.
 * ---------------------------------------------------------------------------------------- */
int  shl_ring_build(shl_ctx* ctx, const char* const* hosts, const uint32_t* ports,
                    const uint32_t* weights, uint32_t n_hosts, int mode);
int  shl_ring_size(shl_ctx* ctx, uint32_t* n_points, uint32_t* n_hosts);
int  shl_ring_export(shl_ctx* ctx, uint32_t* values, uint32_t* nodes, uint32_t cap);

/* ------------------------------------------------------------------------------------------
 * Host-buffer entry points (the drop-in boundary).  Inputs and outputs are HOST pointers
 * (pinned for full speed); each call copies H2D, runs the kernels, copies D2H, all on the ctx
 * stream; results are valid after shl_sync().
 *
 * shl_parse_batch  replaces  HttpParser.parse() driven to completion over each message
 *                  (HttpParser.py:140-202, 304-337; caller loop Server.py:320-325).
 * shl_frame_batch  replaces  HttpParser._parse_chunked / Content-Length framing
 *                  (HttpParser.py:204-301) for the messages shl_parse_batch flagged.
 * shl_place_batch  replaces  the placement inside pylibmc get/set (Server.py:335,432).
 * shl_route_batch  = parse + (key = url(), Server.py:327) + place, fused on the device: the
 *                  request path of Server._read_requests (Server.py:320-335).
 * ---------------------------------------------------------------------------------------- */
int  shl_parse_batch(shl_ctx* ctx, const uint8_t* arena, uint64_t arena_bytes,
                     const uint64_t* off, const uint32_t* len, uint64_t n, const shl_parse_out* out);
int  shl_frame_batch(shl_ctx* ctx, const uint8_t* arena, uint64_t arena_bytes,
                     const uint64_t* off, const uint32_t* len, uint64_t n,
                     const uint32_t* status, const uint32_t* first_line_len, const uint32_t* headers_len,
                     const shl_frame_out* out);
int  shl_place_batch(shl_ctx* ctx, const uint8_t* arena, uint64_t arena_bytes,
                     const uint64_t* key_off, const uint32_t* key_len, uint64_t n, const shl_place_out* out);
int  shl_route_batch(shl_ctx* ctx, const uint8_t* arena, uint64_t arena_bytes,
                     const uint64_t* off, const uint32_t* len, uint64_t n,
                     const shl_parse_out* pout, const shl_place_out* qout);
/* shl_route_batch with the compact record as its output: rec[n], the header records, and optionally
 * the digests (md5 may be NULL) and the per-node histogram (hist may be NULL; ADDED to).  Same kernel,
 * same H2D / kernel / D2H pipelining; 32 + 8 x n_headers bytes come back per request.             */
int  shl_route_compact_batch(shl_ctx* ctx, const uint8_t* arena, uint64_t arena_bytes,
                             const uint64_t* off, const uint32_t* len, uint64_t n,
                             shl_msg_rec* rec, shl_hdr_span* hdr, uint64_t hdr_cap, uint64_t* hdr_used,
                             uint8_t* md5, uint64_t* hist);
int  shl_route_compact_dev(shl_ctx* ctx, const uint8_t* d_arena, const uint64_t* d_off, const uint32_t* d_len,
                           uint64_t n, shl_msg_rec* d_rec, shl_hdr_span* d_hdr, uint64_t hdr_cap,
                           uint64_t* d_hdr_used, uint8_t* d_md5, uint64_t* d_hist);
/* One CHUNK_HEADER_RX.match + _parse_chunk_size at the START of each window (HttpParser.py:253-261):
 * hdr_len[i] = len(match.group()) or -1 when the regex does not match; size[i] = int(size, 16);
 * ok[i] = 0 when that int() raises ValueError.  The primitive an incremental caller needs.       */
int  shl_chunk_header_batch(shl_ctx* ctx, const uint8_t* arena, uint64_t arena_bytes,
                            const uint64_t* off, const uint32_t* len, uint64_t n,
                            int32_t* hdr_len, uint64_t* size, uint8_t* ok);
/* shl_body_batch / shl_body_dev replace the body accumulation of HttpParser._parse_body
 * (HttpParser.py:340-347; callers :213-214 for Content-Length bodies, :296-297 per chunk): the
 * bytes the framing spans select are gathered into one contiguous body per message (identity
 * content-coding; gzip inflate stays with the host's zlib as in the reference, :343-344).
 * shl_body_batch works on the batch whose shl_parse_batch + shl_frame_batch results are still
 * resident on this ctx (call it right after them, same n).                                       */
int  shl_body_batch(shl_ctx* ctx, uint64_t n, const shl_body_out* out);
/* Group the requests of a placed batch by node: what pylibmc does behind get/set before it
 * talks to each memcached node (Server.py:335, 432).  start[k] .. start[k+1] index the requests
 * of node k in order[] (start has n_hosts + 1 entries; requests without a node are left out; the
 * order inside a node is arbitrary).  shl_group_batch works on the batch whose shl_route_batch /
 * shl_place_batch results (with a histogram) are still resident on this ctx.                     */
int  shl_group_batch(shl_ctx* ctx, uint64_t n, uint32_t* start, uint32_t* order);
int  shl_group_dev(shl_ctx* ctx, const uint32_t* d_node, uint64_t n, uint32_t n_hosts, const uint64_t* d_hist,
                   uint32_t* d_start, uint32_t* d_cursor, uint32_t* d_order);
/* Memcached binary-protocol request frames, grouped per node: the bytes the cache client puts on
 * each node's socket after placement.  The reference builds its client with binary=True
 * (pylibmc.Client(..., binary=True), Server.py:81-83) and calls get(key) at Server.py:335 and
 * set(key, obj, time=ttl) at :432; pylibmc / libmemcached are third-party and absent, the frames
 * follow the published protocol (24-byte header, big-endian: magic 0x80, opcode, key length,
 * extras length, data type, vbucket, total body length, opaque, CAS).
 *   SHL_MC_GET  header + key                       (opcode 0x00; opaque = request index)
 *   SHL_MC_SET  header + {flags = 0, expiry = ttl} + key   (opcode 0x01); the value (the re-serialised
 *               response) is appended by the caller; value_len[i] (per request index, may be NULL)
 *               is counted in the total body length.
 * Requests are taken in the node-grouped order of shl_group_*: frame j (j-th entry of order[])
 * lies at frames[frame_off[j] : frame_off[j + 1]], so node k's stream is the contiguous range
 * frames[frame_off[start[k]] : frame_off[start[k + 1]]].  shl_mc_frames_batch works on the batch
 * whose shl_route_batch + shl_group_batch results are still resident on this ctx (n = batch size,
 * n_placed = start[n_hosts]).                                                                    */
#define SHL_MC_GET 0
#define SHL_MC_SET 1
typedef struct shl_mc_out {
    uint64_t* frame_off;       /* [n_placed + 1]                                                   */
    uint8_t*  frames;          /* byte arena, `frames_cap` bytes                                   */
    uint64_t  frames_cap;
    uint64_t* frames_used;     /* one element                                                      */
} shl_mc_out;
int  shl_mc_frames_batch(shl_ctx* ctx, uint64_t n, uint64_t n_placed, int op, uint32_t ttl,
                         const uint32_t* value_len, const shl_mc_out* out);
int  shl_mc_frames_dev(shl_ctx* ctx, const uint8_t* d_arena, const uint64_t* d_off, const uint32_t* d_sp1,
                       const uint32_t* d_sp2, const uint32_t* d_order, uint64_t n_placed, int op, uint32_t ttl,
                       const uint32_t* d_value_len, const shl_mc_out* d_out);
/* In every *_batch call `arena` may be NULL: the bytes uploaded by the previous *_batch call on
 * this ctx are reused (parse, then frame the same arena, costs one H2D copy).                    */

/* ------------------------------------------------------------------------------------------
 * Device-resident entry points: same semantics, every data pointer is a DEVICE pointer and no
 * copy is made.  For callers whose arenas already live in HBM, and for kernel-only timing.
 * d_arena must be 32-byte aligned (SHL_ERR_ARG otherwise) and followed by 64 readable bytes: the
 * kernels read whole 32-byte sectors around a message (cudaMalloc'd buffers from
 * shl_device_alloc(bytes + 64) qualify).  Messages may be indexed in any order and may overlap;
 * groups of 32 consecutive messages that lie in index order are scanned fastest.
 * ---------------------------------------------------------------------------------------- */
int  shl_parse_dev(shl_ctx* ctx, const uint8_t* d_arena, const uint64_t* d_off, const uint32_t* d_len,
                   uint64_t n, const shl_parse_out* d_out);
int  shl_frame_dev(shl_ctx* ctx, const uint8_t* d_arena, const uint64_t* d_off, const uint32_t* d_len,
                   uint64_t n, const uint32_t* d_status, const uint32_t* d_first_line_len,
                   const uint32_t* d_headers_len, const shl_frame_out* d_out);
int  shl_place_dev(shl_ctx* ctx, const uint8_t* d_arena, const uint64_t* d_key_off, const uint32_t* d_key_len,
                   uint64_t n, const shl_place_out* d_out);
int  shl_route_dev(shl_ctx* ctx, const uint8_t* d_arena, const uint64_t* d_off, const uint32_t* d_len,
                   uint64_t n, const shl_parse_out* d_pout, const shl_place_out* d_qout);
int  shl_body_dev(shl_ctx* ctx, const uint8_t* d_arena, const uint64_t* d_off, const uint32_t* d_len,
                  uint64_t n, const shl_parse_out* d_pout, const shl_frame_out* d_fout, const shl_body_out* d_out);

/* Placement without any allocation.  It has no counterpart in the reference: it exists for callers
 * of THIS library that cannot call cudaMalloc (the proxy's reactor thread, where an allocation
 * would take the driver's device-wide lock and stall every batch in flight; a one-off "which node
 * owns this key?" probe in a process that never created a ctx; a GPU whose memory is already fully
 * handed out, where an allocation can fail at run time).  Same results as shl_place_batch, but no
 * ctx: keys and ring come in as HOST arrays (the ring as shl_ring_export returns it), results go
 * to HOST arrays (any of md5/hash/node/hist may be NULL; hist is ADDED to), and the device-side
 * working memory is one fixed array inside the library (about 536 KiB held on every GPU that loads
 * it, until the process exits).  Sizes beyond the SHL_STATIC_MAX_* limits are refused with
 * SHL_ERR_ARG and a message in shl_last_error(NULL).  The call returns when the results are in
 * host memory.  The first call on a device also resolves the array's device address; make that
 * first call before any stream capture starts.  All calls share the array, which still holds the
 * previous call's data (the histogram bins are cleared each time): calls whose GPU work overlaps
 * on one device race, so serialise them per device.  The work runs on the device's legacy default
 * stream: it therefore also waits for, and holds up, any BLOCKING stream another library of the
 * process created on that device (streams made with cudaStreamNonBlocking, like every stream of
 * this library, are not affected).  The caller's current device is restored on return.          */
#define SHL_STATIC_MAX_KEYS       4096
#define SHL_STATIC_MAX_KEY_BYTES  262144
#define SHL_STATIC_MAX_POINTS     16384
#define SHL_STATIC_MAX_HOSTS      1024
int  shl_place_static(int device, const uint8_t* arena, uint64_t arena_bytes, const uint64_t* key_off,
                      const uint32_t* key_len, uint64_t n, const uint32_t* ring_values,
                      const uint32_t* ring_nodes, uint32_t n_points, uint32_t n_hosts,
                      uint8_t* md5, uint32_t* hash, uint32_t* node, uint64_t* hist);

/* Number of kernel launches this ctx has issued (for bench.py's gpu_launches claim). */
uint64_t shl_launch_count(shl_ctx* ctx);

/* Per-outcome counters: SHL_N_CODES bins indexed by SHL_CODE(status), accumulated on the device
 * by every parse/route call (the "parse-error counts" that are all-reduced across GPUs next to
 * the per-node placement histogram).  shl_code_hist copies them out (and optionally zeroes
 * them); the *_dev getters expose the device buffers so a collective can read them in place.    */
#define SHL_N_CODES 64
int   shl_code_hist(shl_ctx* ctx, uint64_t* out, int reset);
void* shl_code_hist_dev(shl_ctx* ctx);   /* uint64_t[SHL_N_CODES] = shl_stats_dev + n_hosts (moves with it) */
void* shl_node_hist_dev(shl_ctx* ctx);   /* uint64_t[n_hosts], filled by *_batch calls with hist   */
/* ------------------------------------------------------------------------------------------
 * Counters across GPUs.  The path shards by request batch, one process per GPU, with no exchange
 * of payloads; the ranks of a job only sum their COUNTERS: uint64[n_hosts + SHL_N_CODES], the
 * per-node placement histogram followed by the per-outcome counts (there is
 * no counterpart in the reference, which is one single-threaded process).  The library owns four
 * such device buffers, the NCCL communicator and a side stream.
 *   shl_stats_dev      the CURRENT buffer: hand it to a *_dev call as place.hist so the kernel adds
 *                      its histogram there; the outcome counts of every parse/route call land
 *                      behind it (shl_code_hist_dev points at them).  Changes with every
 *                      shl_stats_allreduce: ask again after each.
 *   shl_comm_unique_id / shl_comm_init   rank 0 makes the 128-byte id, ships it to the others by
 *                      any means (a file, a socket), every rank calls init (collective, blocks).
 *                      Without a communicator everything below still works, on one rank's counters.
 *   shl_stats_allreduce  closes the current buffer: summed over the ranks (ncclAllReduce, uint64,
 *                      sum) on the side stream, ordered after everything queued on the ctx stream,
 *                      result also copied to pinned host memory; the next buffer, zeroed, becomes
 *                      current.  Returns at once: the kernel of the next batch overlaps the reduce.
 *   shl_stats_join     makes the ctx STREAM wait for the outstanding reduces (call it before
 *                      recording the event that ends a timed region).
 *   shl_stats_fetch    host side: waits for the most recent reduce, copies out[0:n].
 *   shl_comm_allreduce_u64  small blocking all-reduce of n <= 64 host values (op 0 = sum, 1 = max)
 *                      after draining this ctx's streams: a barrier, the max of a timing over ranks.
 * ---------------------------------------------------------------------------------------- */
#define SHL_COMM_ID_BYTES 128
void*    shl_stats_dev(shl_ctx* ctx);
uint32_t shl_stats_len(shl_ctx* ctx);                       /* n_hosts + SHL_N_CODES */
int   shl_comm_unique_id(uint8_t* id, uint32_t cap);        /* cap >= SHL_COMM_ID_BYTES */
int   shl_comm_init(shl_ctx* ctx, const uint8_t* id, int rank, int world);
int   shl_comm_size(shl_ctx* ctx);
int   shl_stats_allreduce(shl_ctx* ctx);
int   shl_stats_join(shl_ctx* ctx);
int   shl_stats_fetch(shl_ctx* ctx, uint64_t* out, uint32_t n);
int   shl_comm_allreduce_u64(shl_ctx* ctx, uint64_t* vals, uint32_t n, int op);

/* Test hook: in a library built with -DSHL_BOUNDS_CHECK the kernels count every index that falls
 * outside its shared-memory bitmap / record slot; this returns the count so far (-1: the checks are
 * compiled out, as in the shipped build).  No reference counterpart. */
int   shl_debug_oob_count(void);
/* Test hook: route every message through the scalar (non-vectorised) device path. */
int   shl_set_force_scalar(shl_ctx* ctx, int on);

/* ------------------------------------------------------------------------------------------
 * Synthetic corpora (BASELINE.json configs 2-5; recipe in SURVEY.md section 8d).  A counter-based
 * generator: message i is a pure function of (seed, i), so the host and device versions produce
 * identical bytes.  shl_corpus_lengths fills len[i]; the caller prefix-sums them into off[] and
 * shl_corpus_fill_* writes the bytes.  kind: 0 = GET requests (config 2/4), 1 = keys only (config 5).
 * ---------------------------------------------------------------------------------------- */
int  shl_corpus_lengths_host(uint64_t seed, uint64_t first, uint64_t n, int kind, uint32_t* len);
int  shl_corpus_fill_host(uint64_t seed, uint64_t first, uint64_t n, int kind,
                          const uint64_t* off, uint8_t* arena);
int  shl_corpus_lengths_dev(shl_ctx* ctx, uint64_t seed, uint64_t first, uint64_t n, int kind, uint32_t* d_len);
int  shl_corpus_fill_dev(shl_ctx* ctx, uint64_t seed, uint64_t first, uint64_t n, int kind,
                         const uint64_t* d_off, uint8_t* d_arena);

#ifdef __cplusplus
}
#endif
#endif /* SHELLAC_B200_H */
