"""ctypes binding of libshellac_b200.so (declared in include/shellac_b200.h).

This is the counterpart of the reference's native boundary: Shellac reaches libmemcached through
a Python binding (`import pylibmc`, Server.py:10; README.md:22 calls it "a ctypes wrapper around
libmemcached").  Here the whole hot path (tokenise, frame, key, place) sits behind one C ABI.

There is NO CPU fallback: if the shared library is missing, or there is no CUDA device, loading
raises and every entry point fails loudly.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('SHL_LIB') or os.path.join(_HERE, 'libshellac_b200.so')

# ---------------------------------------------------------------------------- status word
S_OK, S_EMPTY, S_INCOMPLETE_LINE, S_INCOMPLETE_HEADERS, S_INCOMPLETE_BODY = 0, 1, 2, 3, 4
E_FIRSTLINE_PARTS, E_VERSION_SPLIT, E_VERSION_FLOAT, E_STATUS_INT = 16, 17, 18, 19
E_HDR_NO_SEP, E_HDR_MULTI_SEP, E_CLEN_INT, E_CLEN_LIST, E_CHUNK_HEX = 20, 21, 22, 23, 24
E_LIB_HDR_TOO_LARGE, E_LIB_RECORDS_FULL = 32, 33
F_IS_REQUEST, F_HEADERS_COMPLETE, F_MESSAGE_COMPLETE = 1 << 8, 1 << 9, 1 << 10
F_ON_BODY, F_CHUNKED = 1 << 11, 1 << 12
F_VERSION_SHIFT = 13
F_STATUS_WIDE, F_CLEN_WIDE, F_FIRST_LINE_DONE = 1 << 15, 1 << 16, 1 << 17
F_KEY_TOO_LONG = 1 << 18
MAX_KEY_LEN = 250
FS_NOT_CHUNKED = 0xff
RING_KETAMA, RING_KETAMA_WEIGHTED = 0, 1
MC_GET, MC_SET = 0, 1
N_CODES = 64          # SHL_N_CODES: histogram length for per-outcome counts
COMM_ID_BYTES = 128   # SHL_COMM_ID_BYTES

CODE_NAMES = {
    0: 'OK', 1: 'EMPTY', 2: 'INCOMPLETE_LINE', 3: 'INCOMPLETE_HEADERS', 4: 'INCOMPLETE_BODY',
    16: 'E_FIRSTLINE_PARTS', 17: 'E_VERSION_SPLIT', 18: 'E_VERSION_FLOAT', 19: 'E_STATUS_INT',
    20: 'E_HDR_NO_SEP', 21: 'E_HDR_MULTI_SEP', 22: 'E_CLEN_INT', 23: 'E_CLEN_LIST',
    24: 'E_CHUNK_HEX', 32: 'E_LIB_HDR_TOO_LARGE', 33: 'E_LIB_RECORDS_FULL',
}
# The Python exception type the reference raises for each error code (only the TYPE is portable
# between py2 and py3).
CODE_EXC = {16: ValueError, 17: ValueError, 18: ValueError, 19: ValueError, 20: ValueError,
            21: ValueError, 22: ValueError, 23: TypeError, 24: ValueError,
            32: OverflowError, 33: MemoryError}

HDR_DTYPE = np.dtype([('key_off', '<u2'), ('key_len', '<u2'), ('val_off', '<u2'), ('val_len', '<u2')])
REC_DTYPE = np.dtype([('status', '<u4'), ('hdr_base', '<u4'), ('content_len', '<i8'), ('first_line_len', '<u2'),
                      ('headers_len', '<u2'), ('sp1', '<u2'), ('sp2', '<u2'), ('fl_end', '<u2'), ('slash', '<u2'),
                      ('n_headers', '<u2'), ('tail', '<u2')])      # shl_msg_rec
CHUNK_DTYPE = np.dtype([('hdr_off', '<u4'), ('hdr_len', '<u4'), ('data_off', '<u4'), ('data_len', '<u4')])

_u8p, _u32p, _i32p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_int32)
_u64p, _i64p = C.POINTER(C.c_uint64), C.POINTER(C.c_int64)


class ParseOutStruct(C.Structure):
    _fields_ = [('status', C.c_void_p), ('first_line_len', C.c_void_p), ('headers_len', C.c_void_p),
                ('sp1', C.c_void_p), ('sp2', C.c_void_p), ('fl_end', C.c_void_p), ('slash', C.c_void_p),
                ('status_code', C.c_void_p), ('n_headers', C.c_void_p), ('hdr_base', C.c_void_p),
                ('content_len', C.c_void_p), ('hdr', C.c_void_p), ('hdr_cap', C.c_uint64),
                ('hdr_used', C.c_void_p), ('lkeys', C.c_void_p), ('lkeys_cap', C.c_uint64),
                ('lkeys_used', C.c_void_p), ('lkey_base', C.c_void_p)]


class FrameOutStruct(C.Structure):
    _fields_ = [('fstatus', C.c_void_p), ('n_chunks', C.c_void_p), ('chunk_base', C.c_void_p),
                ('body_end', C.c_void_p), ('body_bytes', C.c_void_p), ('last_call', C.c_void_p),
                ('chunk', C.c_void_p), ('chunk_cap', C.c_uint64), ('chunk_used', C.c_void_p)]


class BodyOutStruct(C.Structure):
    _fields_ = [('body_off', C.c_void_p), ('body_len', C.c_void_p), ('body', C.c_void_p),
                ('body_cap', C.c_uint64), ('body_used', C.c_void_p)]


class McOutStruct(C.Structure):
    _fields_ = [('frame_off', C.c_void_p), ('frames', C.c_void_p), ('frames_cap', C.c_uint64), ('frames_used', C.c_void_p)]


class PlaceOutStruct(C.Structure):
    _fields_ = [('md5', C.c_void_p), ('hash', C.c_void_p), ('node', C.c_void_p), ('hist', C.c_void_p)]


PARSE_FIELDS = (('status', np.uint32), ('first_line_len', np.uint32), ('headers_len', np.uint32),
                ('sp1', np.uint32), ('sp2', np.uint32), ('fl_end', np.uint32), ('slash', np.uint32),
                ('status_code', np.int32), ('n_headers', np.uint32), ('hdr_base', np.uint32),
                ('content_len', np.int64))
FRAME_FIELDS = (('fstatus', np.uint32), ('n_chunks', np.uint32), ('chunk_base', np.uint32),
                ('body_end', np.uint64), ('body_bytes', np.uint64), ('last_call', np.uint32))

_lib = None


class ShellacError(RuntimeError):
    """A library / CUDA failure (negative return code), never a per-message parse outcome."""


def lib():
    """Load libshellac_b200.so once.  Raises if it was not built: build it with
    `python -m shellac_b200.build` (or __graft_entry__.build())."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ShellacError('%s not found: build it (python -m shellac_b200.build); '
                           'there is no CPU fallback' % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, u64, i32 = C.c_void_p, C.c_uint64, C.c_int
    sig = {
        'shl_device_count': (i32, []),
        'shl_device_mem_info': (i32, [i32, _u64p, _u64p]),
        'shl_device_pci_bus_id': (i32, [i32, C.c_char_p, i32]),
        'shl_ctx_create': (i32, [i32, u64, u64, u64, C.POINTER(vp)]),
        'shl_ctx_destroy': (None, [vp]),
        'shl_last_error': (C.c_char_p, [vp]),
        'shl_sync': (i32, [vp]),
        'shl_pinned_alloc': (i32, [C.POINTER(vp), u64]),
        'shl_pinned_free': (i32, [vp]),
        'shl_device_alloc': (i32, [vp, C.POINTER(vp), u64]),
        'shl_device_free': (i32, [vp, vp]),
        'shl_memcpy_h2d': (i32, [vp, vp, vp, u64]),
        'shl_memcpy_d2h': (i32, [vp, vp, vp, u64]),
        'shl_memcpy_d2d': (i32, [vp, vp, vp, u64]),
        'shl_memset_dev': (i32, [vp, vp, i32, u64]),
        'shl_event_record': (i32, [vp, i32]),
        'shl_event_elapsed_ms': (i32, [vp, i32, i32, C.POINTER(C.c_float)]),
        'shl_stream': (vp, [vp]),
        'shl_ring_build': (i32, [vp, C.POINTER(C.c_char_p), vp, vp, C.c_uint32, i32]),
        'shl_ring_size': (i32, [vp, _u32p, _u32p]),
        'shl_ring_export': (i32, [vp, vp, vp, C.c_uint32]),
        'shl_parse_batch': (i32, [vp, vp, u64, vp, vp, u64, C.POINTER(ParseOutStruct)]),
        'shl_frame_batch': (i32, [vp, vp, u64, vp, vp, u64, vp, vp, vp, C.POINTER(FrameOutStruct)]),
        'shl_place_batch': (i32, [vp, vp, u64, vp, vp, u64, C.POINTER(PlaceOutStruct)]),
        'shl_route_batch': (i32, [vp, vp, u64, vp, vp, u64, C.POINTER(ParseOutStruct), C.POINTER(PlaceOutStruct)]),
        'shl_body_batch': (i32, [vp, u64, C.POINTER(BodyOutStruct)]),
        'shl_body_dev': (i32, [vp, vp, vp, vp, u64, C.POINTER(ParseOutStruct), C.POINTER(FrameOutStruct), C.POINTER(BodyOutStruct)]),
        'shl_group_batch': (i32, [vp, u64, vp, vp]),
        'shl_group_dev': (i32, [vp, vp, u64, C.c_uint32, vp, vp, vp, vp]),
        'shl_mc_frames_batch': (i32, [vp, u64, u64, i32, C.c_uint32, vp, C.POINTER(McOutStruct)]),
        'shl_mc_frames_dev': (i32, [vp, vp, vp, vp, vp, vp, u64, i32, C.c_uint32, vp, C.POINTER(McOutStruct)]),
        'shl_chunk_header_batch': (i32, [vp, vp, u64, vp, vp, u64, vp, vp, vp]),
        'shl_parse_dev': (i32, [vp, vp, vp, vp, u64, C.POINTER(ParseOutStruct)]),
        'shl_frame_dev': (i32, [vp, vp, vp, vp, u64, vp, vp, vp, C.POINTER(FrameOutStruct)]),
        'shl_place_dev': (i32, [vp, vp, vp, vp, u64, C.POINTER(PlaceOutStruct)]),
        'shl_route_dev': (i32, [vp, vp, vp, vp, u64, C.POINTER(ParseOutStruct), C.POINTER(PlaceOutStruct)]),
        'shl_route_compact_batch': (i32, [vp, vp, u64, vp, vp, u64, vp, vp, u64, vp, vp, vp]),
        'shl_route_compact_dev': (i32, [vp, vp, vp, vp, u64, vp, vp, u64, vp, vp, vp]),
        'shl_place_static': (i32, [i32, vp, u64, vp, vp, u64, vp, vp, C.c_uint32, C.c_uint32, vp, vp, vp, vp]),
        'shl_launch_count': (u64, [vp]),
        'shl_code_hist': (i32, [vp, vp, i32]),
        'shl_code_hist_dev': (vp, [vp]),
        'shl_node_hist_dev': (vp, [vp]),
        'shl_set_force_scalar': (i32, [vp, i32]),
        'shl_debug_oob_count': (i32, []),
        'shl_stats_dev': (vp, [vp]),
        'shl_stats_len': (C.c_uint32, [vp]),
        'shl_comm_unique_id': (i32, [vp, C.c_uint32]),
        'shl_comm_init': (i32, [vp, vp, i32, i32]),
        'shl_comm_size': (i32, [vp]),
        'shl_stats_allreduce': (i32, [vp]),
        'shl_stats_join': (i32, [vp]),
        'shl_stats_fetch': (i32, [vp, vp, C.c_uint32]),
        'shl_comm_allreduce_u64': (i32, [vp, vp, C.c_uint32, i32]),
        'shl_corpus_lengths_host': (i32, [u64, u64, u64, i32, vp]),
        'shl_corpus_fill_host': (i32, [u64, u64, u64, i32, vp, vp]),
        'shl_corpus_lengths_dev': (i32, [vp, u64, u64, u64, i32, vp]),
        'shl_corpus_fill_dev': (i32, [vp, u64, u64, u64, i32, vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)          # AttributeError here == header/library mismatch: loud
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


EXPORTS = None   # filled lazily by exported_symbols()


def exported_symbols():
    """Names include/shellac_b200.h declares (parsed from the header), for the load test."""
    import re
    hdr = os.path.join(os.path.dirname(_HERE), 'include', 'shellac_b200.h')
    text = open(hdr).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(shl_[a-z0-9_]+)\s*\(', text)))


def ptr(a):
    """Address of a numpy array's first element (or 0 for None)."""
    if a is None:
        return None
    return a.ctypes.data
