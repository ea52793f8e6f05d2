"""Test-side loader for oracle/ (the CPU restatement).  TEST INFRASTRUCTURE — never imported by
the shellac_b200 package."""
import ctypes as C
import gzip
import json
import os
import subprocess

import numpy as np

from shellac_b200 import _native as N
from shellac_b200.batch import Arena, BodyOut, FrameOut, MessageView, ParseOut, PlaceOut

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORC_SRC = os.path.join(ROOT, 'oracle', 'shl_oracle.c')
ORC_LIB = os.path.join(ROOT, 'oracle', '_build', 'liboracle.so')
GOLDEN = os.path.join(ROOT, 'tests', 'golden', 'parser_golden.json.gz')

_orc = None


def build_oracle(force=False):
    if force or not os.path.exists(ORC_LIB) or os.path.getmtime(ORC_LIB) < os.path.getmtime(ORC_SRC):
        os.makedirs(os.path.dirname(ORC_LIB), exist_ok=True)
        subprocess.check_call(['gcc', '-O2', '-shared', '-fPIC', '-o', ORC_LIB, ORC_SRC, '-lm'])
    return ORC_LIB


def oracle():
    global _orc
    if _orc is None:
        L = C.CDLL(build_oracle())
        vp, u64 = C.c_void_p, C.c_uint64
        L.orc_parse.argtypes = [vp, vp, vp, u64, C.POINTER(N.ParseOutStruct)]
        L.orc_frame.argtypes = [vp, vp, vp, u64, vp, vp, vp, C.POINTER(N.FrameOutStruct)]
        L.orc_body.argtypes = [vp, vp, vp, u64, C.POINTER(N.ParseOutStruct), C.POINTER(N.FrameOutStruct), C.POINTER(N.BodyOutStruct)]
        L.orc_chunk_header.argtypes = [vp, vp, vp, u64, vp, vp, vp]
        L.orc_md5.argtypes = [vp, u64, vp]
        L.orc_ring_build.argtypes = [C.POINTER(C.c_char_p), vp, vp, C.c_uint32, C.c_int, vp, vp, C.c_uint32]
        L.orc_place.argtypes = [vp, vp, vp, u64, vp, vp, C.c_uint32, C.c_uint32, C.POINTER(N.PlaceOutStruct)]
        _orc = L
    return _orc


def orc_parse(arena, hdr_cap=None, lkeys=False):
    n = arena.n
    out = ParseOut(n, hdr_cap if hdr_cap is not None else max(16, arena.nbytes // 4 + 16),
                   lkeys_cap=arena.nbytes + 16 if lkeys else 0)
    s = out.struct()
    oracle().orc_parse(arena.data.ctypes.data, arena.off.ctypes.data, arena.len.ctypes.data, n, C.byref(s))
    return out


def orc_frame(arena, pout, chunk_cap=None):
    n = arena.n
    out = FrameOut(n, chunk_cap if chunk_cap is not None else max(16, arena.nbytes // 5 + 16))
    s = out.struct()
    oracle().orc_frame(arena.data.ctypes.data, arena.off.ctypes.data, arena.len.ctypes.data, n,
                       pout.status.ctypes.data, pout.first_line_len.ctypes.data,
                       pout.headers_len.ctypes.data, C.byref(s))
    return out


def orc_body(arena, pout, fout):
    out = BodyOut(arena.n, arena.nbytes + 16 * arena.n + 64)
    ps, fs, bs = pout.struct(), fout.struct(), out.struct()
    oracle().orc_body(arena.data.ctypes.data, arena.off.ctypes.data, arena.len.ctypes.data, arena.n,
                      C.byref(ps), C.byref(fs), C.byref(bs))
    return out


def orc_chunk_header(arena):
    n = arena.n
    hl, sz, ok = np.zeros(max(n, 1), np.int32), np.zeros(max(n, 1), np.uint64), np.zeros(max(n, 1), np.uint8)
    oracle().orc_chunk_header(arena.data.ctypes.data, arena.off.ctypes.data, arena.len.ctypes.data, n,
                              hl.ctypes.data, sz.ctypes.data, ok.ctypes.data)
    return hl[:n], sz[:n], ok[:n]


def orc_md5(data):
    out = np.zeros(16, np.uint8)
    buf = np.frombuffer(bytes(data) + b'\0', dtype=np.uint8)
    oracle().orc_md5(buf.ctypes.data, len(data), out.ctypes.data)
    return out.tobytes()


def orc_ring(hosts, ports, weights=None, mode=0):
    n = len(hosts)
    arr = (C.c_char_p * n)(*[h.encode() for h in hosts])
    p = np.asarray(ports, dtype=np.uint32)
    w = np.asarray(weights, dtype=np.uint32) if weights is not None else None
    cap = n * 640 + 16
    vals, nodes = np.zeros(cap, np.uint32), np.zeros(cap, np.uint32)
    k = oracle().orc_ring_build(arr, p.ctypes.data, w.ctypes.data if w is not None else None, n, mode,
                                vals.ctypes.data, nodes.ctypes.data, cap)
    assert k >= 0
    return vals[:k].copy(), nodes[:k].copy()


def orc_place(arena, vals, nodes, n_hosts, key_off=None, key_len=None):
    key_off = arena.off if key_off is None else key_off
    key_len = arena.len if key_len is None else key_len
    n = len(key_off)
    out = PlaceOut(n, n_hosts)
    s = out.struct()
    oracle().orc_place(arena.data.ctypes.data, key_off.ctypes.data, key_len.ctypes.data, n,
                       vals.ctypes.data, nodes.ctypes.data, len(vals), n_hosts, C.byref(s))
    return out


def load_golden():
    with gzip.open(GOLDEN) as f:
        return json.loads(f.read().decode('ascii'))


def expected_content_len(v):
    """Final HttpParser._content_len for a NON-chunked message, from a MessageView."""
    if not v.headers_complete() or v.code in (N.E_CLEN_INT, N.E_CLEN_LIST):
        return None
    cl = v.content_len()
    if v._body_call():
        rem = len(v.m) - v.body_off()
        return 0 if v.message_complete() else cl - rem
    return cl


def view_observation(v):
    """The same record tests/golden/make_golden.py:observe() builds from the reference object."""
    exc = v.exception()
    ver = v.version()
    return {
        'calls': v.calls(), 'exc': exc.__name__ if exc else None,
        'is_request': v.is_request(), 'method': v.method(), 'url': v.url(),
        'version': repr(ver) if ver is not None else None,
        'status': v.status(), 'message': v.message(), 'headers': v.headers(),
        'headers_complete': v.headers_complete(), 'message_complete': v.message_complete(),
        'on_body': v.on_body(), 'chunked': v.chunked(),
        'body': v.body().getvalue(),
    }


def diff_observation(v, obs):
    got = view_observation(v)
    bad = {k: (got[k], obs[k]) for k in got if got[k] != obs[k]}
    if not v.chunked():
        e = expected_content_len(v)
        if e != obs['content_len']:
            bad['content_len'] = (e, obs['content_len'])
    return bad
