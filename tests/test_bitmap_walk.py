"""The route kernel's bitmap walker (shellac_b200/csrc/bitmap_walk.cuh) run on the CPU against the
oracle: tests/bm_emu.cpp compiles the SAME header the kernel includes and feeds it bitmaps built by
a scalar loop.  What is checked per message: end of the first line (HttpParser.py:155), end of the
header block (:176), number of header lines tokenised before the first line that does not split
into exactly two parts on ': ' and which of the two errors it is (:326-328), and the line start /
separator / line end positions the records are made from."""
import ctypes as C
import os
import subprocess

import numpy as np

from shellac_b200 import _native as N
from shellac_b200.batch import Arena
from tests import oracle_lib as O

SRC = os.path.join(O.ROOT, 'tests', 'bm_emu.cpp')
HDR = os.path.join(O.ROOT, 'shellac_b200', 'csrc', 'bitmap_walk.cuh')
LIB = os.path.join(O.ROOT, 'oracle', '_build', 'libbm_emu.so')
WS = frozenset(b'\t\n\x0b\x0c\r ')


def _emu():
    if not os.path.exists(LIB) or os.path.getmtime(LIB) < max(os.path.getmtime(SRC), os.path.getmtime(HDR)):
        os.makedirs(os.path.dirname(LIB), exist_ok=True)
        subprocess.check_call(['g++', '-O2', '-std=c++17', '-shared', '-fPIC', '-x', 'c++', SRC, '-o', LIB])
    L = C.CDLL(LIB)
    L.bm_emu_group.argtypes = [C.c_void_p] * 3 + [C.c_uint32] + [C.c_void_p] * 7 + [C.c_uint32, C.c_void_p]
    L.bm_emu_group.restype = C.c_int
    return L


def _check(msgs, group=32, align=1):
    L = _emu()
    ar = Arena.pack(msgs, align=align)
    o = O.orc_parse(ar)
    n = ar.n
    idx, T = np.zeros(n, np.int32), np.zeros(n, np.int32)
    n_c, n_hdr, err, base = (np.zeros(n, np.uint32) for _ in range(4))
    cap = ar.nbytes // 2 + 64
    lines, used = np.zeros(3 * cap, np.uint32), np.zeros(1, np.uint32)
    for i0 in range(0, n, group):
        g = min(group, n - i0)
        rc = L.bm_emu_group(ar.data.ctypes.data, ar.off[i0:].ctypes.data, ar.len[i0:].ctypes.data, g,
                            idx[i0:].ctypes.data, T[i0:].ctypes.data, n_c[i0:].ctypes.data, n_hdr[i0:].ctypes.data,
                            err[i0:].ctypes.data, base[i0:].ctypes.data, lines.ctypes.data, cap, used.ctypes.data)
        assert rc == 0
    checked_lines = 0
    for i in range(n):
        m = ar.message(i)
        code = int(o.status[i]) & 0xff
        if code == N.S_EMPTY:
            continue
        if code == N.S_INCOMPLETE_LINE:
            assert idx[i] < 0, (i, m[:80])
            continue
        assert idx[i] + 2 == int(o.first_line_len[i]), (i, m[:80], idx[i])
        if code in (N.E_FIRSTLINE_PARTS, N.E_VERSION_SPLIT, N.E_VERSION_FLOAT, N.E_STATUS_INT):
            continue
        if code == N.S_INCOMPLETE_HEADERS:
            assert T[i] < 0, (i, m[:120])
            continue
        if code == N.E_LIB_HDR_TOO_LARGE:
            continue
        h0 = idx[i] + 2
        assert T[i] + 4 - h0 == int(o.headers_len[i]), (i, m[:120], T[i])
        assert n_c[i] >= 1
        if m[h0] in WS or m[T[i] - 1] in WS:
            continue                      # strip() moves the block's ends: the kernel's scalar path tokenises it
        assert int(n_hdr[i]) == int(o.n_headers[i]), (i, m[:200], n_hdr[i], o.n_headers[i])
        want_err = code if code in (N.E_HDR_NO_SEP, N.E_HDR_MULTI_SEP) else 0
        assert int(err[i]) == want_err, (i, m[:200], err[i], code)
        hb = int(o.hdr_base[i])
        for k in range(int(n_hdr[i])):
            ls, sep, end = (int(x) for x in lines[3 * (int(base[i]) + k):3 * (int(base[i]) + k) + 3])
            r = o.hdr[hb + k]
            assert (ls, sep - ls) == (int(r['key_off']), int(r['key_len'])), (i, k, m[:200])
            assert m[sep:sep + 2] == b': ' and m[end:end + 2] == b'\r\n'
            assert sep + 2 <= int(r['val_off']) and int(r['val_off']) + int(r['val_len']) <= end
            checked_lines += 1
    return checked_lines


def test_walker_on_golden_whole_messages(golden):
    msgs = [w['msg'] for w in golden['whole']]
    assert _check(msgs) > 8000
    assert _check(msgs[::-1][:2000], group=7) > 1000          # other group sizes: other word alignments
    assert _check(msgs[:3000], group=32, align=32) > 1000     # gaps between the messages of a span


def test_walker_on_config2_corpus():
    lib = N.lib()
    n = 4096
    ln = np.zeros(n, np.uint32)
    lib.shl_corpus_lengths_host(20130725, 0, n, 0, ln.ctypes.data)
    off = np.zeros(n, np.uint64)
    np.cumsum(ln[:-1], dtype=np.uint64, out=off[1:])
    data = np.zeros(int(off[-1] + ln[-1]) + 64, np.uint8)
    lib.shl_corpus_fill_host(20130725, 0, n, 0, off.ctypes.data, data.ctypes.data)
    blob = data.tobytes()
    msgs = [blob[int(o):int(o) + int(l)] for o, l in zip(off, ln)]
    assert _check(msgs) == int(O.orc_parse(Arena.pack(msgs)).n_headers[:n].sum())


def test_walker_edges():
    """Pairs that straddle words / messages, terminators at every bit position, truncations."""
    base = b'GET /x HTTP/1.1\r\nA: b\r\nCc: dd\r\n\r\n'
    msgs = []
    for pad in range(0, 70):
        msgs.append(b'GET /' + b'u' * pad + b' HTTP/1.1\r\nA: b\r\nCc: dd\r\n\r\n')
    for cut in range(len(base) + 1):
        msgs.append(base[:cut])
    msgs += [b'\r\n', b'\r\n\r\n', b'\r\n\r\n\r\n', b'GET / HTTP/1.1\r\n\r\n', b'GET / HTTP/1.1\r\n\r\n\r\n',
             b'GET / HTTP/1.1\r\n\r\nA: b\r\n\r\n', b'GET / HTTP/1.1\r\nA: b: c\r\n\r\n', b'GET / HTTP/1.1\r\nA:b\r\nC: d\r\n\r\n',
             b'GET / HTTP/1.1\r\nA: b\r\nC:d\r\nE: f: g\r\n\r\n', b'GET /a: b HTTP/1.1\r\nA: b\r\n\r\n',
             b'GET / HTTP/1.1\r\nA: b\r\n\r\nBODY: x\r\n\r\n', b': \r\n: \r\n\r\n', b'\r', b'\n', b':', b'']
    assert _check(msgs) > 100
    assert _check(msgs, group=1) > 100
    assert _check(msgs, group=5, align=16) > 100
