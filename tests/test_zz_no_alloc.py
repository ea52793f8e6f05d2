"""shl_place_static: the placement entry point for callers that are not allowed to allocate.

Who that is: code on the proxy's reactor thread (a cudaMalloc there takes the driver's device-wide
lock and stalls every batch in flight), a one-off "which node owns this key?" probe in a process
that never created a ctx, or a box whose GPU memory is already fully handed out.  The entry point
keeps all of its device-side working memory in one fixed array inside the library, so this test
drives it WITHOUT an Engine of its own for the calls under test and checks that what it returns
is bit-identical to the regular, ctx-based shl_place_batch on the same keys and ring.

Two calls are made back to back on purpose: the array still holds the first call's keys, results
and histogram when the second call starts, and the second call (fewer, different keys) must not
see any of it.  The file name sorts last so a `pytest -x` run has already executed everything else
when it gets here; it must also pass alone in a fresh process.
"""
import ctypes as C
import hashlib

import numpy as np
import pytest

from shellac_b200 import _native as N
from shellac_b200.batch import Arena

pytestmark = pytest.mark.gpu


def _static_place(lib, ar, vals, nodes, n_hosts):
    n = ar.n
    md5 = np.full(16 * max(n, 1), 0xAA, np.uint8)
    hsh = np.full(max(n, 1), 0xAAAAAAAA, np.uint32)
    node = np.full(max(n, 1), 0xAAAAAAAA, np.uint32)
    hist = np.zeros(n_hosts, np.uint64)
    rc = lib.shl_place_static(0, ar.data.ctypes.data, ar.nbytes, ar.off.ctypes.data, ar.len.ctypes.data, n,
                              vals.ctypes.data, nodes.ctypes.data, len(vals), n_hosts,
                              md5.ctypes.data, hsh.ctypes.data, node.ctypes.data, hist.ctypes.data)
    assert rc == 0, lib.shl_last_error(None)
    return md5[:16 * n], hsh[:n], node[:n], hist


def test_static_placement_matches_ctx_placement_twice_in_a_row():
    from shellac_b200 import build as B
    from shellac_b200.engine import Engine
    B.build()
    lib = N.lib()
    hosts = ['10.0.0.%d' % (i + 1) for i in range(16)]
    # expected values from the existing entry point (its own ctx, closed before the calls under test)
    eng = Engine(device=0, max_msgs=1 << 13, max_bytes=1 << 21)
    eng.ring_build(hosts)
    vals, nodes = (x.copy() for x in eng.ring_export())
    rng = np.random.default_rng(20130730)
    batches = []
    for n in (4000, 1500):                                   # second call smaller: stale tail must not leak
        keys = [b'/' + rng.integers(97, 123, int(rng.integers(1, 60)), dtype=np.uint8).tobytes() for _ in range(n)]
        keys[0] = b''                                        # empty key is legal for MD5
        ar = Arena.pack(keys)                                # plain (pageable) numpy memory
        ref = eng.place(eng.pack(keys))
        batches.append((ar, ref.md5[:16 * n].copy(), ref.hash[:n].copy(), ref.node[:n].copy(), ref.hist.copy()))
    eng.close()
    free0 = _free_bytes(lib)
    for ar, e_md5, e_hash, e_node, e_hist in batches:        # two calls in a row through the C ABI
        md5, hsh, node, hist = _static_place(lib, ar, vals, nodes, len(hosts))
        assert (md5 == e_md5).all()
        assert (hsh == e_hash).all()
        assert (node == e_node).all()
        assert hist.tolist() == e_hist.tolist() and int(hist.sum()) == ar.n
        # ... and against an expectation that shares no code with the library: hashlib + searchsorted
        keys = [ar.message(i) for i in range(ar.n)]
        dig = [hashlib.md5(k).digest() for k in keys]
        assert md5.tobytes() == b''.join(dig)
        h = np.array([int.from_bytes(d[:4], 'little') for d in dig], np.uint32)
        j = np.searchsorted(vals, h, side='left')
        assert (node == nodes[np.where(j == len(vals), 0, j)]).all() and (hsh == h).all()
    # the property in the entry point's name: the calls left the device's free memory where it was
    assert _free_bytes(lib) == free0


def _free_bytes(lib):
    free, total = C.c_uint64(), C.c_uint64()
    assert lib.shl_device_mem_info(0, C.byref(free), C.byref(total)) == 0
    return free.value


def test_static_placement_edges():
    """No keys at all, a ring of one host, and a histogram that already holds counts (it is added to)."""
    from shellac_b200 import build as B
    B.build()
    lib = N.lib()
    vals = (np.arange(1, 201, dtype=np.uint64) * 21000000).astype(np.uint32)
    nodes = (np.arange(200) % 4).astype(np.uint32)
    keys = [b'/k%d' % i for i in range(300)]
    ar = Arena.pack(keys)
    md5, hsh, node, hist = _static_place(lib, ar, vals, nodes, 4)
    again = np.full(4, 7, np.uint64)                          # pre-filled: counts are ADDED
    rc = lib.shl_place_static(0, ar.data.ctypes.data, ar.nbytes, ar.off.ctypes.data, ar.len.ctypes.data, ar.n,
                              vals.ctypes.data, nodes.ctypes.data, len(vals), 4, None, None, None, again.ctypes.data)
    assert rc == 0 and (again == hist + 7).all()
    h = np.array([int.from_bytes(hashlib.md5(k).digest()[:4], 'little') for k in keys], np.uint32)
    j = np.searchsorted(vals, h, side='left')
    assert (node == nodes[np.where(j == len(vals), 0, j)]).all()
    empty = Arena.pack([])
    e_hist = np.full(4, 3, np.uint64)
    rc = lib.shl_place_static(0, ar.data.ctypes.data, 0, empty.off.ctypes.data, empty.len.ctypes.data, 0,
                              vals.ctypes.data, nodes.ctypes.data, len(vals), 4, None, None, None, e_hist.ctypes.data)
    assert rc == 0 and (e_hist == 3).all()                    # nothing placed, nothing touched
    one = np.zeros(len(vals), np.uint32)
    _, _, node1, hist1 = _static_place(lib, ar, vals, one, 1)  # libmemcached: a single host takes every key
    assert (node1 == 0).all() and hist1.tolist() == [len(keys)]


def test_static_placement_refuses_oversize_inputs():
    from shellac_b200 import build as B
    B.build()
    lib = N.lib()
    vals = np.arange(1, 101, dtype=np.uint32) * 1000
    nodes = np.zeros(100, np.uint32)
    big = Arena.pack([b'k'] * 4097)
    rc = lib.shl_place_static(0, big.data.ctypes.data, big.nbytes, big.off.ctypes.data, big.len.ctypes.data, big.n,
                              vals.ctypes.data, nodes.ctypes.data, 100, 1, None, None, None, None)
    assert rc == -2 and b'SHL_STATIC_MAX_KEYS' in lib.shl_last_error(None)
    fat = Arena.pack([b'x' * 200000, b'y' * 100000])
    rc = lib.shl_place_static(0, fat.data.ctypes.data, fat.nbytes, fat.off.ctypes.data, fat.len.ctypes.data, fat.n,
                              vals.ctypes.data, nodes.ctypes.data, 100, 1, None, None, None, None)
    assert rc == -2 and b'SHL_STATIC_MAX_KEY_BYTES' in lib.shl_last_error(None)
    ok = Arena.pack([b'/a'])
    rc = lib.shl_place_static(0, ok.data.ctypes.data, ok.nbytes, ok.off.ctypes.data, ok.len.ctypes.data, ok.n,
                              vals.ctypes.data, nodes.ctypes.data, 20000, 1, None, None, None, None)
    assert rc == -2 and b'points' in lib.shl_last_error(None)
