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
    for ar, e_md5, e_hash, e_node, e_hist in batches:        # two calls in a row through the C ABI
        md5, hsh, node, hist = _static_place(lib, ar, vals, nodes, len(hosts))
        assert (md5 == e_md5).all()
        assert (hsh == e_hash).all()
        assert (node == e_node).all()
        assert hist.tolist() == e_hist.tolist() and int(hist.sum()) == ar.n


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
