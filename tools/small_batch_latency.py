#!/usr/bin/env python
"""Per-call latency of the host-buffer entry point for reactor-sized batches (one epoll tick)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from shellac_b200.batch import ParseOut, PlaceOut  # noqa: E402
from shellac_b200.engine import Engine  # noqa: E402

eng = Engine(device=0, max_msgs=1 << 16, max_bytes=1 << 26)
hosts = ['10.0.0.%d' % (i + 1) for i in range(16)]
eng.ring_build(hosts)
for n in (1, 32, 256, 1024, 4096, 16384):
    ar = eng.corpus_host(20130725, 0, n, 0)
    pout = ParseOut(n, n * 24, alloc=eng.pinned)
    qout = PlaceOut(n, len(hosts), alloc=eng.pinned)
    for _ in range(20):
        eng.route(ar, pout=pout, qout=qout, sync=True)
    ts = []
    for _ in range(200):
        t0 = time.perf_counter()
        eng.route(ar, pout=pout, qout=qout, sync=True)
        ts.append(time.perf_counter() - t0)
    ts = np.array(ts) * 1e6
    print('n=%6d  median %.1f us  p90 %.1f us  -> %.2f M req/s' % (n, np.median(ts), np.percentile(ts, 90), n / np.median(ts)))

# the same call without the Python wrapper's per-call work (structs built once)
import ctypes as C
n = 256
ar = eng.corpus_host(20130725, 0, n, 0)
pout = ParseOut(n, n * 24, alloc=eng.pinned)
qout = PlaceOut(n, len(hosts), alloc=eng.pinned)
ps, qs = pout.struct(), qout.struct()
args = (eng.ctx, ar.data.ctypes.data, ar.nbytes + 64, ar.off.ctypes.data, ar.len.ctypes.data, n, C.byref(ps), C.byref(qs))
for _ in range(50):
    assert eng.lib.shl_route_batch(*args) == 0
ts = []
for _ in range(300):
    t0 = time.perf_counter()
    eng.lib.shl_route_batch(*args)
    ts.append(time.perf_counter() - t0)
print('n=%d bare C ABI call: median %.1f us' % (n, np.median(np.array(ts) * 1e6)))

# the reactor shim's router (what one epoll tick pays end to end, packing included)
from shellac_b200.reactor import EngineRouter
router = EngineRouter(eng)
for n in (1, 64, 512):
    msgs = [ar.message(i % ar.n) for i in range(n)]
    for _ in range(5):
        router.route(msgs)
    ts = []
    for _ in range(50):
        t0 = time.perf_counter()
        router.route(msgs)
        ts.append(time.perf_counter() - t0)
    print('EngineRouter.route n=%d: median %.1f us' % (n, np.median(np.array(ts) * 1e6)))
