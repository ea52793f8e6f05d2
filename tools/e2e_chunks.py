#!/usr/bin/env python
"""End-to-end rate of shl_route_compact_batch on the config-2 corpus for the chunking in the
environment (SHL_CHUNK_MSGS messages per chunk, SHL_MAX_CHUNKS chunks per batch)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from shellac_b200.engine import Engine  # noqa: E402
n = 1 << 20
eng = Engine(device=0, max_msgs=n, max_bytes=n * 640, max_records=n * 24)
eng.ring_build(['10.0.0.%d' % (i + 1) for i in range(16)])
ar = eng.corpus_host(20130725, 0, n, 0)
out = eng.route_compact(ar, hdr_cap=n * 24)
ts = []
for _ in range(12):
    out[4][:] = 0
    t0 = time.perf_counter()
    eng.route_compact(ar, out=out, sync=True)
    ts.append(time.perf_counter() - t0)
print('chunk_msgs=%s max_chunks=%s  %.3f ms  %.1f M requests/s' % (os.environ.get('SHL_CHUNK_MSGS', '65536'), os.environ.get('SHL_MAX_CHUNKS', '16'), np.median(ts[2:]) * 1e3, n / np.median(ts[2:]) / 1e6))
