#!/usr/bin/env python
"""Kernel-only timing of shl_route_dev on the config-2 corpus (for A/B-ing library builds:
SHL_LIB=path/to/variant.so python tools/quick_kernel_time.py [n_messages])."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from shellac_b200.engine import Engine  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 19
eng = Engine(device=0, max_msgs=1 << 12, max_bytes=1 << 20)
eng.ring_build(['10.0.0.%d' % (i + 1) for i in range(16)])
db = eng.upload(eng.corpus_host(20130725, 0, n, 0))
db.parse_struct(); db.place_struct(16)
ts = []
for it in range(25):
    eng.memset(db.parse_struct().hdr_used, 0, 8)
    eng.event(0)
    if os.environ.get('SHL_PARSE_ONLY'):
        assert eng.lib.shl_parse_dev(eng.ctx, db.d_arena, db.d_off, db.d_len, db.n, C.byref(db.parse_struct())) == 0
    else:
        db.route(16)
    eng.event(1)
    eng.sync()
    if it >= 5:
        ts.append(eng.elapsed_ms(0, 1))
hist = db.fetch(db.d_hist, np.uint64, 16)
assert os.environ.get('SHL_PARSE_ONLY') or int(hist.sum()) == 25 * n
print('%-40s n=%d  kernel %.4f ms (min %.4f)  %.3f G req/s' % (os.environ.get('SHL_LIB', 'default'), n, float(np.mean(ts)), min(ts), n / np.mean(ts) / 1e6))
