#!/usr/bin/env python
"""Full-size runs of BASELINE.json configs 3, 4 and 5 on one GPU (bench.py covers config 2).
Prints one JSON object per config; meant to be run under gpurun with the output saved in profiles/.

    python tools/run_configs.py [c3] [c4] [c5]
"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from shellac_b200 import _native as N          # noqa: E402
from shellac_b200.batch import Arena, FrameOut, ParseOut  # noqa: E402
from shellac_b200.engine import Engine         # noqa: E402


def c3(eng):
    """8 192 upstream responses, 50 % Content-Length / 50 % chunked, bodies log-uniform 4 KiB-1 MiB,
    chunk sizes U{512..16384}, 10 % of chunks with one of the reference test's extension forms."""
    rng = np.random.default_rng(20130726)
    n = 8192
    pool = rng.integers(0, 256, 4 << 20, dtype=np.uint8).tobytes()
    exts = [b''] * 27 + [b';ext=foo', b';ext="foo"', b';ext=foo7']
    firsts = [b'HTTP/1.1 200 OK\r\n', b'HTTP/1.1 302 Not Modified\r\n', b'HTTP/1.1 500 Internal Server Error\r\n']
    parts, off, lens, pos, n_chunks, body_total = [], [], [], 0, 0, 0
    for i in range(n):
        size = int(np.exp(rng.uniform(np.log(4096), np.log(1 << 20))))
        s0 = int(rng.integers(0, len(pool) - size))
        body = pool[s0:s0 + size]
        head = firsts[i % 3] + b''.join(b'X-H%d: %s\r\n' % (k, b'v' * int(rng.integers(1, 40))) for k in range(int(rng.integers(3, 11))))
        if i % 2 == 0:
            msg = [head, b'Content-Length: %d\r\n\r\n' % size, body]
        else:
            msg = [head, b'Transfer-Encoding: chunked\r\n\r\n']
            p = 0
            while p < size:
                c = min(int(rng.integers(512, 16385)), size - p)
                msg += [(b'%x' if rng.random() < 0.5 else b'%X') % c, exts[int(rng.integers(0, 30))], b'\r\n', body[p:p + c], b'\r\n']
                p += c
                n_chunks += 1
            msg.append(b'0\r\n\r\n')
        ln = sum(len(x) for x in msg)
        parts += msg; off.append(pos); lens.append(ln); pos += ln; body_total += size
    data = eng.pinned(pos + 64, np.uint8)
    data[:pos] = np.frombuffer(b''.join(parts), np.uint8)
    a_off, a_len = eng.pinned(n, np.uint64), eng.pinned(n, np.uint32)
    a_off[:], a_len[:] = off, lens
    ar = Arena(data, a_off, a_len)
    db = eng.upload(ar)
    ps = db.parse_struct(hdr_cap=n * 16)
    fs = N.FrameOutStruct()
    ffields = {}
    for name, dt in N.FRAME_FIELDS:
        ffields[name] = db._alloc(n * np.dtype(dt).itemsize)
        setattr(fs, name, ffields[name])
    cap = n_chunks + 1024
    d_chunk, d_cused = db._alloc(cap * 16), db._alloc(8)
    fs.chunk, fs.chunk_cap, fs.chunk_used = d_chunk, cap, d_cused
    lib, ctx = eng.lib, eng.ctx

    def step():
        eng.memset(db.d_used, 0, 8); eng.memset(d_cused, 0, 8)
        eng.event(0)
        assert lib.shl_parse_dev(ctx, db.d_arena, db.d_off, db.d_len, n, C.byref(ps)) == 0
        eng.event(1)
        assert lib.shl_frame_dev(ctx, db.d_arena, db.d_off, db.d_len, n, db.p_fields['status'], db.p_fields['first_line_len'],
                                 db.p_fields['headers_len'], C.byref(fs)) == 0
        eng.event(2)
        eng.sync()
        return eng.elapsed_ms(0, 1), eng.elapsed_ms(1, 2)
    for _ in range(3):
        step()
    t = np.array([step() for _ in range(10)])
    fstat = db.fetch(ffields['fstatus'], np.uint32, n)
    nch = db.fetch(ffields['n_chunks'], np.uint32, n)
    bb = db.fetch(ffields['body_bytes'], np.uint64, n)
    st = db.field('status')
    assert int(nch.sum()) == n_chunks and ((fstat[1::2]) == 0).all() and ((st[0::2] & N.F_MESSAGE_COMPLETE) != 0).all()
    cl_bytes = int(db.field('content_len')[0::2].sum())
    assert int(bb.sum()) + cl_bytes == body_total
    # e2e through the host-buffer entry points (one H2D of the arena, reused by the framer)
    pout, fout = ParseOut(n, n * 16, alloc=eng.pinned), FrameOut(n, cap, alloc=eng.pinned)
    e = []
    for _ in range(4):
        t0 = time.perf_counter()
        eng.parse(ar, out=pout, sync=False)
        eng.frame(ar, pout, reuse_arena=True, out=fout, sync=True)
        e.append(time.perf_counter() - t0)
    # N1: body gather on the resident batch
    bstruct = N.BodyOutStruct()
    d_boff, d_blen, d_body, d_bused = db._alloc(n * 8), db._alloc(n * 8), db._alloc(pos + 16 * n + 64), db._alloc(8)
    bstruct.body_off, bstruct.body_len, bstruct.body, bstruct.body_cap, bstruct.body_used = d_boff, d_blen, d_body, pos + 16 * n, d_bused
    bt = []
    for it in range(8):
        eng.memset(d_bused, 0, 8)
        eng.event(3)
        assert lib.shl_body_dev(ctx, db.d_arena, db.d_off, db.d_len, n, C.byref(ps), C.byref(fs), C.byref(bstruct)) == 0
        eng.event(4)
        eng.sync()
        if it >= 3:
            bt.append(eng.elapsed_ms(3, 4))
    blen = db.fetch(d_blen, np.uint64, n)
    assert int(blen.sum()) == body_total
    bm = float(np.mean(bt))
    pm, fm = float(t[:, 0].mean()), float(t[:, 1].mean())
    return {'config': 3, 'responses': n, 'arena_gb': pos / 1e9, 'chunked': n // 2, 'chunks': n_chunks,
            'parse_ms': pm, 'frame_ms': fm, 'responses_per_s_resident': n / ((pm + fm) / 1e3),
            'chunks_per_s': n_chunks / (fm / 1e3), 'body_gb_s_skipped_by_arithmetic': pos / ((pm + fm) / 1e3) / 1e9,
            'body_gather_ms': bm, 'body_gather_gb_s_algorithmic_2x': 2 * body_total / (bm / 1e3) / 1e9, 'body_bytes': body_total,
            'e2e_ms_best': min(e) * 1e3, 'e2e_responses_per_s': n / min(e), 'e2e_gb_s': pos / min(e) / 1e9,
            'note': 'framing is a pointer chase over chunk headers: latency-bound, bodies are never read; e2e is the H2D of the bodies'}


def c4(eng, total=100_000_000, per=1 << 20):
    hosts = ['10.1.%d.%d' % (i // 250, i % 250 + 1) for i in range(256)]
    out = {}
    for mode, label in ((N.RING_KETAMA_WEIGHTED, '256x160 weighted'), (N.RING_KETAMA, '256x100 reference mode')):
        eng.ring_build(hosts, mode=mode)
        hist = np.zeros(256, np.uint64)
        k_ms, done, nbytes = 0.0, 0, 0
        t0 = time.perf_counter()
        b = 0
        while done < total:
            m = min(per, total - done)
            db = eng.corpus_dev(20130727, done, m, 0)
            db.parse_struct(); db.place_struct(256, md5=False)
            eng.memset(db.d_used, 0, 8)
            eng.event(0)
            db.route(256)
            eng.event(1)
            eng.sync()
            k_ms += eng.elapsed_ms(0, 1)
            if b % 16 == 0:
                assert ((db.field('status') & 0xff) == 0).all()
            hist += db.fetch(db.d_hist, np.uint64, 256)
            nbytes += db.nbytes
            done += m; b += 1
            db.free()
        assert int(hist.sum()) == total
        out[label] = {'requests': total, 'batches': b, 'kernel_ms_total': k_ms, 'requests_per_s_kernel': total / (k_ms / 1e3),
                      'header_gb_s': nbytes / (k_ms / 1e3) / 1e9, 'wall_s_incl_generation': time.perf_counter() - t0,
                      'hist_min': int(hist.min()), 'hist_max': int(hist.max()), 'hist_sum': int(hist.sum())}
        if mode == N.RING_KETAMA:
            break
    return {'config': 4, 'gpus': 1, 'rings': out}


def c5(eng, total=100_000_000, per=1 << 22):
    base_hosts = ['10.%d.%d.%d' % (2 + i // 62500, (i // 250) % 250, i % 250 + 1) for i in range(1024)]
    variants = [('base', base_hosts)] + [('drop%d+new' % k, base_hosts[:k] + base_hosts[k + 1:] + ['10.9.9.%d' % (k % 250 + 1)]) for k in (0, 511, 1023)]
    nb = (total + per - 1) // per
    base_nodes, res = [], {}
    for vi, (name, hosts) in enumerate(variants):
        eng.ring_build(hosts)
        names = np.array(hosts, dtype=object)
        k_ms, moved, done = 0.0, 0, 0
        hist = np.zeros(len(hosts), np.uint64)
        for b in range(nb):
            m = min(per, total - done)
            db = eng.corpus_dev(20130728, done, m, 1)
            db.place_struct(len(hosts), md5=False)
            eng.event(0)
            db.place(len(hosts))
            eng.event(1)
            eng.sync()
            k_ms += eng.elapsed_ms(0, 1)
            node = db.fetch(db.d_node, np.uint32, m)
            hist += db.fetch(db.d_hist, np.uint64, len(hosts))
            if vi == 0:
                base_nodes.append(node.astype(np.uint16))
            else:
                old = np.array(variants[0][1], dtype=object)[base_nodes[b]]
                moved += int((names[node] != old).sum())
            done += m
            db.free()
        res[name] = {'keys': total, 'kernel_ms_total': k_ms, 'keys_per_s_kernel': total / (k_ms / 1e3),
                     'moved_fraction': moved / total if vi else 0.0, 'expected_about': 2.0 / 1024 if vi else 0.0,
                     'hist_sum': int(hist.sum())}
    return {'config': 5, 'ring_points': 102400, 'variants': res}


if __name__ == '__main__':
    which = sys.argv[1:] or ['c3', 'c4', 'c5']
    eng = Engine(device=0, max_msgs=1 << 16, max_bytes=(1 << 31) - (1 << 20), max_records=1 << 22)
    for w in which:
        t0 = time.perf_counter()
        r = {'c3': c3, 'c4': c4, 'c5': c5}[w](eng)
        r['wall_s'] = time.perf_counter() - t0
        print(json.dumps(r), flush=True)
    eng.close()
