#!/usr/bin/env python
"""bench.py — BASELINE.json's metric, one JSON line on stdout.

    python bench.py [--config 2|3|4|5] [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

--config 2 (default, the headline): a step is one pass of the hot path (tokenise + key + MD5 +
    ketama placement, ONE fused kernel) over 2^20 synthetic GET requests (~512 B, 8-20 header lines,
    16x100-point ring) per GPU; ranks take disjoint slices of the same counter-based corpus (weak
    scaling) and sum the per-node histogram + per-outcome counts after EVERY step (the library's
    own NCCL all-reduce on a side stream, overlapped with the next step's kernel).
--config 3: 8 192 upstream responses per GPU (50 % Content-Length, 50 % chunked, bodies 4 KiB-1 MiB):
    a step = tokenise + chunk framing + body gather; value = responses/s.
--config 4: ONE job of 100 M requests, rank g owns [g*T/G, (g+1)*T/G) (strong scaling), resident in
    HBM, 256-node x 160-point weighted ring, counters all-reduced per 2^20-request batch.
--config 5: remap sweep: 100 M keys re-placed on a 1 024-node ring and on three "one node failed,
    one added" variants of it; value = keys/s; moved fraction and node identity checked.

  value : whole-job throughput with the inputs already resident in HBM (kernels + all-reduce).
  e2e   : the same metric through the C ABI's host-buffer entry point (pinned HOST arrays in, H2D,
          kernels, D2H of every result) inside the timed region.
  roofline : algorithmic bytes of the dominant kernel / its CUDA-event duration vs the measured HBM
          copy bandwidth in MEASURED_PEAKS.json.
  cpu_baseline : the reference's CPU path on ONE host core over a bounded sample (N=1 only).
  --impl reference : the reference's own CPU implementation on all host cores: the UNMODIFIED
          HttpParser from oracle/_ref (bytecode built by oracle/build_ref.py where /root/reference is
          mounted; kind "reference") or, without it, the CPython port (kind "port"); placement is
          the hashlib + bisect restatement either way (pylibmc / libmemcached are absent).  That
          arm never loads libshellac_b200.so: its corpus comes from oracle/_build/libcorpus.so.
No PyTorch anywhere: ranks find each other through the launcher's environment and a file.
"""
import argparse
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = 'parsed+routed requests/s (HTTP/1.1 tokenise + cache key + ketama/MD5 placement)'
BATCH = 1 << 20
RING_NOTE = 'libmemcached 0.44 semantics, restated (parity unpinned: pylibmc/libmemcached absent)'


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------------- configs
def hosts_for(cfg):
    if cfg == 2:
        return ['10.0.0.%d' % (i + 1) for i in range(16)], 0
    if cfg == 4:
        return ['10.1.%d.%d' % (i // 250, i % 250 + 1) for i in range(256)], 1       # ketama_weighted: 160 points
    if cfg == 5:
        return ['10.%d.%d.%d' % (2 + i // 62500, (i // 250) % 250, i % 250 + 1) for i in range(1024)], 0
    return [], 0


def c5_variants():
    base, _ = hosts_for(5)
    out = [('base', base)]
    for k in (0, 511, 1023):
        out.append(('drop%d+new' % k, base[:k] + base[k + 1:] + ['10.9.9.%d' % (k % 250 + 1)]))
    return out


SEEDS = {2: 20130725, 3: 20130726, 4: 20130727, 5: 20130728}


def config_dict(args, world):
    """The same object in both arms (the driver compares them)."""
    c = args.config
    if c == 2:
        return {'workload': 'config 2: 2^20 synthetic HTTP/1.1 GET requests per GPU per step, ~512 B, 8-20 header lines, '
                            '16-node x 100-point ketama ring', 'batch_per_gpu': args.batch, 'ring': '16x100 ketama/MD5, ' + RING_NOTE,
                'l2': 'inputs (537 MB per GPU) exceed the 126 MB L2; no flush needed'}
    if c == 3:
        return {'workload': 'config 3: 8 192 upstream responses per GPU per step, 50 % Content-Length / 50 % chunked, bodies '
                            'log-uniform 4 KiB-1 MiB, chunk sizes U{512..16384}, 10 % of chunks with an extension',
                'batch_per_gpu': args.batch, 'l2': 'inputs (1.5 GB per GPU) exceed the 126 MB L2; no flush needed'}
    if c == 4:
        return {'workload': 'config 4: ONE job of %d synthetic GET requests sharded over the GPUs (rank g owns [g*T/G, (g+1)*T/G)), '
                            'resident in HBM, 2^20-request batches, counters all-reduced per batch' % args.total,
                'total_requests': args.total, 'batch': args.batch, 'ring': '256 nodes x 160 points (ketama_weighted), ' + RING_NOTE,
                'l2': 'every batch (537 MB) exceeds the 126 MB L2 and is read once; no flush needed'}
    return {'workload': 'config 5: remap sweep, %d keys re-placed on a 1 024-node x 100-point ring and on 3 variants '
                        '(node 0 / 511 / 1023 removed, one node added)' % args.total, 'total_keys': args.total, 'batch': args.batch,
            'ring': '1024x100 ketama/MD5 (102 400 points), ' + RING_NOTE,
            'l2': 'every batch of keys (%.0f MB) exceeds the 126 MB L2 and is read once; no flush needed' % (args.batch * 60 / 1e6)}


def metric_unit(cfg):
    if cfg == 3:
        return 'framed upstream responses/s (tokenise + Content-Length / chunked framing + body gather)', 'responses/s'
    if cfg == 5:
        return 'placed keys/s (MD5 + ketama continuum lookup), remap sweep', 'keys/s'
    return METRIC, 'requests/s'


# ------------------------------------------------------------------------------------- plumbing
class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (via NVML)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        self.armed = threading.Event()      # NVML start-up costs milliseconds: do it BEFORE the barrier
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:      # noqa: BLE001
            self.nv = None
            log('clock sampling unavailable: %r' % (e,))

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: 'hw_slowdown',
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: 'hw_thermal_slowdown',
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: 'sw_thermal_slowdown',
                 nv.nvmlClocksThrottleReasonSwPowerCap: 'sw_power_cap',
                 nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: 'hw_power_brake'}
        self.armed.wait()
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:       # noqa: BLE001
                pass
            time.sleep(0.001)

    def result(self):
        self.stop_flag = True
        self.armed.set()
        if not self.samples:
            return {'sm_mhz': None, 'sm_max_mhz': self.max_mhz, 'reasons': sorted(self.reasons)}
        return {'sm_mhz': float(np.median(self.samples)), 'sm_max_mhz': self.max_mhz,
                'reasons': sorted(self.reasons), 'samples': len(self.samples)}


def measured_peak():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        try:
            return float(json.load(open(p))['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
        except Exception:           # noqa: BLE001
            pass
    return 6650.0, 'fallback (B200_PROFILING.md 6.65 TB/s)'


def ncu_traffic(key):
    """dram bytes per launch of a kernel from the committed ncu --set full capture (profiles/traffic.json)."""
    p = os.path.join(ROOT, 'profiles', 'traffic.json')
    if os.path.exists(p):
        try:
            return json.load(open(p)).get(key)
        except Exception:           # noqa: BLE001
            return None
    return None


# ------------------------------------------------------------------------------------- corpora on the host
_corpus = None


def corpus_lib():
    """oracle/_build/libcorpus.so: the counter-based generator as a host library (NOT the product)."""
    global _corpus
    if _corpus is None:
        sys.path.insert(0, os.path.join(ROOT, 'oracle'))
        import build_ref
        L = C.CDLL(build_ref.build_corpus_lib())
        L.orc_corpus_lengths.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, C.c_void_p]
        L.orc_corpus_fill.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, C.c_void_p, C.c_void_p]
        _corpus = L
    return _corpus


def host_corpus(seed, first, n, kind):
    """(blob bytes, off u64[n], len u32[n]) of messages [first, first + n)."""
    L = corpus_lib()
    ln = np.zeros(n, np.uint32)
    L.orc_corpus_lengths(seed, first, n, kind, ln.ctypes.data)
    off = np.zeros(n, np.uint64)
    np.cumsum(ln[:-1], dtype=np.uint64, out=off[1:])
    data = np.zeros(int(off[-1] + ln[-1]) + 64, np.uint8)
    L.orc_corpus_fill(seed, first, n, kind, off.ctypes.data, data.ctypes.data)
    return data, off, ln


def host_messages(seed, first, n, kind, text):
    data, off, ln = host_corpus(seed, first, n, kind)
    blob = data.tobytes()
    if text:
        blob = blob.decode('latin-1')
    return [blob[int(o):int(o) + int(l)] for o, l in zip(off, ln)], int(ln.sum())


def c3_messages(n, rank=0):
    """Config 3's responses: list of bytes, plus chunk and body totals."""
    rng = np.random.default_rng(SEEDS[3] + rank)
    pool = rng.integers(0, 256, 4 << 20, dtype=np.uint8).tobytes()
    exts = [b''] * 27 + [b';ext=foo', b';ext="foo"', b';ext=foo7']
    firsts = [b'HTTP/1.1 200 OK\r\n', b'HTTP/1.1 302 Not Modified\r\n', b'HTTP/1.1 500 Internal Server Error\r\n']
    msgs, n_chunks, body_total = [], 0, 0
    for i in range(n):
        size = int(np.exp(rng.uniform(np.log(4096), np.log(1 << 20))))
        s0 = int(rng.integers(0, len(pool) - size))
        body = pool[s0:s0 + size]
        head = firsts[i % 3] + b''.join(b'X-H%d: %s\r\n' % (k, b'v' * int(rng.integers(1, 40))) for k in range(int(rng.integers(3, 11))))
        if i % 2 == 0:
            parts = [head, b'Content-Length: %d\r\n\r\n' % size, body]
        else:
            parts = [head, b'Transfer-Encoding: chunked\r\n\r\n']
            p = 0
            while p < size:
                c = min(int(rng.integers(512, 16385)), size - p)
                parts += [(b'%x' if rng.random() < 0.5 else b'%X') % c, exts[int(rng.integers(0, 30))], b'\r\n', body[p:p + c], b'\r\n']
                p += c
                n_chunks += 1
            parts.append(b'0\r\n\r\n')
        msgs.append(b''.join(parts))
        body_total += size
    return msgs, n_chunks, body_total


# ------------------------------------------------------------------------------------- CPU arms
def _ref():
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import py_port
    import ref_driver
    return py_port, ref_driver


def _drive_responses(msgs, cls, port):
    """Config 3 on the CPU: every response driven to completion like test/HttpParserTests.py:36-39."""
    done = 0
    for data in msgs:
        p = cls()
        if port:
            while data and not p.message_complete:
                c = p.parse(data, len(data))
                data = data[c:]
            done += bool(p.message_complete)
        else:
            while data and not p.message_complete():
                c = p.parse(data, len(data))
                data = data[c:]
            done += bool(p.message_complete())
    return done


_W = {}


def _cpu_unit(cfg, items, ring, cls):
    """One slice of CPU work; returns the units completed."""
    pp, rd = _ref()
    port = cls is None
    if cfg == 3:
        return _drive_responses(items, pp.PortParser if port else cls, port)
    v, nd, nh = ring
    if cfg == 5:
        return len(rd.place_keys(items, v, nd, nh))
    if port:
        return pp.route_messages(items, v, nd, nh)[0]
    return rd.route_reference(items, v, nd, nh, cls)[0]


def _worker_run(span):
    a, b = span
    return _cpu_unit(_W['cfg'], _W['items'][a:b], _W['ring'], _W['cls'])


def cpu_items(cfg, n, text, rank=0, first=0):
    """n units of config cfg's corpus for a CPU arm: (items, bytes)."""
    if cfg == 3:
        msgs, _, _ = c3_messages(n, rank)
        nb = sum(len(m) for m in msgs)
        return ([m.decode('latin-1') for m in msgs] if text else msgs), nb
    if cfg == 5:
        return host_messages(SEEDS[5], first, n, 1, False)
    return host_messages(SEEDS[cfg], first, n, 0, text)


def cpu_ring(cfg):
    if cfg == 3:
        return None
    pp, _ = _ref()
    hosts, mode = hosts_for(cfg)
    v, nd = pp.build_ring(hosts, mode=mode)
    return v, nd, len(hosts)


def cpu_kind(cfg):
    """(parser class or None, kind): the unmodified reference when oracle/_ref travelled here."""
    if cfg == 5:
        return None, 'port'            # placement only: the restatement IS the CPU path (libmemcached absent)
    _, rd = _ref()
    return rd.load_reference()


CPU_WHAT = {'reference': 'UNMODIFIED reference HttpParser.py (bytecode in oracle/_ref) + url() + hashlib/bisect ketama restatement',
            'port': 'oracle/py_port.py (CPython restatement of HttpParser.py + url() + hashlib/bisect ketama)'}


def cpu_baseline_single(cfg, unit):
    """The reference's CPU path on ONE core over a bounded sample of the same corpus."""
    n = {2: 400000, 3: 4096, 4: 400000, 5: 8000000}[cfg]
    cls, kind = cpu_kind(cfg)
    items, nbytes = cpu_items(cfg, n, text=kind == 'reference')
    ring = cpu_ring(cfg)
    t0 = time.perf_counter()
    done = _cpu_unit(cfg, items, ring, cls)
    dt = time.perf_counter() - t0
    assert done == n, (done, n)
    return {'value': n / dt, 'unit': unit, 'cores': 1, 'kind': kind,
            'sample': '%d units of the config-%d corpus (%.1f MB), %s, %.1f s' % (n, cfg, nbytes / 1e6, CPU_WHAT[kind], dt),
            'gb_s': nbytes / dt / 1e9}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on all host cores."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return 0
    import multiprocessing as mp
    cfg = args.config
    metric, unit = metric_unit(cfg)
    cores = len(os.sched_getaffinity(0))
    n = {2: args.batch, 3: min(args.batch, 1024), 4: args.batch, 5: args.batch}[cfg]
    cls, kind = cpu_kind(cfg)
    t0 = time.perf_counter()
    items, nbytes = cpu_items(cfg, n, text=kind == 'reference')
    log('reference arm: %d units (%.1f MB) generated in %.1f s; kind %s, %d cores' % (n, nbytes / 1e6, time.perf_counter() - t0, kind, cores))
    _W.update({'cfg': cfg, 'items': items, 'ring': cpu_ring(cfg), 'cls': cls})
    per = (n + cores * 4 - 1) // (cores * 4)
    spans = [(a, min(n, a + per)) for a in range(0, n, per)]
    with mp.get_context('fork').Pool(cores) as pool:
        for _ in range(args.warmup):
            pool.map(_worker_run, spans)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            done = sum(pool.map(_worker_run, spans))
            assert done == n, (done, n)
        dt = time.perf_counter() - t0
    v = n * args.steps / dt
    sample = 'each step = %d units of the config-%d corpus over %d cores; %s' % (n, cfg, cores, CPU_WHAT[kind])
    line = {'impl': 'reference', 'metric': metric, 'value': v, 'unit': unit, 'n_gpus': args.gpus,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': dt / args.steps * 1e3,
            'higher_is_better': True, 'scaling': 'strong' if cfg in (4, 5) else 'weak', 'vs_baseline': None, 'dtype': 'u8',
            'data': 'synthetic', 'config': config_dict(args, 1), 'gb_s': nbytes * args.steps / dt / 1e9,
            'cpu_baseline': {'value': v, 'unit': unit, 'cores': cores, 'kind': kind, 'sample': sample},
            'e2e': {'value': v, 'unit': unit, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------- GPU arm: plumbing
class Job(object):
    """One rank of the job: its engine, the communicator, barrier / max-over-ranks and the timed region."""

    def __init__(self, args, max_msgs, max_bytes, max_records):
        from shellac_b200 import build as B
        from shellac_b200 import dist as D
        from shellac_b200.engine import Engine
        self.rank, self.local, self.world = D.env_rank()
        if self.rank == 0:
            B.build()
        self.numa = None
        if not os.environ.get('SHL_NO_BIND'):
            from shellac_b200.hostmem import bind_near_gpu
            self.numa = bind_near_gpu(self.local)          # before any pinned arena is allocated
        self.eng = Engine(device=self.local, max_msgs=max_msgs, max_bytes=max_bytes, max_records=max_records)
        self.lib, self.ctx = self.eng.lib, self.eng.ctx
        D.join_job(self.eng, self.rank, self.world)          # the library's own NCCL communicator (no torch)
        log('[rank %d/%d] numa binding: %r, %d cpus allowed' % (self.rank, self.world, self.numa, len(os.sched_getaffinity(0))))

    def barrier(self):
        self.eng.sync()
        self.eng.barrier()

    def max_over_ranks(self, x):
        return float(self.eng.allreduce_u64([int(x * 1e6)], op='max')[0]) / 1e6

    def sum_over_ranks(self, x):
        return int(self.eng.allreduce_u64([int(x)], op='sum')[0])

    def timed(self, fn, steps, warmup, flush=None, regions=1):
        """W untimed steps, then exactly K steps between two barriers, timed with CUDA events on the
        ctx stream; returns (ms = max over ranks, clocks, kernel launches).  regions > 1 repeats the
        timed region (K steps each, its own barriers) and reports the MEDIAN region, so that a
        millisecond-long region is not at the mercy of one hiccup and the clock sampler sees more
        than a handful of samples; self.region_ms keeps every region's time."""
        eng = self.eng
        for i in range(warmup):
            fn(i)
        if flush:
            flush()
        sampler = ClockSampler(self.local)
        sampler.start()
        times, launches = [], 0
        for _ in range(regions):
            self.barrier()
            sampler.armed.set()
            l0 = eng.launches()
            eng.event(0)
            for i in range(steps):
                fn(i)
            if flush:
                flush()
            eng.event(1)
            self.barrier()
            times.append(self.max_over_ranks(eng.elapsed_ms(0, 1)))
            launches = eng.launches() - l0
        clocks = sampler.result()
        self.region_ms = times
        return float(np.median(times)), clocks, launches

    def pcie_probe(self, d_buf, h_buf, nbytes):
        """Pinned copies of the step's bytes: each direction alone, then both at once on two streams,
        every rank at the same time (the ranks of one box share the host's memory system)."""
        from shellac_b200.engine import Engine
        eng = self.eng
        out = {}
        for name, fn in (('h2d', lambda e: e.h2d(d_buf, h_buf, nbytes)), ('d2h', lambda e: e.d2h(h_buf, d_buf, nbytes))):
            best = 1e9
            for _ in range(3):
                self.barrier()
                eng.event(4); fn(eng); eng.event(5); eng.sync()
                best = min(best, eng.elapsed_ms(4, 5))
            out[name + '_gb_s'] = nbytes / (best / 1e3) / 1e9
        e2 = Engine(device=self.local, max_msgs=64, max_bytes=1 << 16)
        h2 = eng.pinned(nbytes, np.uint8)
        d2 = eng.dev_alloc(nbytes + 64)
        best = 1e9
        for _ in range(3):
            self.barrier()
            eng.event(4); e2.event(4)
            eng.h2d(d_buf, h_buf, nbytes); e2.d2h(h2, d2, nbytes)
            eng.event(5); e2.event(5)
            eng.sync(); e2.sync()
            best = min(best, max(eng.elapsed_ms(4, 5), e2.elapsed_ms(4, 5)))
        out['concurrent'] = {'h2d_and_d2h_gb_s_each': nbytes / (best / 1e3) / 1e9,
                             'note': 'H2D and D2H of the same bytes in flight together, all %d ranks at once' % self.world}
        eng.dev_free(d2)
        e2.close()
        return out


# ------------------------------------------------------------------------------------- GPU arm: configs 2 and 4
def run_route(args):
    """Config 2 (weak scaling, the headline) and config 4 (one 100 M-request job, strong scaling):
    the fused tokenise + key + MD5 + ketama kernel over HBM-resident batches."""
    from shellac_b200 import _native as N
    from shellac_b200 import dist as D
    from shellac_b200.batch import ParseOut, PlaceOut
    cfg, B = args.config, args.batch
    job = Job(args, max_msgs=B, max_bytes=B * 640, max_records=B * 24)
    eng, lib, ctx, rank, world = job.eng, job.lib, job.ctx, job.rank, job.world
    hosts, mode = hosts_for(cfg)
    eng.ring_build(hosts, mode=mode)
    nh = len(hosts)
    t0 = time.perf_counter()
    if cfg == 2:
        first, n_share = rank * B, B
        ar = eng.corpus_host(SEEDS[2], first, B, 0)          # pinned host arena, disjoint slice per rank
        db = eng.upload(ar)
    else:
        first, hi = D.shard_range(args.total, rank, world)
        n_share = hi - first
        db = eng.corpus_dev(SEEDS[4], first, n_share, 0)     # the rank's whole share, generated in HBM
        ar = eng.corpus_host(SEEDS[4], first, min(B, n_share), 0)      # e2e sample: the share's first batch
    batches = (n_share + B - 1) // B
    log('[rank %d] corpus: %d requests resident (%.1f MB) in %d batch(es), %.1f s' % (rank, n_share, db.nbytes / 1e6, batches, time.perf_counter() - t0))

    hdr_cap = B * 24
    ps = N.ParseOutStruct()
    for name, dt in N.PARSE_FIELDS:
        setattr(ps, name, eng.dev_alloc(B * np.dtype(dt).itemsize))
    d_hdr, d_used = eng.dev_alloc(hdr_cap * 8), eng.dev_alloc(8)
    ps.hdr, ps.hdr_cap, ps.hdr_used = d_hdr, hdr_cap, d_used
    qs = N.PlaceOutStruct()
    want_md5 = cfg == 2
    qs.md5 = eng.dev_alloc(16 * B) if want_md5 else None
    qs.hash, qs.node = eng.dev_alloc(4 * B), eng.dev_alloc(4 * B)
    n_stat = nh + N.N_CODES
    st = {'since': 0, 'units': 0}

    def launch(b):
        m = min(B, n_share - b * B)
        qs.hist = eng.stats_dev()                            # the library's CURRENT counters buffer
        eng.memset(d_used, 0, 8)
        rc = lib.shl_route_dev(ctx, db.d_arena, db.d_off + 8 * b * B, db.d_len + 4 * b * B, m, C.byref(ps), C.byref(qs))
        assert rc == 0, lib.shl_last_error(ctx)
        return m

    def step(i):
        st['units'] += launch(i % batches)
        st['since'] += 1
        if args.stats_every and st['since'] >= args.stats_every:
            eng.stats_allreduce()                            # NCCL on the side stream; the next kernel overlaps it
            st['since'] = 0

    def flush():
        if st['since']:
            eng.stats_allreduce()
            st['since'] = 0
        eng.stats_join()                                     # the timed region ends when the last all-reduce has

    steps = args.steps if args.steps > 0 else batches
    st['units'] = 0
    for i in range(args.warmup):
        step(i)
    flush()
    st['units'] = 0
    regions = 5 if cfg == 2 else 1
    ms, clocks, launches = job.timed(step, steps, 0, flush, regions=regions)
    region_ms = list(job.region_ms)
    units = job.sum_over_ranks(st['units']) // regions

    # kernel-only duration for the roofline: events around the kernel alone
    eng.sync()
    kms = []
    for _ in range(max(5, min(steps, 20))):
        eng.memset(d_used, 0, 8)
        qs.hist = eng.stats_dev()
        eng.event(2)
        lib.shl_route_dev(ctx, db.d_arena, db.d_off, db.d_len, min(B, n_share), C.byref(ps), C.byref(qs))
        eng.event(3)
        eng.sync()
        kms.append(eng.elapsed_ms(2, 3))
    k_ms = float(np.median(kms))
    total_hdrs = np.zeros(1, np.uint64)
    eng.d2h(total_hdrs, d_used); eng.sync()
    n_hdr = int(total_hdrs[0])
    m0 = min(B, n_share)
    b0_bytes = int(db.off[m0 - 1] + db.len[m0 - 1])
    log('[rank %d] kernel alone %.4f ms (median of %d), step %.4f ms (max over ranks)' % (rank, k_ms, len(kms), ms / steps))

    # ---- check (outside the timed region): every batch of the share once, ONE all-reduce at the end
    eng.stats_allreduce()                                    # closes whatever the timing loops left behind
    for b in range(batches):
        launch(b)
    eng.stats_allreduce()
    hs = eng.stats_fetch()
    expect = job.sum_over_ranks(n_share)
    assert int(hs[:nh].sum()) == expect, (int(hs[:nh].sum()), expect)
    assert int(hs[nh + N.S_OK]) == expect
    check = {'histogram_sum': int(hs[:nh].sum()), 'expected': expect, 'ok_count': int(hs[nh + N.S_OK]),
             'histogram_md5': hashlib.md5(hs[:nh].astype('<u8').tobytes()).hexdigest(),
             'hist_min': int(hs[:nh].min()), 'hist_max': int(hs[:nh].max())}
    if cfg == 4:
        ref = os.path.join(ROOT, 'profiles', 'r02', 'c4_histogram.json')
        if os.path.exists(ref):
            r = json.load(open(ref))
            if r.get('total') == args.total:
                check['equals_n1_histogram'] = r.get('histogram_md5') == check['histogram_md5']
        elif world == 1 and rank == 0 and args.total == 100_000_000:
            check['n1_histogram_written'] = True
            os.makedirs(os.path.dirname(ref), exist_ok=True)
            json.dump({'total': args.total, 'histogram_md5': check['histogram_md5'], 'histogram': hs[:nh].tolist()}, open(ref, 'w'))
    if rank == 0 and not args.no_cpu:
        check['oracle'] = oracle_check_route(eng, db, ps, qs, hosts, mode, first, B, n_share, launch, SEEDS[cfg])

    # ---- e2e: host buffers through shl_route_batch
    pout = ParseOut(ar.n, hdr_cap, alloc=eng.pinned)
    qout = PlaceOut(ar.n, nh, alloc=eng.pinned, want_md5=want_md5)

    def step_e2e(i):
        qout.hist[:] = 0
        eng.route(ar, pout=pout, qout=qout, sync=True)

    e_steps = steps if cfg == 2 else min(steps, 10)
    e_ms, e_clocks, _ = job.timed(step_e2e, e_steps, max(args.warmup, 1))
    assert int(qout.hist.sum()) == ar.n and (pout.code()[:ar.n] == 0).all()
    nbytes = ar.nbytes
    # the same call with the compact 32-byte record as its output (no digests): fewer bytes back over PCIe
    cout = eng.route_compact(ar, hdr_cap=hdr_cap, want_md5=False)

    def step_compact(i):
        cout[4][:] = 0
        eng.route_compact(ar, out=cout, sync=True)

    c_ms, _, _ = job.timed(step_compact, e_steps, 1)
    assert int(cout[4].sum()) == ar.n and ((cout[0]['status'][:ar.n] & 0xff) == 0).all()
    pcie = job.pcie_probe(db.d_arena, ar.data, nbytes)
    e_hdr = int(pout.n_headers[:ar.n].sum())
    h2d = nbytes + ar.n * 12
    d2h = ar.n * 48 + e_hdr * 8 + ar.n * (24 if want_md5 else 8) + nh * 8 + 8

    metric, unit = metric_unit(cfg)
    value = units / (ms / 1e3)
    e2e = ar.n * world * e_steps / (e_ms / 1e3)
    e2e_c = ar.n * world * e_steps / (c_ms / 1e3)
    fixed = 12 + 48 + (24 if want_md5 else 8)
    alg_bytes = b0_bytes + m0 * fixed + 8 * n_hdr
    peak, peak_src = measured_peak()
    achieved = alg_bytes / (k_ms / 1e3) / 1e9
    conf = config_dict(args, world)
    line = {
        'metric': metric, 'value': value, 'unit': unit, 'n_gpus': world, 'steps': steps,
        'warmup': args.warmup, 'ms_per_step': ms / steps, 'higher_is_better': True,
        'scaling': 'weak' if cfg == 2 else 'strong', 'vs_baseline': None, 'dtype': 'u8', 'data': 'synthetic', 'config': conf,
        'value_note': 'device-resident: the arena is already in HBM when the timed region starts (kernel + counter all-reduce); '
                      'the end-to-end number through the C ABI with host buffers is `e2e`; vs_baseline stays null because the '
                      'reference publishes no number for this metric (BASELINE.md)',
        'parallelism': 'dp%d (batch-sharded, no data-path collective; %d uint64 counters summed by the library\'s NCCL '
                       'all-reduce every %s step(s), overlapped with the next kernel)' % (world, n_stat, args.stats_every or 'K'),
        'mean_request_bytes': b0_bytes / m0, 'mean_header_lines': n_hdr / m0,
        'header_gb_s': value * b0_bytes / m0 / 1e9,
        'clocks': clocks, 'gpu_launches': launches, 'check': check,
        'timed_regions': {'count': regions, 'ms_each': region_ms, 'reported': 'median region; every region is exactly K steps between two barriers'},
        'e2e': {'value': e2e_c, 'unit': unit, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': ar.n * 32 + e_hdr * 8 + nh * 8 + 8,
                'header_gb_s': nbytes * world * e_steps / (c_ms / 1e3) / 1e9, 'steps': e_steps,
                'pcie_gb_s_h2d': h2d * e_steps / (c_ms / 1e3) / 1e9, 'ms_per_step': c_ms / e_steps,
                'wide': {'value': e2e, 'ms_per_step': e_ms / e_steps, 'd2h_bytes_per_step': d2h,
                         'note': 'shl_route_batch: the struct-of-arrays layout (48 + 24 bytes of arrays per request, digests included)'},
                'pcie_probe': pcie, 'frac_of_pcie_h2d_roofline': (h2d * e_steps / (c_ms / 1e3) / 1e9) / pcie['h2d_gb_s'],
                'note': 'shl_route_compact_batch on pinned host arrays, %d requests per rank per step (one 32-byte record + 8 bytes per '
                        'header line come back per request): chunked over 4 streams so H2D, kernel and D2H overlap; bound by the H2D '
                        'link.  The GPU arm returns spans; the CPU arm materialises a string and a header dict per request.' % ar.n,
                'host_numa': 'process bound to NUMA node %d (%d cpus) of its GPU' % job.numa if job.numa else 'unbound',
                'clocks': e_clocks},
        'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                     'traffic': ncu_traffic('k_route_dram_bytes_per_launch'), 'kernel': 'shl::k_route<true, true>', 'kernel_ms': k_ms,
                     'algorithmic_bytes_per_launch': alg_bytes, 'bytes_per_request': alg_bytes / m0,
                     'frac_with_survey_692B': (692.1 * m0 / (k_ms / 1e3) / 1e9) / peak if cfg == 2 else None,
                     'peak_source': peak_src, 'frac_of_nominal_8TBs': achieved / 8000.0},
    }
    if rank == 0 and world == 1 and not args.no_cpu:
        line['cpu_baseline'] = cpu_baseline_single(cfg, unit)
        line['vs_cpu_baseline_1core'] = {'resident': value / line['cpu_baseline']['value'], 'e2e': e2e_c / line['cpu_baseline']['value']}
    if cfg == 2 and rank == 0 and not args.no_cpu:
        try:
            line['e2e_facade'] = facade_rates(eng, ar)
        except Exception as e:      # noqa: BLE001
            line['e2e_facade'] = {'error': repr(e)}
    job.barrier()
    eng.close()
    return line if rank == 0 else None


def oracle_check_route(eng, db, ps, qs, hosts, mode, first, B, n_share, launch, seed):
    """The C oracle as the checker: 2^17 requests from the share's first and last batch, regenerated
    on the host, parsed + placed by oracle/shl_oracle.c, compared with what the kernel left."""
    from shellac_b200.batch import Arena
    from tests import oracle_lib as O
    vals, nodes = O.orc_ring(hosts, [11211] * len(hosts), None, mode)
    batches = (n_share + B - 1) // B
    out = {'compared': 0, 'mismatches': 0}
    for b in sorted({0, batches - 1}):
        m = min(B, n_share - b * B, 1 << 17)
        launch(b)
        eng.sync()
        node = np.zeros(m, np.uint32); status = np.zeros(m, np.uint32); nhd = np.zeros(m, np.uint32)
        eng.d2h(node, qs.node); eng.d2h(status, ps.status); eng.d2h(nhd, ps.n_headers); eng.sync()
        data, off, ln = host_corpus(seed, first + b * B, m, 0)
        ar = Arena(data, off, ln)
        o = O.orc_parse(ar)
        oq = O.orc_place(ar, vals, nodes, len(hosts), (ar.off + o.sp1[:m] + 1).astype(np.uint64), (o.sp2[:m] - o.sp1[:m] - 1).astype(np.uint32))
        bad = int((node != oq.node[:m]).sum() + (status != o.status[:m]).sum() + (nhd != o.n_headers[:m]).sum())
        out['compared'] += m
        out['mismatches'] += bad
    assert out['mismatches'] == 0, out
    return out                     # (no collective in here: only rank 0 runs the checker)


# ------------------------------------------------------------------------------------- GPU arm: config 3
def run_frame(args):
    """Config 3: upstream responses through tokenise + Content-Length / chunked framing + body gather."""
    from shellac_b200 import _native as N
    from shellac_b200.batch import Arena, BodyOut, FrameOut, ParseOut
    n = args.batch
    job = Job(args, max_msgs=max(n, 64), max_bytes=(1 << 31) - (1 << 20), max_records=1 << 22)
    eng, lib, ctx, rank, world = job.eng, job.lib, job.ctx, job.rank, job.world
    t0 = time.perf_counter()
    msgs, n_chunks, body_total = c3_messages(n, rank)
    ar = eng.pack(msgs)
    del msgs
    pos = ar.nbytes
    db = eng.upload(ar)
    log('[rank %d] corpus: %d responses, %.1f MB, %d chunks, built in %.1f s' % (rank, n, pos / 1e6, n_chunks, time.perf_counter() - t0))
    ps = db.parse_struct(hdr_cap=n * 16)
    fs = N.FrameOutStruct()
    ff = {}
    for name, dt in N.FRAME_FIELDS:
        ff[name] = db._alloc(n * np.dtype(dt).itemsize)
        setattr(fs, name, ff[name])
    cap = n_chunks + 1024
    d_chunk, d_cused = db._alloc(cap * 16), db._alloc(8)
    fs.chunk, fs.chunk_cap, fs.chunk_used = d_chunk, cap, d_cused
    bs = N.BodyOutStruct()
    d_boff, d_blen, d_body, d_bused = db._alloc(n * 8), db._alloc(n * 8), db._alloc(pos + 16 * n + 64), db._alloc(8)
    bs.body_off, bs.body_len, bs.body, bs.body_cap, bs.body_used = d_boff, d_blen, d_body, pos + 16 * n, d_bused

    def step(i, ev=False):
        eng.memset(db.d_used, 0, 8); eng.memset(d_cused, 0, 8); eng.memset(d_bused, 0, 8)
        if ev:
            eng.event(2)
        assert lib.shl_parse_dev(ctx, db.d_arena, db.d_off, db.d_len, n, C.byref(ps)) == 0
        if ev:
            eng.event(3)
        assert lib.shl_frame_dev(ctx, db.d_arena, db.d_off, db.d_len, n, db.p_fields['status'], db.p_fields['first_line_len'],
                                 db.p_fields['headers_len'], C.byref(fs)) == 0
        if ev:
            eng.event(6)
        assert lib.shl_body_dev(ctx, db.d_arena, db.d_off, db.d_len, n, C.byref(ps), C.byref(fs), C.byref(bs)) == 0
        if ev:
            eng.event(7)
        eng.stats_allreduce()                                # per-outcome counts, summed over the ranks per batch

    ms, clocks, launches = job.timed(step, args.steps, args.warmup, eng.stats_join)
    parts = []
    for i in range(5):
        step(i, ev=True)
        eng.sync()
        parts.append((eng.elapsed_ms(2, 3), eng.elapsed_ms(3, 6), eng.elapsed_ms(6, 7)))
    pm, fm, bm = (float(x) for x in np.median(np.array(parts), axis=0))
    # checks
    fstat, nch = db.fetch(ff['fstatus'], np.uint32, n), db.fetch(ff['n_chunks'], np.uint32, n)
    blen = db.fetch(d_blen, np.uint64, n)
    stt = db.field('status')
    assert int(nch.sum()) == n_chunks and (fstat[1::2] == 0).all() and ((stt[0::2] & N.F_MESSAGE_COMPLETE) != 0).all()
    assert int(blen.sum()) == body_total
    check = {'chunks': int(nch.sum()), 'body_bytes': int(blen.sum()), 'expected_body_bytes': body_total}
    if rank == 0 and not args.no_cpu:
        from tests import oracle_lib as O
        k = min(n, 512)
        sub = Arena(ar.data, ar.off[:k], ar.len[:k])
        o = O.orc_parse(sub); fo = O.orc_frame(sub, o); bo = O.orc_body(sub, o, fo)
        goff = db.fetch(d_boff, np.uint64, n)                # bodies are handed out in completion order: fetch each
        bad = 0
        for i in range(k):
            gb = np.zeros(max(int(blen[i]), 1), np.uint8)
            if blen[i]:
                eng.d2h(gb, d_body + int(goff[i]), int(blen[i])); eng.sync()
            bad += gb[:int(blen[i])].tobytes() != bo.get(i)
        bad += int((stt[:k] != o.status[:k]).sum() + (nch[:k] != fo.n_chunks[:k]).sum())
        assert bad == 0, bad
        check['oracle'] = {'compared_responses': k, 'mismatches': bad, 'what': 'status, chunk counts, gathered body bytes'}

    # e2e: host buffers through parse_batch + frame_batch + body_batch (one H2D of the arena)
    pout, fout = ParseOut(n, n * 16, alloc=eng.pinned), FrameOut(n, cap, alloc=eng.pinned)
    bout = BodyOut(n, pos + 16 * n + 64, alloc=eng.pinned)

    def step_e2e(i):
        eng.parse(ar, out=pout, sync=False)
        eng.frame(ar, pout, reuse_arena=True, out=fout, sync=False)
        eng.body(ar, out=bout, sync=True)

    e_steps = min(args.steps, 8)
    e_ms, e_clocks, _ = job.timed(step_e2e, e_steps, 1)
    assert int(bout.body_len[:n].sum()) == body_total
    metric, unit = metric_unit(3)
    value = n * world * args.steps / (ms / 1e3)
    e2e = n * world * e_steps / (e_ms / 1e3)
    peak, peak_src = measured_peak()
    achieved = 2 * body_total / (bm / 1e3) / 1e9
    h2d = pos + n * 12
    d2h = n * 48 + n * 16 * 8 + n * 32 + n_chunks * 16 + n * 16 + body_total
    line = {'metric': metric, 'value': value, 'unit': unit, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
            'ms_per_step': ms / args.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'u8',
            'data': 'synthetic', 'config': config_dict(args, world), 'clocks': clocks, 'gpu_launches': launches, 'check': check,
            'arena_gb': pos / 1e9, 'chunks_per_s': n_chunks * world * args.steps / (ms / 1e3),
            'kernel_ms': {'k_route<false> (tokenise)': pm, 'k_frame (chunk walk)': fm, 'k_body_plan + k_body_copy (gather)': bm},
            'e2e': {'value': e2e, 'unit': unit, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h, 'steps': e_steps,
                    'ms_per_step': e_ms / e_steps, 'gb_s_h2d': h2d * e_steps / (e_ms / 1e3) / 1e9, 'clocks': e_clocks,
                    'note': 'shl_parse_batch + shl_frame_batch + shl_body_batch on pinned host arrays: the bodies cross PCIe both ways'},
            'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                         'traffic': ncu_traffic('k_body_copy_dram_bytes_per_launch'), 'kernel': 'shl::k_body_copy', 'kernel_ms': bm,
                         'algorithmic_bytes_per_launch': 2 * body_total, 'peak_source': peak_src,
                         'note': 'the framing kernels skip bodies by arithmetic (latency-bound pointer chase): responses/s and '
                                 'chunks/s are their figures; the gather moves 2 x body bytes'}}
    if rank == 0 and world == 1 and not args.no_cpu:
        line['cpu_baseline'] = cpu_baseline_single(3, unit)
    job.barrier()
    eng.close()
    return line if rank == 0 else None


# ------------------------------------------------------------------------------------- GPU arm: config 5
def run_remap(args):
    """Config 5: every key of the job placed on the base ring and on each "one node failed, one
    added" variant; keys resident in HBM, sharded over the ranks, counters all-reduced per batch."""
    from shellac_b200 import _native as N
    from shellac_b200 import dist as D
    from shellac_b200.batch import Arena, PlaceOut
    B = args.batch
    job = Job(args, max_msgs=B, max_bytes=B * 160, max_records=1024)
    eng, lib, ctx, rank, world = job.eng, job.lib, job.ctx, job.rank, job.world
    first, hi = D.shard_range(args.total, rank, world)
    n_share = hi - first
    t0 = time.perf_counter()
    db = eng.corpus_dev(SEEDS[5], first, n_share, 1)
    batches = (n_share + B - 1) // B
    log('[rank %d] %d keys resident (%.1f MB), %d batch(es), %.1f s' % (rank, n_share, db.nbytes / 1e6, batches, time.perf_counter() - t0))
    variants = c5_variants()
    base_hosts = variants[0][1]
    d_node = [eng.dev_alloc(4 * n_share), eng.dev_alloc(4 * n_share)]        # base ring / current variant
    d_hash = eng.dev_alloc(4 * B)
    qs = N.PlaceOutStruct()
    qs.md5 = None
    qs.hash = d_hash

    def place_all(slot, timed_units):
        for b in range(batches):
            m = min(B, n_share - b * B)
            qs.node = d_node[slot] + 4 * b * B
            qs.hist = eng.stats_dev()
            rc = lib.shl_place_dev(ctx, db.d_arena, db.d_off + 8 * b * B, db.d_len + 4 * b * B, m, C.byref(qs))
            assert rc == 0, lib.shl_last_error(ctx)
            eng.stats_allreduce()
            timed_units[0] += m

    results, total_ms, total_units, all_clocks, launches = {}, 0.0, 0, [], 0
    node_base = None
    k_ms = None
    for vi, (name, hosts) in enumerate(variants):
        eng.ring_build(hosts)                                # full rebuild, as a server-list edit does (Server.py:81-83)
        units = [0]
        for _ in range(args.warmup if vi == 0 else 0):       # warm-up passes on the first ring only
            place_all(0, [0])
        eng.stats_join()
        ms, clocks, ln = job.timed(lambda i: place_all(0 if vi == 0 else 1, units), 1, 0, eng.stats_join)
        total_ms += ms; total_units += job.sum_over_ranks(units[0]); all_clocks.append(clocks); launches += ln
        node = np.zeros(n_share, np.uint32)
        eng.d2h(node, d_node[0 if vi == 0 else 1]); eng.sync()
        hist = np.bincount(node, minlength=len(hosts))
        res = {'ms': ms, 'keys_per_s': job.sum_over_ranks(n_share) / (ms / 1e3), 'hist_sum': job.sum_over_ranks(int(hist.sum()))}
        names = np.array(hosts, dtype=object)
        if vi == 0:
            node_base = node
            eng.event(2)
            qs.node, qs.hist = d_node[1], eng.stats_dev()
            lib.shl_place_dev(ctx, db.d_arena, db.d_off, db.d_len, min(B, n_share), C.byref(qs))
            eng.event(3); eng.sync()
            k_ms = eng.elapsed_ms(2, 3)
            eng.stats_allreduce()
        else:
            moved = int((names[node] != np.array(base_hosts, dtype=object)[node_base]).sum())      # node IDENTITY (host string), not index
            res['moved_fraction'] = job.sum_over_ranks(moved) / args.total
            res['expected_about'] = 2.0 / 1024
        if rank == 0 and not args.no_cpu:                    # the C oracle as the checker: a strided sample
            from tests import oracle_lib as O
            idx = np.arange(0, min(n_share, 1 << 21), 32)
            data, off, ln_ = host_corpus(SEEDS[5], first, int(idx[-1]) + 1, 1)
            vals, nodes = O.orc_ring(hosts, [11211] * len(hosts))
            oq = O.orc_place(Arena(data, off, ln_), vals, nodes, len(hosts), np.ascontiguousarray(off[idx]), np.ascontiguousarray(ln_[idx]))
            bad = int((oq.node[:len(idx)] != node[idx]).sum())
            assert bad == 0, (name, bad)
            res['oracle'] = {'compared_keys': len(idx), 'mismatches': bad}
        results[name] = res
    # e2e: host keys through shl_place_batch
    m = min(B, n_share)
    ar = eng.corpus_host(SEEDS[5], first, m, 1)
    qout = PlaceOut(m, len(variants[-1][1]), alloc=eng.pinned, want_md5=False)

    def step_e2e(i):
        qout.hist[:] = 0
        eng.place(ar, out=qout, sync=True)

    e_steps = 5
    e_ms, e_clocks, _ = job.timed(step_e2e, e_steps, 1)
    assert int(qout.hist.sum()) == m
    metric, unit = metric_unit(5)
    value = total_units / (total_ms / 1e3)
    e2e = m * world * e_steps / (e_ms / 1e3)
    peak, peak_src = measured_peak()
    kb = int(db.off[m - 1] + db.len[m - 1])
    alg = kb + m * (12 + 8)
    achieved = alg / (k_ms / 1e3) / 1e9
    steps = batches * len(variants)
    line = {'metric': metric, 'value': value, 'unit': unit, 'n_gpus': world, 'steps': steps, 'warmup': args.warmup * batches,
            'ms_per_step': total_ms / steps, 'higher_is_better': True, 'scaling': 'strong', 'vs_baseline': None, 'dtype': 'u8',
            'data': 'synthetic', 'config': config_dict(args, world), 'clocks': all_clocks[0], 'gpu_launches': launches,
            'variants': results, 'mean_key_bytes': kb / m,
            'e2e': {'value': e2e, 'unit': unit, 'h2d_bytes_per_step': ar.nbytes + m * 12, 'd2h_bytes_per_step': m * 8 + 8 * 1024,
                    'steps': e_steps, 'ms_per_step': e_ms / e_steps, 'clocks': e_clocks,
                    'note': 'shl_place_batch on pinned host arrays, %d keys per rank per step' % m},
            'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                         'traffic': ncu_traffic('k_place_dram_bytes_per_launch'), 'kernel': 'shl::k_place', 'kernel_ms': k_ms,
                         'algorithmic_bytes_per_launch': alg, 'bytes_per_key': alg / m, 'peak_source': peak_src,
                         'note': 'MD5 is ~400 integer ops per 64-byte block: the kernel is issue-bound, the HBM fraction is what it is'}}
    if rank == 0 and world == 1 and not args.no_cpu:
        line['cpu_baseline'] = cpu_baseline_single(5, unit)
    job.barrier()
    eng.close()
    return line if rank == 0 else None


# ------------------------------------------------------------------------------------- the drop-in class surface
def facade_rates(eng, ar, n=1 << 16):
    """requests/s through the reference-shaped Python surface on the config-2 corpus: one
    HttpParser.parse_many call (batch entry of the drop-in class) and the reactor shim's tick."""
    from shellac_b200.reactor import EngineRouter, ReactorShim
    from shellac_b200.server import HttpParser
    from shellac_b200.server.HttpParser import DeviceBackend
    n = min(n, ar.n)
    blob = ar.data[:int(ar.off[n - 1] + ar.len[n - 1])].tobytes()
    msgs = [blob[int(o):int(o) + int(l)] for o, l in zip(ar.off[:n], ar.len[:n])]
    be = DeviceBackend(eng)
    HttpParser.parse_many(msgs[:1024], backend=be)
    t0 = time.perf_counter()
    ps = HttpParser.parse_many(msgs, backend=be)
    dt = time.perf_counter() - t0
    t1 = time.perf_counter()
    touched = sum(1 for p in ps[::16] if p.message_complete() and p.url() is not None and len(p.headers()) > 0)
    dt_touch = time.perf_counter() - t1
    assert touched == len(ps[::16])
    t3 = time.perf_counter()
    for m in msgs[:1000]:                       # the reference's own driving loop (test/HttpParserTests.py:36-39), one object per message
        p, data = HttpParser(backend=be), m
        while data and not p.message_complete():
            c = p.parse(data, len(data))
            data = data[c:]
    dt_inc = time.perf_counter() - t3
    shim = ReactorShim(EngineRouter(eng, msgs=512, nbytes=1 << 19))
    for fd in range(512):
        shim.feed(fd, msgs[fd])
    shim.tick()
    ticks = []
    for rep in range(5):
        for fd in range(512):
            shim.feed(fd, msgs[512 * (rep + 1) + fd])
        t2 = time.perf_counter()
        done = shim.tick()
        ticks.append(time.perf_counter() - t2)
        assert len(done) == 512
    return {'parse_many_requests_per_s': n / dt, 'parse_many_n': n,
            'parse_many_then_touch_every_16th_us_per_touched': dt_touch / max(touched, 1) * 1e6,
            'incremental_parse_us_per_message': dt_inc / 1000 * 1e6,
            'reactor_tick_512_requests_us': float(np.median(ticks)) * 1e6,
            'reactor_tick_us_per_request': float(np.median(ticks)) * 1e6 / 512,
            'note': 'host Python on one core + the device call; parser objects materialise their fields on first access'}


# ------------------------------------------------------------------------------------- main
class StdoutToStderr(object):
    """Everything a library prints on fd 1 while we work (NCCL's version banner, for one) goes to
    stderr, so that stdout carries exactly ONE line: the JSON result."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)
        return self

    def __exit__(self, *exc):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=None, help='timed steps (configs 4/5 default to the whole job)')
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--config', type=int, default=2, choices=[2, 3, 4, 5])
    ap.add_argument('--batch', type=int, default=None)
    ap.add_argument('--total', type=int, default=100_000_000, help='configs 4/5: units in the whole job')
    ap.add_argument('--stats-every', type=int, default=1,
                    help='all-reduce the per-node / per-outcome counters every S steps (0: once, at the end of the timed steps)')
    ap.add_argument('--no-cpu', action='store_true', help='skip the cpu_baseline leg and the oracle checks')
    args = ap.parse_args()
    if args.batch is None:
        args.batch = {2: BATCH, 3: 8192, 4: BATCH, 5: 1 << 22}[args.config]
    if args.steps is None:
        args.steps = {2: 20, 3: 10, 4: 0, 5: 0}[args.config] if args.impl == 'ours' else 5
    if args.impl == 'reference':
        if args.steps <= 0:
            args.steps = 5
        return run_reference(args)
    args.warmup = max(args.warmup, 3)
    with StdoutToStderr():
        line = {2: run_route, 3: run_frame, 4: run_route, 5: run_remap}[args.config](args)
    if line is not None:
        print(json.dumps(line), flush=True)
    return 0


if __name__ == '__main__':
    sys.exit(main())
