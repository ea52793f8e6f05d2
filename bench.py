#!/usr/bin/env python
"""bench.py — BASELINE.json's metric on BASELINE.json's config 2, one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path (tokenise + key + MD5 + ketama placement, fused in one
kernel) over one batch of 2^20 synthetic GET requests (~512 B, 8-20 header lines, 16x100-point
ring) per GPU; ranks take disjoint slices of the same counter-based corpus (weak scaling) and
all-reduce the per-node placement histogram + per-outcome counts (NCCL) once per step.

  value : whole-job requests/s with the arena already resident in HBM (kernel + all-reduce).
  e2e   : the same metric through shl_route_batch with HOST (pinned) buffers: H2D of the arena and
          index, the kernel, D2H of every result array, all inside the timed region.
  roofline : algorithmic bytes of the route kernel / its CUDA-event duration vs the measured HBM
          copy bandwidth in MEASURED_PEAKS.json.
  cpu_baseline : oracle/py_port.py (the reference's algorithm in CPython, as the reference is) on
          one host core over a bounded sample; the C oracle's figure is given beside it.
  --impl reference : the CPython port on all host cores (the reference itself cannot travel to
          the GPU box: /root/reference does not exist there).
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED = 20130725
BATCH = 1 << 20
HOSTS = ['10.0.0.%d' % (i + 1) for i in range(16)]
METRIC = 'parsed+routed requests/s (HTTP/1.1 tokenise + cache key + ketama/MD5 placement)'
WORKLOAD = ('config 2: 2^20 synthetic HTTP/1.1 GET requests per GPU per step, ~512 B, 8-20 header lines, '
            '16-node x 100-point ketama ring')
# algorithmic bytes per request (DESIGN.md "Data layout"): message bytes + 12 B index in,
# 8 B per header record + 48 B fixed record + 24 B placement (md5 16, hash 4, node 4) out
FIXED_BYTES = 12 + 48 + 24


def log(*a):
    print(*a, file=sys.stderr, flush=True)


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
            time.sleep(0.002)

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


def ncu_traffic():
    """dram bytes per launch of the route kernel from the committed ncu --set full capture."""
    p = os.path.join(ROOT, 'profiles', 'traffic.json')
    if os.path.exists(p):
        try:
            return json.load(open(p)).get('k_route_dram_bytes_per_launch')
        except Exception:           # noqa: BLE001
            return None
    return None


# ------------------------------------------------------------------------------------- CPU arms
def _cpu_messages(n, first=0):
    from shellac_b200 import _native as N
    lib = N.lib()
    ln = np.zeros(n, np.uint32)
    lib.shl_corpus_lengths_host(SEED, first, n, 0, ln.ctypes.data)
    off = np.zeros(n, np.uint64)
    np.cumsum(ln[:-1], dtype=np.uint64, out=off[1:])
    data = np.zeros(int(off[-1] + ln[-1]) + 64, np.uint8)
    lib.shl_corpus_fill_host(SEED, first, n, 0, off.ctypes.data, data.ctypes.data)
    blob = data.tobytes()
    return [blob[int(o):int(o) + int(l)] for o, l in zip(off, ln)], int(ln.sum())


def _port():
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import py_port
    return py_port


_W = {}


def _worker_init(msgs):
    pp = _port()
    _W['pp'] = pp
    _W['ring'] = pp.build_ring(HOSTS)
    _W['msgs'] = msgs


def _worker_run(span):
    a, b = span
    v, nd = _W['ring']
    done, hist = _W['pp'].route_messages(_W['msgs'][a:b], v, nd, len(HOSTS))
    return done


def cpu_baseline_single(sample=300000):
    """The CPython port on ONE core over a bounded sample of the same corpus (+ the C oracle)."""
    pp = _port()
    msgs, nbytes = _cpu_messages(sample)
    v, nd = pp.build_ring(HOSTS)
    t0 = time.perf_counter()
    done, hist = pp.route_messages(msgs, v, nd, len(HOSTS))
    dt = time.perf_counter() - t0
    assert done == sample
    out = {'value': sample / dt, 'unit': 'requests/s', 'cores': 1, 'kind': 'port',
           'sample': '%d requests of the config-2 corpus (%.1f MB), oracle/py_port.py (CPython restatement of '
                     'HttpParser.py + url() + hashlib/bisect ketama), %.1f s' % (sample, nbytes / 1e6, dt),
           'header_gb_s': nbytes / dt / 1e9}
    try:    # the C restatement beside it: a much stronger CPU baseline than the reference's own language
        from tests import oracle_lib as O
        from shellac_b200.batch import Arena
        n = 1 << 18
        ar = Arena.pack(msgs[:n])
        vals, nodes = O.orc_ring(HOSTS, [11211] * len(HOSTS))
        t0 = time.perf_counter()
        o = O.orc_parse(ar, hdr_cap=n * 24)
        koff = (ar.off + o.sp1[:n] + 1).astype(np.uint64)
        klen = (o.sp2[:n] - o.sp1[:n] - 1).astype(np.uint32)
        O.orc_place(ar, vals, nodes, len(HOSTS), koff, klen)
        dtc = time.perf_counter() - t0
        out['c_port_value'] = n / dtc
        out['c_port_note'] = 'oracle/shl_oracle.c, 1 core, %d requests' % n
    except Exception as e:          # noqa: BLE001
        out['c_port_value'] = None
        log('C oracle timing failed: %r' % (e,))
    return out


def run_reference(args):
    """--impl reference: the reference's CPU algorithm (CPython port) on all host cores."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return 0
    import multiprocessing as mp
    cores = len(os.sched_getaffinity(0))
    per_core = 12000
    n = per_core * cores
    msgs, nbytes = _cpu_messages(n)
    spans = [(i * per_core, (i + 1) * per_core) for i in range(cores)]
    ctx = mp.get_context('fork')
    with ctx.Pool(cores, initializer=_worker_init, initargs=(msgs,)) as pool:
        for _ in range(args.warmup):
            pool.map(_worker_run, spans)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            done = sum(pool.map(_worker_run, spans))
            assert done == n
        dt = time.perf_counter() - t0
    v = n * args.steps / dt
    sample = ('each step = %d requests of the config-2 corpus (%d per core x %d cores), oracle/py_port.py '
              '(CPython restatement; the reference is pure Python and /root/reference is absent here)' % (n, per_core, cores))
    line = {'impl': 'reference', 'metric': METRIC, 'value': v, 'unit': 'requests/s', 'n_gpus': args.gpus,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': dt / args.steps * 1e3,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'u8', 'data': 'synthetic',
            'config': {'workload': WORKLOAD, 'ring': '16x100 ketama/MD5', 'batch_per_step': n},
            'header_gb_s': nbytes * args.steps / dt / 1e9,
            'cpu_baseline': {'value': v, 'unit': 'requests/s', 'cores': cores, 'kind': 'port', 'sample': sample},
            'e2e': {'value': v, 'unit': 'requests/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------- GPU arm
def run_ours(args):
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    from shellac_b200 import _native as N
    from shellac_b200 import build as B
    from shellac_b200.batch import ParseOut, PlaceOut
    from shellac_b200.engine import Engine
    import ctypes as C
    if rank == 0:
        B.build()
    torch = dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
        dist.barrier()
    n = args.batch
    numa = None
    if not os.environ.get('SHL_NO_BIND'):
        from shellac_b200.hostmem import bind_near_gpu
        numa = bind_near_gpu(local)          # before any pinned arena is allocated
    log('[rank %d] numa binding: %r, %d cpus allowed' % (rank, numa, len(os.sched_getaffinity(0))))
    eng = Engine(device=local, max_msgs=n, max_bytes=n * 640, max_records=n * 24)
    lib, ctx = eng.lib, eng.ctx
    eng.ring_build(HOSTS)
    t0 = time.perf_counter()
    ar = eng.corpus_host(SEED, rank * n, n, 0)           # pinned host arena, disjoint slice per rank
    log('[rank %d] corpus: %d requests, %.1f MB, generated in %.1f s' % (rank, n, ar.nbytes / 1e6, time.perf_counter() - t0))
    nbytes = ar.nbytes

    # ---- HBM-resident copies + device SoA outputs for the kernel-only number
    d_arena = eng.dev_alloc(nbytes + 64)
    d_off, d_len = eng.dev_alloc(n * 8), eng.dev_alloc(n * 4)
    eng.h2d(d_arena, ar.data, nbytes); eng.h2d(d_off, ar.off); eng.h2d(d_len, ar.len)
    hdr_cap = n * 24
    fields = {}
    for name, dt in N.PARSE_FIELDS:
        fields[name] = eng.dev_alloc(n * np.dtype(dt).itemsize)
    d_hdr, d_used = eng.dev_alloc(hdr_cap * 8), eng.dev_alloc(8)
    ps = N.ParseOutStruct()
    for name, _ in N.PARSE_FIELDS:
        setattr(ps, name, fields[name])
    ps.hdr, ps.hdr_cap, ps.hdr_used = d_hdr, hdr_cap, d_used
    qs = N.PlaceOutStruct()
    qs.md5, qs.hash, qs.node = eng.dev_alloc(16 * n), eng.dev_alloc(4 * n), eng.dev_alloc(4 * n)
    n_stat = len(HOSTS) + N.N_CODES
    if world > 1:
        from shellac_b200 import dist as D
        # two stats buffers [hist | code counts]: the all-reduce of step k runs on its own stream while
        # the kernel of step k+1 is already scanning (collective overlapped with compute)
        stats2 = [torch.zeros(n_stat, dtype=torch.int64, device='cuda:%d' % local) for _ in range(2)]
        ext = D.external_stream(eng, local)
        comm = torch.cuda.Stream(device='cuda:%d' % local)
        ev_ready = [torch.cuda.Event() for _ in range(2)]       # stats[k] written by the kernel
        ev_free = [torch.cuda.Event() for _ in range(2)]        # stats[k] reduced, may be overwritten
        for e in ev_free:
            e.record(comm)
        d_stats_list = [t.data_ptr() for t in stats2]
    else:
        d_stats_list = [eng.dev_alloc(n_stat * 8)]
    d_codes = lib.shl_code_hist_dev(ctx)
    # The path has no exchange step (requests are independent): ranks only share the per-node and
    # per-outcome COUNTERS.  They accumulate on the device over `--stats-every` steps (0 = the whole
    # timed region) and are then all-reduced on a side stream while the next kernel already runs.
    st = {'k': 0, 'since': 0, 'fresh': True}

    def reduce_stats():
        k = st['k']
        lib.shl_memcpy_d2d(ctx, d_stats_list[k] + len(HOSTS) * 8, d_codes, N.N_CODES * 8)
        st['last'] = k
        st['since'] = 0
        st['fresh'] = True
        if world > 1:
            ev_ready[k].record(ext)
            comm.wait_event(ev_ready[k])
            with torch.cuda.stream(comm):
                D.allreduce_stats(stats2[k])
                ev_free[k].record(comm)
            st['k'] = (k + 1) % len(d_stats_list)

    def step_resident():
        k = st['k']
        if st['fresh']:
            if world > 1:
                ext.wait_event(ev_free[k])
            eng.memset(d_stats_list[k], 0, n_stat * 8)
            eng.memset(d_codes, 0, N.N_CODES * 8)
            st['fresh'] = False
        qs.hist = d_stats_list[k]
        eng.memset(d_used, 0, 8)
        rc = lib.shl_route_dev(ctx, d_arena, d_off, d_len, n, C.byref(ps), C.byref(qs))
        assert rc == 0, lib.shl_last_error(ctx)
        st['since'] += 1
        if args.stats_every and st['since'] >= args.stats_every:
            reduce_stats()

    def flush_resident():
        if st['since']:
            reduce_stats()
        if world > 1:                       # the timed region ends when the last all-reduce has finished
            for e in ev_free:
                ext.wait_event(e)

    def barrier():
        eng.sync()
        if world > 1:
            torch.cuda.synchronize()
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps, warmup, flush=None):
        for _ in range(warmup):
            fn()
        if flush:
            flush()
        sampler = ClockSampler(local)
        sampler.start()
        barrier()
        sampler.armed.set()
        l0 = eng.launches()
        eng.event(0)
        for _ in range(steps):
            fn()
        if flush:
            flush()
        eng.event(1)
        barrier()
        ms = eng.elapsed_ms(0, 1)
        clocks = sampler.result()
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device='cuda:%d' % local)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, clocks, eng.launches() - l0

    ms, clocks, launches = timed(step_resident, args.steps, args.warmup, flush_resident)
    # kernel-only duration for the roofline: events around the kernel alone, averaged over the steps
    eng.sync()
    kms = []
    for _ in range(args.steps):
        eng.memset(d_used, 0, 8)
        qs.hist = d_stats_list[0]
        eng.event(2)
        lib.shl_route_dev(ctx, d_arena, d_off, d_len, n, C.byref(ps), C.byref(qs))
        eng.event(3)
        eng.sync()
        kms.append(eng.elapsed_ms(2, 3))
    k_ms = float(np.mean(kms))
    log('[rank %d] kernel alone %.4f ms, step %.4f ms (max over ranks)' % (rank, k_ms, ms / args.steps))

    # ---- check the resident result (outside the timed region): histogram adds up
    st['fresh'] = True
    step_resident()
    flush_resident()
    barrier()
    hs = np.zeros(n_stat, np.uint64)
    eng.d2h(hs, d_stats_list[st['last']]); eng.sync()
    total_hdrs = np.zeros(1, np.uint64)
    eng.d2h(total_hdrs, d_used); eng.sync()
    expect = n * (world if world > 1 else 1)
    assert int(hs[:len(HOSTS)].sum()) == expect, (hs[:len(HOSTS)].sum(), expect)
    assert int(hs[len(HOSTS) + N.S_OK]) == expect

    # ---- e2e: host buffers through shl_route_batch
    pout = ParseOut(n, hdr_cap, alloc=eng.pinned)
    qout = PlaceOut(n, len(HOSTS), alloc=eng.pinned)

    def step_e2e():
        qout.hist[:] = 0
        eng.route(ar, pout=pout, qout=qout, sync=True)

    e_ms, e_clocks, _ = timed(step_e2e, args.steps, max(args.warmup, 1))
    # PCIe roofline of this box, measured in the same run: one pinned copy each way of the step's bytes
    eng.sync()
    pcie = {}
    for name, fn in (('h2d', lambda: eng.h2d(d_arena, ar.data, nbytes)), ('d2h', lambda: eng.d2h(ar.data, d_arena, nbytes))):
        best = 1e9
        for _ in range(3):
            eng.event(4); fn(); eng.event(5); eng.sync()
            best = min(best, eng.elapsed_ms(4, 5))
        pcie[name + '_gb_s'] = nbytes / (best / 1e3) / 1e9
    assert int(qout.hist.sum()) == n and (pout.code() == 0).all()
    n_hdr = int(total_hdrs[0])
    h2d = nbytes + n * 12
    d2h = n * 48 + n_hdr * 8 + n * 24 + len(HOSTS) * 8 + 8

    total_req = n * world
    value = total_req * args.steps / (ms / 1e3)
    e2e = total_req * args.steps / (e_ms / 1e3)
    alg_bytes = nbytes + n * FIXED_BYTES + 8 * n_hdr
    peak, peak_src = measured_peak()
    achieved = alg_bytes / (k_ms / 1e3) / 1e9
    line = {
        'metric': METRIC, 'value': value, 'unit': 'requests/s', 'n_gpus': world, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': ms / args.steps, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'u8', 'data': 'synthetic',
        'config': {'workload': WORKLOAD, 'batch_per_gpu': n, 'mean_request_bytes': nbytes / n,
                   'mean_header_lines': n_hdr / n, 'ring': '16x100 ketama/MD5 (libmemcached 0.44 "ketama": True)',
                   'parallelism': 'dp%d (batch-sharded, no data-path collective; %d int64 counters all-reduced %s)' % (
                       world, n_stat, 'every %d steps' % args.stats_every if args.stats_every else 'once per timed region'),
                   'l2': 'inputs (%.0f MB per GPU) exceed the 126 MB L2; no flush needed' % (nbytes / 1e6)},
        'header_gb_s': nbytes * world * args.steps / (ms / 1e3) / 1e9,
        'clocks': clocks, 'gpu_launches': launches,
        'e2e': {'value': e2e, 'unit': 'requests/s', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                'header_gb_s': nbytes * world * args.steps / (e_ms / 1e3) / 1e9,
                'pcie_gb_s_h2d': h2d * args.steps / (e_ms / 1e3) / 1e9, 'ms_per_step': e_ms / args.steps,
                'pcie_probe': pcie, 'frac_of_pcie_h2d_roofline': (h2d * args.steps / (e_ms / 1e3) / 1e9) / pcie['h2d_gb_s'],
                'note': 'chunked over 4 streams: H2D, kernel and D2H of different chunks overlap; bound by the H2D link',
                'host_numa': 'process bound to NUMA node %d (%d cpus) of its GPU' % numa if numa else 'unbound',
                'clocks': e_clocks},
        'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak,
                     'traffic': ncu_traffic(), 'kernel': 'shl::k_route<true, true>', 'kernel_ms': k_ms,
                     'algorithmic_bytes_per_launch': alg_bytes, 'bytes_per_request': alg_bytes / n,
                     'peak_source': peak_src, 'frac_of_nominal_8TBs': achieved / 8000.0},
    }
    if rank == 0 and world == 1 and not args.no_cpu:
        line['cpu_baseline'] = cpu_baseline_single()
    eng.close()
    if world > 1:
        dist.destroy_process_group()
    return line if rank == 0 else None


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
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--batch', type=int, default=BATCH)
    ap.add_argument('--stats-every', type=int, default=0,
                    help='all-reduce the per-node / per-outcome counters every S steps (0: once, at the end of the timed steps)')
    ap.add_argument('--no-cpu', action='store_true', help='skip the cpu_baseline leg')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == 'ours' else args.warmup
    if args.impl == 'reference':
        return run_reference(args)
    with StdoutToStderr():
        line = run_ours(args)
    if line is not None:
        print(json.dumps(line), flush=True)
    return 0


if __name__ == '__main__':
    sys.exit(main())
