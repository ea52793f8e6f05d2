"""Engine: the Python owner of one shl_ctx (one GPU), the object a Shellac server would hold where
it holds `self._mc = pylibmc.Client([...], binary=True, behaviors={'ketama': True})` today
(Server.py:81-83).  It packs batches into pinned arenas and calls the C ABI through ctypes.

No CPU fallback: constructing an Engine without the built library or without a CUDA device raises.
"""
import ctypes as C

import numpy as np

from . import _native as N
from .batch import Arena, BodyOut, FrameOut, ParseOut, PlaceOut  # noqa: F401


def _check(lib, ctx, rc, what):
    if rc != 0:
        msg = lib.shl_last_error(ctx)
        raise N.ShellacError('%s failed (%d): %s' % (what, rc, msg.decode() if msg else '?'))


class Engine(object):
    def __init__(self, device=0, max_msgs=1 << 16, max_bytes=1 << 26, max_records=0):
        self.lib = N.lib()
        if self.lib.shl_device_count() <= 0:
            raise N.ShellacError('no CUDA device: shellac_b200 has no CPU fallback')
        h = C.c_void_p()
        rc = self.lib.shl_ctx_create(device, max_msgs, max_bytes, max_records, C.byref(h))
        _check(self.lib, None, rc, 'shl_ctx_create')
        self.ctx = h
        self.device, self.max_msgs, self.max_bytes = device, max_msgs, max_bytes
        self.max_records = max_records if max_records else max(max_bytes // 16, 1024)
        self._pinned = []
        self.n_hosts = 0

    # -- memory -----------------------------------------------------------------------------
    def pinned(self, n, dtype):
        """A page-locked numpy array (zero-filled) that lives as long as the engine.  The calls
        below that are not given `out=` allocate their result arrays this way on EVERY call: a
        long-lived caller passes `out=` (see PinnedBatch) or frees with release_pinned(mark)."""
        dt = np.dtype(dtype)
        nbytes = max(int(n), 1) * dt.itemsize
        p = C.c_void_p()
        _check(self.lib, self.ctx, self.lib.shl_pinned_alloc(C.byref(p), nbytes), 'shl_pinned_alloc')
        self._pinned.append(p)
        buf = (C.c_uint8 * nbytes).from_address(p.value)
        a = np.frombuffer(buf, dtype=np.uint8)
        a[:] = 0
        return a.view(dt)

    def release_pinned(self, mark):
        """Free every pinned array allocated after `mark = len(self._pinned)` (the arrays die)."""
        self.sync()
        while len(self._pinned) > mark:
            self.lib.shl_pinned_free(self._pinned.pop())

    def pack(self, messages, align=1):
        return Arena.pack(messages, alloc=self.pinned, align=align)

    def dev_alloc(self, nbytes):
        p = C.c_void_p()
        _check(self.lib, self.ctx, self.lib.shl_device_alloc(self.ctx, C.byref(p), int(nbytes)), 'shl_device_alloc')
        return p.value

    def dev_free(self, dptr):
        _check(self.lib, self.ctx, self.lib.shl_device_free(self.ctx, dptr), 'shl_device_free')

    def h2d(self, dptr, arr, nbytes=None):
        n = arr.nbytes if nbytes is None else int(nbytes)
        _check(self.lib, self.ctx, self.lib.shl_memcpy_h2d(self.ctx, dptr, arr.ctypes.data, n), 'shl_memcpy_h2d')

    def d2h(self, arr, dptr, nbytes=None):
        n = arr.nbytes if nbytes is None else int(nbytes)
        _check(self.lib, self.ctx, self.lib.shl_memcpy_d2h(self.ctx, arr.ctypes.data, dptr, n), 'shl_memcpy_d2h')

    def memset(self, dptr, value, nbytes):
        _check(self.lib, self.ctx, self.lib.shl_memset_dev(self.ctx, dptr, value, int(nbytes)), 'shl_memset_dev')

    def sync(self):
        _check(self.lib, self.ctx, self.lib.shl_sync(self.ctx), 'shl_sync')

    def event(self, slot):
        _check(self.lib, self.ctx, self.lib.shl_event_record(self.ctx, slot), 'shl_event_record')

    def elapsed_ms(self, a, b):
        ms = C.c_float()
        _check(self.lib, self.ctx, self.lib.shl_event_elapsed_ms(self.ctx, a, b, C.byref(ms)), 'shl_event_elapsed_ms')
        return float(ms.value)

    def launches(self):
        return int(self.lib.shl_launch_count(self.ctx))

    def set_force_scalar(self, on):
        _check(self.lib, self.ctx, self.lib.shl_set_force_scalar(self.ctx, 1 if on else 0), 'shl_set_force_scalar')

    # -- ring (pylibmc.Client(servers, behaviors={'ketama': True}), Server.py:81-83) ----------
    def ring_build(self, hosts, ports=None, weights=None, mode=N.RING_KETAMA):
        n = len(hosts)
        arr = (C.c_char_p * n)(*[h.encode() if isinstance(h, str) else h for h in hosts])
        p = np.asarray(ports if ports is not None else [11211] * n, dtype=np.uint32)
        w = np.asarray(weights, dtype=np.uint32) if weights is not None else None
        rc = self.lib.shl_ring_build(self.ctx, arr, p.ctypes.data, w.ctypes.data if w is not None else None, n, mode)
        _check(self.lib, self.ctx, rc, 'shl_ring_build')
        self.n_hosts = n

    def ring_export(self):
        npts, nh = C.c_uint32(), C.c_uint32()
        _check(self.lib, self.ctx, self.lib.shl_ring_size(self.ctx, C.byref(npts), C.byref(nh)), 'shl_ring_size')
        v, nd = np.zeros(max(npts.value, 1), np.uint32), np.zeros(max(npts.value, 1), np.uint32)
        _check(self.lib, self.ctx, self.lib.shl_ring_export(self.ctx, v.ctypes.data, nd.ctypes.data, npts.value), 'shl_ring_export')
        return v[:npts.value], nd[:npts.value]

    # -- host-buffer entry points ------------------------------------------------------------
    def _hdr_cap(self, arena, cap):
        return int(cap) if cap else min(self.max_records, max(64, arena.nbytes // 4 + 64))

    def parse(self, arena, hdr_cap=0, sync=True, out=None):
        out = out or ParseOut(arena.n, self._hdr_cap(arena, hdr_cap), alloc=self.pinned)
        s = out.struct()
        rc = self.lib.shl_parse_batch(self.ctx, arena.data.ctypes.data, arena.nbytes, arena.off.ctypes.data,
                                      arena.len.ctypes.data, arena.n, C.byref(s))
        _check(self.lib, self.ctx, rc, 'shl_parse_batch')
        if sync:
            self.sync()
        return out

    def frame(self, arena, pout, chunk_cap=0, reuse_arena=False, sync=True, out=None):
        cap = int(chunk_cap) if chunk_cap else min(self.max_records, max(64, arena.nbytes // 8 + 64))
        out = out or FrameOut(arena.n, cap, alloc=self.pinned)
        s = out.struct()
        rc = self.lib.shl_frame_batch(self.ctx, None if reuse_arena else arena.data.ctypes.data, arena.nbytes,
                                      arena.off.ctypes.data, arena.len.ctypes.data, arena.n,
                                      pout.status.ctypes.data, pout.first_line_len.ctypes.data,
                                      pout.headers_len.ctypes.data, C.byref(s))
        _check(self.lib, self.ctx, rc, 'shl_frame_batch')
        if sync:
            self.sync()
        return out

    def route(self, arena, hdr_cap=0, sync=True, pout=None, qout=None):
        pout = pout or ParseOut(arena.n, self._hdr_cap(arena, hdr_cap), alloc=self.pinned)
        qout = qout or PlaceOut(arena.n, self.n_hosts, alloc=self.pinned)
        ps, qs = pout.struct(), qout.struct()
        rc = self.lib.shl_route_batch(self.ctx, arena.data.ctypes.data, arena.nbytes, arena.off.ctypes.data,
                                      arena.len.ctypes.data, arena.n, C.byref(ps), C.byref(qs))
        _check(self.lib, self.ctx, rc, 'shl_route_batch')
        if sync:
            self.sync()
        return pout, qout

    def route_compact(self, arena, hdr_cap=0, want_md5=False, sync=True, out=None):
        """shl_route_compact_batch: (rec[n] of N.REC_DTYPE, hdr records, hdr_used, md5 or None, hist).
        `out` = the tuple a previous call returned (the pinned arrays are reused)."""
        n = arena.n
        cap = self._hdr_cap(arena, hdr_cap)
        if out is None:
            out = (self.pinned(n, N.REC_DTYPE), self.pinned(cap, N.HDR_DTYPE), self.pinned(1, np.uint64),
                   self.pinned(16 * n, np.uint8) if want_md5 else None, self.pinned(max(self.n_hosts, 1), np.uint64))
        rec, hdr, used, md5, hist = out
        rc = self.lib.shl_route_compact_batch(self.ctx, arena.data.ctypes.data, arena.nbytes, arena.off.ctypes.data,
                                              arena.len.ctypes.data, n, rec.ctypes.data, hdr.ctypes.data, len(hdr), used.ctypes.data,
                                              md5.ctypes.data if md5 is not None else None, hist.ctypes.data)
        _check(self.lib, self.ctx, rc, 'shl_route_compact_batch')
        if sync:
            self.sync()
        return out

    def place(self, arena, key_off=None, key_len=None, sync=True, out=None):
        key_off = arena.off if key_off is None else key_off
        key_len = arena.len if key_len is None else key_len
        n = len(key_off)
        out = out or PlaceOut(n, self.n_hosts, alloc=self.pinned)
        s = out.struct()
        rc = self.lib.shl_place_batch(self.ctx, arena.data.ctypes.data, arena.nbytes, key_off.ctypes.data,
                                      key_len.ctypes.data, n, C.byref(s))
        _check(self.lib, self.ctx, rc, 'shl_place_batch')
        if sync:
            self.sync()
        return out

    def body(self, arena, body_cap=0, sync=True, out=None):
        """Gather the bodies of the batch just parsed + framed on this engine (shl_body_batch)."""
        cap = int(body_cap) if body_cap else arena.nbytes + 16 * arena.n + 64
        out = out or BodyOut(arena.n, cap, alloc=self.pinned)
        s = out.struct()
        _check(self.lib, self.ctx, self.lib.shl_body_batch(self.ctx, arena.n, C.byref(s)), 'shl_body_batch')
        if sync:
            self.sync()
        return out

    def group(self, n):
        """Requests of the batch just routed/placed (with a histogram), grouped by node:
        (start[n_hosts + 1], order[n]); node k owns order[start[k]:start[k + 1]]."""
        g = getattr(self, '_group_bufs', None)          # pinned outputs are kept and only grow (a reactor calls this every tick)
        if g is None or len(g[0]) < self.n_hosts + 1 or len(g[1]) < n:
            g = (self.pinned(max(self.n_hosts + 1, 64), np.uint32), self.pinned(max(n, 1024), np.uint32))
            self._group_bufs = g
        start, order = g[0][:self.n_hosts + 1], g[1]
        _check(self.lib, self.ctx, self.lib.shl_group_batch(self.ctx, n, start.ctypes.data, order.ctypes.data), 'shl_group_batch')
        self.sync()
        return start, order[:n]

    def mc_frames(self, n, n_placed, op=N.MC_GET, ttl=0, value_len=None, frames_cap=0):
        """Memcached binary-protocol request frames of the batch just routed + grouped, in the
        node-grouped order: (frame_off[n_placed + 1], frames bytes).  Node k's stream is
        frames[frame_off[start[k]]:frame_off[start[k + 1]]] with `start` from group()."""
        cap = int(frames_cap) if frames_cap else self.max_bytes + 32 * self.max_msgs
        bufs = getattr(self, '_mc_bufs', None)
        if bufs is None or bufs[0] < n_placed + 1 or bufs[1] < cap:
            bufs = (max(n_placed + 1, 1024), cap, self.pinned(max(n_placed + 1, 1024), np.uint64), self.pinned(cap, np.uint8),
                    self.pinned(1, np.uint64))
            self._mc_bufs = bufs
        _, _, foff, frames, used = bufs
        s = N.McOutStruct()
        s.frame_off, s.frames, s.frames_cap, s.frames_used = foff.ctypes.data, frames.ctypes.data, cap, used.ctypes.data
        vl = None
        if value_len is not None:
            vl = np.ascontiguousarray(value_len, dtype=np.uint32)
        rc = self.lib.shl_mc_frames_batch(self.ctx, n, n_placed, op, ttl, vl.ctypes.data if vl is not None else None, C.byref(s))
        _check(self.lib, self.ctx, rc, 'shl_mc_frames_batch')
        self.sync()
        return foff[:n_placed + 1], frames[:int(used[0])]

    def chunk_header(self, arena, out=None):
        n = arena.n
        hl, sz, ok = out if out is not None else (self.pinned(n, np.int32), self.pinned(n, np.uint64), self.pinned(n, np.uint8))
        rc = self.lib.shl_chunk_header_batch(self.ctx, arena.data.ctypes.data, arena.nbytes, arena.off.ctypes.data,
                                             arena.len.ctypes.data, n, hl.ctypes.data, sz.ctypes.data, ok.ctypes.data)
        _check(self.lib, self.ctx, rc, 'shl_chunk_header_batch')
        self.sync()
        return hl[:n], sz[:n], ok[:n]

    # -- counters across GPUs (shl_stats_* / shl_comm_*: the library owns the NCCL communicator) --
    def comm_init(self, comm_id, rank, world):
        """Join the job's communicator: `comm_id` is the 128 bytes rank 0 got from comm_unique_id()."""
        buf = (C.c_uint8 * N.COMM_ID_BYTES).from_buffer_copy(bytes(comm_id))
        _check(self.lib, self.ctx, self.lib.shl_comm_init(self.ctx, buf, int(rank), int(world)), 'shl_comm_init')

    def comm_unique_id(self):
        buf = (C.c_uint8 * N.COMM_ID_BYTES)()
        _check(self.lib, self.ctx, self.lib.shl_comm_unique_id(buf, N.COMM_ID_BYTES), 'shl_comm_unique_id')
        return bytes(buf)

    def stats_dev(self):
        """Device address of the CURRENT counters buffer (changes with every stats_allreduce)."""
        return self.lib.shl_stats_dev(self.ctx)

    def stats_allreduce(self):
        _check(self.lib, self.ctx, self.lib.shl_stats_allreduce(self.ctx), 'shl_stats_allreduce')

    def stats_join(self):
        _check(self.lib, self.ctx, self.lib.shl_stats_join(self.ctx), 'shl_stats_join')

    def stats_fetch(self):
        """[placement histogram n_hosts | outcome counts 64] of the most recent stats_allreduce."""
        n = int(self.lib.shl_stats_len(self.ctx))
        out = np.zeros(n, np.uint64)
        _check(self.lib, self.ctx, self.lib.shl_stats_fetch(self.ctx, out.ctypes.data, n), 'shl_stats_fetch')
        return out

    def allreduce_u64(self, values, op='sum'):
        """Small blocking all-reduce over the job's ranks (after draining this engine's streams)."""
        v = np.ascontiguousarray(values, dtype=np.uint64).copy()
        rc = self.lib.shl_comm_allreduce_u64(self.ctx, v.ctypes.data, len(v), 1 if op == 'max' else 0)
        _check(self.lib, self.ctx, rc, 'shl_comm_allreduce_u64')
        return v

    def barrier(self):
        self.allreduce_u64([1])

    def code_hist(self, reset=True):
        out = np.zeros(N.N_CODES, np.uint64)
        _check(self.lib, self.ctx, self.lib.shl_code_hist(self.ctx, out.ctypes.data, 1 if reset else 0), 'shl_code_hist')
        return out

    # -- synthetic corpora (BASELINE.json configs 2/4/5) --------------------------------------
    def corpus_host(self, seed, first, n, kind=0):
        """Generate messages [first, first+n) on the HOST into a pinned arena."""
        ln = self.pinned(n, np.uint32)
        _check(self.lib, self.ctx, self.lib.shl_corpus_lengths_host(seed, first, n, kind, ln.ctypes.data), 'corpus_lengths')
        off = self.pinned(n, np.uint64)
        if n > 1:
            np.cumsum(ln[:n - 1], dtype=np.uint64, out=off[1:n])
        total = int(off[n - 1] + ln[n - 1]) if n else 0
        data = self.pinned(total + 64, np.uint8)
        _check(self.lib, self.ctx, self.lib.shl_corpus_fill_host(seed, first, n, kind, off.ctypes.data, data.ctypes.data), 'corpus_fill')
        return Arena(data, off[:n], ln[:n])

    # -- device-resident batches (arenas that already live in HBM; kernel-only timing) ---------
    def corpus_dev(self, seed, first, n, kind=0):
        """Generate messages [first, first+n) directly in HBM.  Returns a DeviceBatch."""
        d_len = self.dev_alloc(n * 4)
        _check(self.lib, self.ctx, self.lib.shl_corpus_lengths_dev(self.ctx, seed, first, n, kind, d_len), 'corpus_lengths_dev')
        ln = np.zeros(n, np.uint32)
        self.d2h(ln, d_len)
        self.sync()
        off = np.zeros(n, np.uint64)
        if n > 1:
            np.cumsum(ln[:n - 1], dtype=np.uint64, out=off[1:n])
        total = int(off[n - 1] + ln[n - 1]) if n else 0
        d_off, d_arena = self.dev_alloc(n * 8), self.dev_alloc(total + 64)
        self.h2d(d_off, off)
        self.memset(d_arena + total, 0, 64)
        _check(self.lib, self.ctx, self.lib.shl_corpus_fill_dev(self.ctx, seed, first, n, kind, d_off, d_arena), 'corpus_fill_dev')
        self.sync()
        return DeviceBatch(self, d_arena, d_off, d_len, n, total, off, ln)

    def upload(self, arena):
        """Copy a host Arena into HBM.  Returns a DeviceBatch."""
        n, total = arena.n, arena.nbytes
        d_arena, d_off, d_len = self.dev_alloc(total + 64), self.dev_alloc(n * 8), self.dev_alloc(n * 4)
        self.memset(d_arena + total, 0, 64)
        self.h2d(d_arena, arena.data, total); self.h2d(d_off, arena.off); self.h2d(d_len, arena.len)
        self.sync()
        return DeviceBatch(self, d_arena, d_off, d_len, n, total, np.array(arena.off), np.array(arena.len))

    def close(self):
        if getattr(self, 'ctx', None):
            self.lib.shl_ctx_destroy(self.ctx)
            self.ctx = None
            for p in self._pinned:
                self.lib.shl_pinned_free(p)
            self._pinned = []

    def __del__(self):
        try:
            self.close()
        except Exception:   # noqa: BLE001 - interpreter shutdown
            pass


class PinnedBatch(object):
    """A pinned arena plus result arrays that live between calls and only grow.  Page-locking
    memory costs more than parsing a reactor tick's worth of requests, so the callers that run
    many small batches (the reactor shim, the incremental HttpParser facade) pack into this."""

    def __init__(self, eng, msgs=64, nbytes=1 << 16):
        self.eng = eng
        self._cap_n = self._cap_b = 0
        self._hosts = -1
        self._fout = None
        self.grow(msgs, nbytes)

    def grow(self, n, nbytes):
        eng = self.eng
        if n > self._cap_n or nbytes > self._cap_b:
            n, nbytes = max(n, 2 * self._cap_n, 64), max(nbytes, 2 * self._cap_b, 1 << 16)
            self.data = eng.pinned(nbytes + 64, np.uint8)
            self.off, self.len = eng.pinned(n, np.uint64), eng.pinned(n, np.uint32)
            self.pout = ParseOut(n, 24 * n + 64, alloc=eng.pinned)
            self._cap_n, self._cap_b, self._hosts, self._fout = n, nbytes, -1, None
        if eng.n_hosts != self._hosts:
            self.qout = PlaceOut(self._cap_n, max(eng.n_hosts, 1), alloc=eng.pinned)
            self._hosts = eng.n_hosts

    def bout(self, cap):
        """Pinned body-gather outputs, kept and only growing."""
        b = getattr(self, '_bout', None)
        if b is None or b.body_cap < cap or b.n < self._cap_n:
            b = BodyOut(self._cap_n, max(cap, 1 << 16), alloc=self.eng.pinned)
            self._bout = b
        return b

    @property
    def fout(self):
        if self._fout is None:
            self._fout = FrameOut(self._cap_n, self._cap_b // 5 + 64, alloc=self.eng.pinned)
        return self._fout

    def pack(self, messages):
        msgs = [m.encode('latin-1') if isinstance(m, str) else bytes(m) for m in messages]
        n = len(msgs)
        lens = np.fromiter((len(m) for m in msgs), dtype=np.uint32, count=n)
        total = int(lens.sum())
        self.grow(n, total)
        self.data[:total] = np.frombuffer(b''.join(msgs), dtype=np.uint8)
        self.data[total:total + 64] = 0
        if n:
            self.len[:n] = lens
            self.off[0] = 0
        if n > 1:
            np.cumsum(lens[:-1], dtype=np.uint64, out=self.off[1:n])
        return Arena(self.data, self.off[:n], self.len[:n])


class DeviceBatch(object):
    """An arena + index resident in HBM, with device-side result arrays on demand."""

    def __init__(self, eng, d_arena, d_off, d_len, n, nbytes, off, ln):
        self.eng, self.d_arena, self.d_off, self.d_len = eng, d_arena, d_off, d_len
        self.n, self.nbytes, self.off, self.len = n, nbytes, off, ln
        self._bufs = [d_arena, d_off, d_len]
        self.p = self.q = None

    def _alloc(self, nbytes):
        p = self.eng.dev_alloc(nbytes)
        self._bufs.append(p)
        return p

    def parse_struct(self, hdr_cap=0):
        if self.p is None:
            self.hdr_cap = int(hdr_cap) if hdr_cap else self.n * 24
            self.p = N.ParseOutStruct()
            self.p_fields = {}
            for name, dt in N.PARSE_FIELDS:
                self.p_fields[name] = self._alloc(self.n * np.dtype(dt).itemsize)
                setattr(self.p, name, self.p_fields[name])
            self.d_hdr, self.d_used = self._alloc(self.hdr_cap * 8), self._alloc(8)
            self.p.hdr, self.p.hdr_cap, self.p.hdr_used = self.d_hdr, self.hdr_cap, self.d_used
        return self.p

    def place_struct(self, n_hosts, md5=True):
        if self.q is None:
            self.q = N.PlaceOutStruct()
            self.d_md5 = self._alloc(16 * self.n) if md5 else None
            self.d_hash, self.d_node = self._alloc(4 * self.n), self._alloc(4 * self.n)
            self.d_hist = self._alloc(8 * max(n_hosts, 1))
            self.n_hosts = n_hosts
            self.q.md5, self.q.hash, self.q.node, self.q.hist = self.d_md5, self.d_hash, self.d_node, self.d_hist
            self.eng.memset(self.d_hist, 0, 8 * max(n_hosts, 1))
        return self.q

    def route(self, n_hosts):
        """shl_route_dev on the resident batch (asynchronous)."""
        e = self.eng
        ps, qs = self.parse_struct(), self.place_struct(n_hosts)
        e.memset(self.d_used, 0, 8)
        _check(e.lib, e.ctx, e.lib.shl_route_dev(e.ctx, self.d_arena, self.d_off, self.d_len, self.n, C.byref(ps), C.byref(qs)), 'shl_route_dev')

    def place(self, n_hosts, d_key_off=None, d_key_len=None, md5=False):
        e = self.eng
        qs = self.place_struct(n_hosts, md5=md5)
        _check(e.lib, e.ctx, e.lib.shl_place_dev(e.ctx, self.d_arena, d_key_off or self.d_off, d_key_len or self.d_len, self.n, C.byref(qs)), 'shl_place_dev')

    def fetch(self, dptr, dtype, count):
        a = np.zeros(count, dtype)
        self.eng.d2h(a, dptr)
        self.eng.sync()
        return a

    def field(self, name):
        dt = dict(N.PARSE_FIELDS)[name]
        return self.fetch(self.p_fields[name], dt, self.n)

    def free(self):
        self.eng.sync()
        for p in self._bufs:
            self.eng.dev_free(p)
        self._bufs = []
