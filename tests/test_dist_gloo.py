"""The N>1 host logic on CPU: world_size-2 gloo run of the batch sharding and the histogram /
outcome-counter all-reduce (the only collective the path has), and the file rendezvous the ranks
use to share the library's NCCL id.  The per-rank placement comes from the oracle here and the sum
goes through gloo; on GPUs the library sums the kernel's device counters itself
(shl_stats_allreduce, NCCL)."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_total, q):
    import torch.distributed as dist
    from shellac_b200 import _native as N
    from shellac_b200 import dist as D
    from shellac_b200.batch import Arena
    from tests import oracle_lib as O
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    lo, hi = D.shard_range(n_total, rank, world)
    lib = N.lib()
    n = hi - lo
    ln = np.zeros(n, np.uint32)
    lib.shl_corpus_lengths_host(20130725, lo, n, 0, ln.ctypes.data)      # message i depends on (seed, i) only
    off = np.zeros(n, np.uint64)
    np.cumsum(ln[:-1], dtype=np.uint64, out=off[1:])
    data = np.zeros(int(off[-1] + ln[-1]) + 64, np.uint8)
    lib.shl_corpus_fill_host(20130725, lo, n, 0, off.ctypes.data, data.ctypes.data)
    ar = Arena(data, off, ln)
    po = O.orc_parse(ar)
    hosts = ['10.0.0.%d' % (i + 1) for i in range(16)]
    vals, nodes = O.orc_ring(hosts, [11211] * 16)
    koff = (ar.off + po.sp1[:n] + 1).astype(np.uint64)
    klen = (po.sp2[:n] - po.sp1[:n] - 1).astype(np.uint32)
    pl = O.orc_place(ar, vals, nodes, 16, koff, klen)
    codes = np.bincount(po.status[:n] & 63, minlength=N.N_CODES)
    import torch
    stats = D.pack_stats(pl.hist, codes)
    dist.all_reduce(torch.from_numpy(stats), op=dist.ReduceOp.SUM)      # in place: the tensor shares the array
    if rank == 0:
        q.put(stats.tolist())
    dist.barrier()
    dist.destroy_process_group()


def test_shard_ranges_cover_the_batch():
    from shellac_b200 import dist as D
    for n, w in ((10, 3), (1 << 20, 8), (5, 8), (0, 2)):
        r = [D.shard_range(n, k, w) for k in range(w)]
        assert r[0][0] == 0 and r[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(r, r[1:]))


def test_two_rank_gloo_allreduce_equals_single_rank():
    import torch.multiprocessing as mp
    from shellac_b200 import _native as N
    from shellac_b200 import build as B
    from shellac_b200.batch import Arena
    from tests import oracle_lib as O
    B.build()
    O.build_oracle()
    n_total, world, port = 6000, 2, _free_port()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, world, port, n_total, q)) for r in range(world)]
    for p in ps:
        p.start()
    got = q.get(timeout=120)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    # single-rank truth over the whole batch
    lib = N.lib()
    ln = np.zeros(n_total, np.uint32)
    lib.shl_corpus_lengths_host(20130725, 0, n_total, 0, ln.ctypes.data)
    off = np.zeros(n_total, np.uint64)
    np.cumsum(ln[:-1], dtype=np.uint64, out=off[1:])
    data = np.zeros(int(off[-1] + ln[-1]) + 64, np.uint8)
    lib.shl_corpus_fill_host(20130725, 0, n_total, 0, off.ctypes.data, data.ctypes.data)
    ar = Arena(data, off, ln)
    po = O.orc_parse(ar)
    vals, nodes = O.orc_ring(['10.0.0.%d' % (i + 1) for i in range(16)], [11211] * 16)
    pl = O.orc_place(ar, vals, nodes, 16, (ar.off + po.sp1[:n_total] + 1).astype(np.uint64),
                     (po.sp2[:n_total] - po.sp1[:n_total] - 1).astype(np.uint32))
    assert got[:16] == pl.hist.tolist() and sum(got[:16]) == n_total
    assert got[16 + N.S_OK] == n_total


class _StubEngine(object):
    """Stands in for Engine in the rendezvous test: records what join_job hands to the library."""

    def __init__(self, q, bar):
        self.q, self.bar = q, bar

    def comm_unique_id(self):
        return bytes(range(128))

    def comm_init(self, cid, rank, world):
        self.q.put((rank, world, bytes(cid)))

    def barrier(self):
        self.bar.wait(60)        # (ncclCommInitRank and the library's barrier are collective: so is this stand-in)


def _rdzv_worker(rank, world, port, ppid_dir, q, bar):
    from shellac_b200 import dist as D
    os.environ.update({'RANK': str(rank), 'LOCAL_RANK': str(rank), 'WORLD_SIZE': str(world), 'MASTER_PORT': str(port),
                       'SHL_RDZV_DIR': ppid_dir})
    assert D.join_job(_StubEngine(q, bar)) == world


def test_two_ranks_share_the_communicator_id(tmp_path):
    import torch.multiprocessing as mp
    ctx = mp.get_context('spawn')
    q, bar = ctx.Queue(), ctx.Barrier(2)
    ps = [ctx.Process(target=_rdzv_worker, args=(r, 2, 29999, str(tmp_path), q, bar)) for r in (1, 0)]   # rank 1 first: it has to wait
    for p in ps:
        p.start()
    got = sorted(q.get(timeout=120) for _ in range(2))
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [g[:2] for g in got] == [(0, 2), (1, 2)] and got[0][2] == got[1][2] == bytes(range(128))
    assert not [f for f in os.listdir(str(tmp_path)) if f.endswith('.id')]        # rank 0 cleans up
