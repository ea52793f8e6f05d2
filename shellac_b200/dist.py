"""Multi-GPU plumbing: the path shards by request batch, one process per GPU,
no inter-GPU movement of payloads.  The only exchange is a sum all-reduce of the per-node placement
histogram and the per-outcome counters, `uint64[n_hosts + SHL_N_CODES]`, once per batch; the library
does it itself (shl_stats_allreduce: NCCL on a side stream, see include/shellac_b200.h).  This
module holds what is left for the host: who owns which requests, and how the ranks of a job find
each other.  No PyTorch: the launcher's environment (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_PORT,
as `python -m torch.distributed.run` sets them) and a file are enough.
"""
import os
import time

import numpy as np

from . import _native as N


def shard_range(n_total, rank, world):
    """Contiguous slice [lo, hi) of a batch of n_total requests owned by `rank`."""
    per = (n_total + world - 1) // world
    lo = min(n_total, rank * per)
    return lo, min(n_total, lo + per)


def pack_stats(hist, codes):
    """[placement histogram | outcome counters] as one vector (the all-reduce payload)."""
    out = np.zeros(len(hist) + N.N_CODES, np.int64)
    out[:len(hist)] = hist
    out[len(hist):len(hist) + len(codes)] = codes
    return out


def env_rank():
    """(rank, local_rank, world) from the launcher's environment (1 process: 0, 0, 1)."""
    return (int(os.environ.get('RANK', '0')), int(os.environ.get('LOCAL_RANK', '0')),
            int(os.environ.get('WORLD_SIZE', '1')))


def _id_path():
    # one launcher (torchrun agent) is the parent of every rank: port + its pid name the job
    return os.path.join(os.environ.get('SHL_RDZV_DIR', '/tmp'),
                        'shl_comm_%s_%d.id' % (os.environ.get('MASTER_PORT', '0'), os.getppid()))


def join_job(engine, rank=None, world=None, timeout=180.0):
    """Give `engine` the job's NCCL communicator.  Rank 0 asks the library for a unique id and
    publishes it in a file (written whole, then renamed); the others wait for the file.  Returns
    the world size.  A single-rank job needs no communicator."""
    r, _, w = env_rank()
    rank = r if rank is None else rank
    world = w if world is None else world
    if world <= 1:
        return 1
    path = _id_path()
    if rank == 0:
        cid = engine.comm_unique_id()
        with open(path + '.tmp', 'wb') as f:
            f.write(cid)
        os.replace(path + '.tmp', path)
    else:
        t0 = time.time()
        while not os.path.exists(path):
            if time.time() - t0 > timeout:
                raise N.ShellacError('rank %d: no communicator id at %s after %.0f s' % (rank, path, timeout))
            time.sleep(0.01)
        with open(path, 'rb') as f:
            cid = f.read()
    engine.comm_init(cid, rank, world)
    engine.barrier()
    if rank == 0:
        try:
            os.remove(path)
        except OSError:
            pass
    return world
