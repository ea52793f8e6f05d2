"""Multi-GPU plumbing: the path shards by request batch, one process per GPU,
no inter-GPU movement of payloads.  The only exchange is a sum all-reduce of the per-node placement
histogram and the per-outcome counters, `int64[n_hosts + SHL_N_CODES]`, once per batch.

torch.distributed is used for exactly that (NCCL on GPUs, gloo in the CPU tests); nothing else in
the package imports torch.
"""
import numpy as np

from . import _native as N


def shard_range(n_total, rank, world):
    """Contiguous slice [lo, hi) of a batch of n_total requests owned by `rank`."""
    per = (n_total + world - 1) // world
    lo = min(n_total, rank * per)
    return lo, min(n_total, lo + per)


def pack_stats(hist, codes):
    """[placement histogram | outcome counters] as one int64 vector (the all-reduce payload)."""
    out = np.zeros(len(hist) + N.N_CODES, np.int64)
    out[:len(hist)] = hist
    out[len(hist):len(hist) + len(codes)] = codes
    return out


def allreduce_stats(stats, group=None):
    """Sum-all-reduce a stats tensor in place.  `stats` is a torch tensor (CUDA for NCCL, CPU for
    gloo) or a numpy int64 array (reduced through a CPU tensor)."""
    import torch
    import torch.distributed as dist
    if isinstance(stats, np.ndarray):
        t = torch.from_numpy(stats)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        return stats
    dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return stats


def external_stream(engine, device_index):
    """The engine's CUDA stream as a torch stream, so a collective can be queued right behind the
    kernel that produced its input without a host synchronisation."""
    import torch
    return torch.cuda.ExternalStream(engine.lib.shl_stream(engine.ctx), device='cuda:%d' % device_index)
