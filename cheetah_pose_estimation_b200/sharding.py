"""Multi-GPU plumbing: sequences (independent NLPs) are partitioned over ranks, one process per GPU; the inner
loop has NO inter-GPU traffic.  The only collectives are the final statistics reduce and the result gather
(SURVEY.md section 8e).  Works on any torch.distributed backend (nccl on the GPUs, gloo in the CPU tests).

The reference runs its clips strictly sequentially (run_dataset.py:656-658, 741-743, 1143-1231); its loop over
clips / monocular camera choices is the axis that is sharded here.
"""
from dataclasses import dataclass

import numpy as np
import torch
import torch.distributed as dist

STAT_KEYS = ("cost_meas", "cost_model", "cost_eq", "cost_in", "iterations", "n_converged", "n_failed", "n_sequences")


def partition_contiguous(n_sequences, world_size):
    """[(start, stop)] per rank: contiguous blocks whose sizes differ by at most one (equal-length sequences)."""
    base, rem = divmod(n_sequences, world_size)
    out, s = [], 0
    for r in range(world_size):
        n = base + (1 if r < rem else 0)
        out.append((s, s + n))
        s += n
    return out


def partition_balanced(work, world_size):
    """Ragged clips: sort by work (frames x cameras) and deal greedily to the least-loaded rank.
    Returns a list of index arrays (one per rank, each sorted ascending).  Deterministic."""
    work = np.asarray(work, dtype=np.float64)
    order = np.argsort(-work, kind="stable")
    load = np.zeros(world_size)
    bins = [[] for _ in range(world_size)]
    for i in order:
        r = int(np.argmin(load))          # ties -> lowest rank
        bins[r].append(int(i))
        load[r] += work[i]
    return [np.array(sorted(b), dtype=np.int64) for b in bins]


@dataclass
class ShardInfo:
    rank: int
    world_size: int
    index: np.ndarray          # global sequence indices owned by this rank


def my_shard(n_sequences, work=None, rank=None, world_size=None):
    if rank is None:
        rank = dist.get_rank() if dist.is_initialized() else 0
    if world_size is None:
        world_size = dist.get_world_size() if dist.is_initialized() else 1
    if work is None:
        a, b = partition_contiguous(n_sequences, world_size)[rank]
        idx = np.arange(a, b, dtype=np.int64)
    else:
        idx = partition_balanced(work, world_size)[rank]
    return ShardInfo(rank, world_size, idx)


def local_stats(cost, counters=None, converged=None, failed=None):
    """FP64 statistics vector of this rank's sequences, in STAT_KEYS order.  cost[S,4]."""
    dev = cost.device
    v = torch.zeros(len(STAT_KEYS), dtype=torch.float64, device=dev)
    v[0:4] = cost.to(torch.float64).sum(dim=0)
    if counters is not None:
        v[4] = counters[:, 0].to(torch.float64).sum()
    if converged is not None:
        v[5] = converged.to(torch.float64).sum()
    if failed is not None:
        v[6] = failed.to(torch.float64).sum()
    v[7] = float(cost.shape[0])
    return v


def reduce_stats(stats, group=None):
    """Sum-all-reduce of the statistics vector (the path's one reduce).  Returns a dict on every rank."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return {k: float(x) for k, x in zip(STAT_KEYS, stats.tolist())}


def reduce_max(value, device, group=None):
    """Max over ranks of a device-timed duration (ms)."""
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def gather_results(shard: ShardInfo, n_sequences, tensors, dst=0, group=None):
    """Gather per-sequence results to rank `dst` (None = to every rank) in GLOBAL sequence order.

    tensors: dict name -> tensor [S_local, ...] (same trailing shape and dtype on every rank).  Shards may differ in
    size, so each rank's block is padded to the largest shard for the all_gather and trimmed afterwards.
    Returns dict name -> tensor [n_sequences, ...] on `dst` (on every rank when dst is None), None elsewhere."""
    world = shard.world_size
    if world == 1 or not dist.is_initialized():
        out = {}
        for k, t in tensors.items():
            full = torch.empty((n_sequences,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
            full[torch.as_tensor(shard.index, device=t.device)] = t
            out[k] = full
        return out
    any_t = next(iter(tensors.values()))
    dev = any_t.device
    n_loc = torch.tensor([len(shard.index)], dtype=torch.int64, device=dev)
    sizes = [torch.zeros_like(n_loc) for _ in range(world)]
    dist.all_gather(sizes, n_loc, group=group)
    sizes = [int(s.item()) for s in sizes]
    n_max = max(sizes)
    idx_pad = torch.full((n_max,), -1, dtype=torch.int64, device=dev)
    idx_pad[:len(shard.index)] = torch.as_tensor(shard.index, device=dev)
    idx_all = [torch.empty_like(idx_pad) for _ in range(world)]
    dist.all_gather(idx_all, idx_pad, group=group)
    keep = dst is None or shard.rank == dst
    out = {} if keep else None
    for k, t in tensors.items():
        pad = torch.zeros((n_max,) + tuple(t.shape[1:]), dtype=t.dtype, device=dev)
        pad[:t.shape[0]] = t
        parts = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(parts, pad, group=group)
        if keep:
            full = torch.empty((n_sequences,) + tuple(t.shape[1:]), dtype=t.dtype, device=dev)
            for r in range(world):
                full[idx_all[r][:sizes[r]]] = parts[r][:sizes[r]]
            out[k] = full
    return out
