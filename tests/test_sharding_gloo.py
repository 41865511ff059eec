"""CPU tests of the multi-GPU host logic (world_size 2 and 3, gloo backend): partitioning, the final statistics
reduce and the result gather in global sequence order."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cheetah_pose_estimation_b200 import sharding


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_seq, ragged, q_out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        work = (np.arange(n_seq) % 5 + 1) * 30.0 if ragged else None
        sh = sharding.my_shard(n_seq, work)
        # per-sequence "results" that are a pure function of the global index: must not depend on the rank count
        gidx = torch.as_tensor(sh.index, dtype=torch.float64)
        cost = torch.stack([gidx + 1.0, 2.0 * gidx, 0.5 * gidx, torch.ones_like(gidx)], dim=1)
        counters = torch.stack([gidx.to(torch.int32) + 3, gidx.to(torch.int32)], dim=1)
        q = gidx[:, None, None] + torch.zeros(len(sh.index), 4, 54, dtype=torch.float64)
        stats = sharding.reduce_stats(sharding.local_stats(cost, counters, converged=torch.ones(len(sh.index))))
        tmax = sharding.reduce_max(10.0 + rank, torch.device("cpu"))
        out = sharding.gather_results(sh, n_seq, {"q": q, "cost": cost, "status": counters[:, 1].contiguous()})
        everywhere = sharding.gather_results(sh, n_seq, {"q": q}, dst=None)       # the run_dataset driver's form
        assert torch.equal(everywhere["q"][:, 0, 0], torch.arange(n_seq, dtype=torch.float64))
        if rank == 0:
            q_out.put((stats, tmax, out["q"][:, 0, 0].numpy().copy(), out["cost"].numpy().copy(),
                       out["status"].numpy().copy(), len(sh.index)))
        else:
            assert out is None
            q_out.put(("n_local", len(sh.index)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_seq,ragged", [(2, 10, False), (2, 7, True), (3, 8, False)])
def test_partition_reduce_gather(world, n_seq, ragged):
    ctx = mp.get_context("spawn")
    q_out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_seq, ragged, q_out)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q_out.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    main = next(g for g in got if len(g) == 6)
    stats, tmax, qg, cost, status, _ = main
    g = np.arange(n_seq, dtype=np.float64)
    assert np.array_equal(qg, g)                                   # global order restored
    assert np.array_equal(cost[:, 0], g + 1) and np.array_equal(status, g.astype(np.int32))
    assert stats["n_sequences"] == n_seq and stats["n_converged"] == n_seq
    assert stats["cost_meas"] == float(np.sum(g + 1)) and stats["iterations"] == float(np.sum(g + 3))
    assert tmax == 10.0 + world - 1                                # max over ranks


def test_partitions_cover_every_sequence_once():
    for n, w in ((4096, 8), (10, 3), (5, 8)):
        parts = sharding.partition_contiguous(n, w)
        assert parts[0][0] == 0 and parts[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
        assert max(b - a for a, b in parts) - min(b - a for a, b in parts) <= 1
    work = np.array([57 * 6, 30 * 6, 45 * 4, 33 * 6, 40 * 5, 52 * 6, 31 * 4, 36 * 6, 48 * 6, 30 * 4], float)
    bins = sharding.partition_balanced(work, 4)
    assert sorted(np.concatenate(bins).tolist()) == list(range(10))
    loads = [work[b].sum() for b in bins]
    assert max(loads) - min(loads) <= work.max()
