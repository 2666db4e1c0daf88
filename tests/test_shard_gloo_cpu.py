"""Multi-GPU host logic on CPU: worlds are sharded by index with no data-path collective
(SURVEY.md §8e); the only cross-rank traffic is the benchmark's barrier + max-over-ranks
timing.  Two gloo ranks check the shard plan (disjoint, covering, RNG keyed by global world
index) and the aggregation bench.py uses."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, total_worlds, q):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "marl-hideandseek_b200"))
    sys.path.insert(0, root)
    import gpu_hideseek as gh
    import bench
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    assert bench.dist_env() == (rank, world, rank)
    base, n = gh.shard_worlds(total_worlds, world)[rank]
    # every rank learns every shard: ranges must tile [0, total) exactly
    mine = torch.tensor([base, n], dtype=torch.int64)
    all_ = [torch.zeros(2, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(all_, mine)
    cover = []
    for b, c in (t.tolist() for t in all_):
        cover.extend(range(b, b + c))
    assert cover == list(range(total_worlds))
    # weak-scaling aggregate: value = total worlds * steps / max-over-ranks time
    ms = torch.tensor([10.0 + 5.0 * rank], dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    value = bench.aggregate_value(n_per_rank=n, world=world, steps=7, ms_total=float(ms.item()))
    if rank == 0:
        q.put((float(ms.item()), value))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shard_plan_and_aggregation():
    world, total = 2, 32000
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    ms, value = q.get()
    assert ms == 15.0
    assert value == pytest.approx(16000 * 2 * 7 / 0.015)


def test_shard_plan_properties():
    import gpu_hideseek as gh
    for total, shards in ((16000, 1), (128000, 8), (10, 4), (3, 8)):
        plan = gh.shard_worlds(total, shards)
        assert len(plan) == shards
        assert sum(n for _, n in plan) == total
        nxt = 0
        for base, n in plan:
            assert base == nxt and n >= 0
            nxt += n
    with pytest.raises(ValueError):
        gh.shard_worlds(10, 0)
