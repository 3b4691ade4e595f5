"""Host-side data-parallel logic on CPU with two gloo ranks: bucket plan, overlapped bucketed all-reduce equals the sum
of the per-rank gradients, identical gate on every rank, batch sharding covers the batch exactly once."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200"))


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from srb200 import dp
    n = 100_003
    g = torch.Generator().manual_seed(100 + rank)
    flat = torch.randn(n, generator=g)
    local = flat.clone()
    ar = dp.BucketedAllReduce(flat, bucket_bytes=64 * 1024)
    assert len(ar.buckets) == 7 and ar.buckets[0][1] == n and ar.buckets[-1][0] == 0
    # backward finishes the tail of the buffer first
    for lo in (90_000, 50_001, 50_000, 12):
        ar.mark_ready(lo)
    ar.finish()
    others = [torch.randn(n, generator=torch.Generator().manual_seed(100 + r)) for r in range(world)]
    torch.testing.assert_close(flat, sum(others), rtol=1e-6, atol=1e-6)
    gate = dp.sync_gate(torch.tensor([0.25 * (rank + 1)]))
    assert abs(gate.item() - 0.375) < 1e-7
    lo, hi = dp.shard_bounds(11, world, rank)
    out[rank] = (lo, hi, float(local.sum()))
    dist.destroy_process_group()


def test_bucketed_allreduce_two_ranks():
    world = 2
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_worker, args=(world, 29500 + os.getpid() % 2000, out), nprocs=world, join=True)
        bounds = [out[r][:2] for r in range(world)]
    assert bounds == [(0, 6), (6, 11)]


def test_bucket_plan_edges():
    from srb200 import dp
    assert dp.make_buckets(0, 10) == []
    assert dp.make_buckets(5, 10) == [(0, 5)]
    assert dp.make_buckets(25, 10) == [(15, 25), (5, 15), (0, 5)]
    assert [dp.shard_bounds(10, 4, r) for r in range(4)] == [(0, 3), (3, 6), (6, 8), (8, 10)]
