"""Per-phase CUDA-event breakdown of the eager GAN train step, max over ranks (run plain for N = 1 or under torchrun).
Usage: [torchrun --nproc-per-node N ...] python tools/phase_breakdown.py [batch]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
import bench
from oracle import oracle as O

B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
pg = None
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
    pg = dist.group.WORLD
cfg = bench.CONFIGS["gan"]
G, Ds = bench.build_models(dev, cfg)
tr = bench.make_trainer(G, Ds, cfg, pg)
hr, lr = O.synthetic_jets(B, 80, 4, 1234 + rank)
hr, lr = hr.to(dev), lr.to(dev)
for _ in range(3):
    tr.step(lr, hr)
torch.cuda.synchronize()
acc = {}
R = 5
for _ in range(R):
    if world > 1:
        dist.barrier()
    tr.marks = []
    tr.step(lr, hr)
    torch.cuda.synchronize()
    m, tr.marks = tr.marks, None
    for (n0, e0), (n1, e1) in zip(m[:-1], m[1:]):
        acc[n1] = acc.get(n1, 0.0) + e0.elapsed_time(e1) / R
    acc["TOTAL"] = acc.get("TOTAL", 0.0) + m[0][1].elapsed_time(m[-1][1]) / R
names = list(acc)
t = torch.tensor([acc[n] for n in names], device=dev)
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print(f"# eager GAN step, per-GPU batch {B}, {world} GPU(s), d_streams={tr.d_streams}, NCCL_MAX_CTAS={os.environ.get('NCCL_MAX_CTAS')}: ms per phase (main stream, max over ranks, mean of {R} steps)")
    for n, v in zip(names, t.tolist()):
        print(f"  {v:8.3f} ms  {n}")
if world > 1:
    dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(0)
