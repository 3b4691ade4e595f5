"""Kernel-time breakdown of one GAN train step (torch.profiler / CUPTI).  Usage: python tools/profile_step.py [batch]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
import bench
from oracle import oracle as O
from srb200.train_step import GANTrainStep
from torch.profiler import profile, ProfilerActivity

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dev = torch.device("cuda:0")
G, Ds = bench.build_models(dev, bench.CONFIGS['gan'])
tr = GANTrainStep(G, Ds, factor=4)
hr, lr = O.synthetic_jets(B, 80, 4, 1)
hr, lr = hr.to(dev), lr.to(dev)
for _ in range(2):
    tr.step(lr, hr)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    tr.step(lr, hr)
    torch.cuda.synchronize()
rows = [(e.key, e.device_time_total / 1e3, e.count) for e in prof.key_averages() if e.device_time_total > 0]
rows.sort(key=lambda r: -r[1])
tot = sum(r[1] for r in rows if "srb" in r[0] or "kernel" in r[0].lower())
print(f"batch {B}")
for k, ms, n in rows[:60]:
    print(f"{ms:9.3f} ms  {n:5d}x  {k[:110]}")
