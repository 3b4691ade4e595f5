"""ncu target: eager GAN train steps at the bench configuration (nb=10, F=64, x4, per-GPU batch 512 by default).
Two warm-up steps, then ONE step between cudaProfilerStart/Stop (use ncu --profile-from-start off to restrict the
capture to it; without that flag every launch is visible and -k / -s / -c select).
Usage: python tools/ncu_step.py [batch] [steps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
import bench
from oracle import oracle as O

B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
dev = torch.device("cuda:0")
cfg = bench.CONFIGS["gan"]
G, Ds = bench.build_models(dev, cfg)
tr = bench.make_trainer(G, Ds, cfg, None)
hr, lr = O.synthetic_jets(B, 80, 4, 1)
hr, lr = hr.to(dev), lr.to(dev)
for _ in range(2):
    tr.step(lr, hr)
torch.cuda.synchronize()
torch.cuda.profiler.start()
for _ in range(steps):
    tr.step(lr, hr)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("done")
