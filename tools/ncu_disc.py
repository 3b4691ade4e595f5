"""Two eager GAN train steps at a small batch: the ncu target for the discriminator kernels
(ncu -k regex:'dmma|dl1' --launch-skip 148 --launch-count 148 ...).  Usage: python tools/ncu_disc.py [batch]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
import bench
from oracle import oracle as O
from srb200.train_step import GANTrainStep

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dev = torch.device("cuda:0")
G, Ds = bench.build_models(dev, bench.CONFIGS['gan'])
tr = GANTrainStep(G, Ds, factor=4)
hr, lr = O.synthetic_jets(B, 80, 4, 1)
hr, lr = hr.to(dev), lr.to(dev)
for _ in range(2):
    tr.step(lr, hr)
torch.cuda.synchronize()
print("done")
