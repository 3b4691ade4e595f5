"""Where does the time between kernels go?  Replays the CUDA graph of the generator-only train step (and of the full GAN
step) under torch.profiler (CUPTI kernel records), then reports busy time (union of kernel intervals), idle time, and
the idle time grouped by (previous kernel -> next kernel).  Usage: python tools/timeline_gaps.py [batch]"""
import os, sys
from collections import defaultdict
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
import bench
from oracle import oracle as O
from srb200.train_step import GeneratorTrainStep, GraphedStep
from torch.profiler import profile, ProfilerActivity

B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
dev = torch.device("cuda:0")
cfg = bench.CONFIGS["gan"]
G, Ds = bench.build_models(dev, cfg)
tr = bench.make_trainer(G, Ds, cfg, None)
hr, lr = O.synthetic_jets(B, 80, 4, 1)
hr, lr = hr.to(dev), lr.to(dev)


def short(n):
    n = n.replace("void ", "").replace("srb::", "")
    return n.split("(")[0][:40]


def analyse(tag, replay, reps=3):
    for _ in range(3):
        replay()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(reps):
            replay()
        torch.cuda.synchronize()
    ev = [(e.time_range.start, e.time_range.end, e.name) for e in prof.events()
          if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range.end > e.time_range.start]
    ev.sort()
    t0, t1 = ev[0][0], max(e[1] for e in ev)
    busy, cur_end, gaps, prev = 0.0, ev[0][0], defaultdict(lambda: [0.0, 0]), None
    for s, e, n in ev:
        if s > cur_end:
            g = gaps[(short(prev) if prev else "-", short(n))]
            g[0] += s - cur_end; g[1] += 1
            busy += e - s
            cur_end = e
        else:
            if e > cur_end:
                busy += e - cur_end
                cur_end = e
        if e >= cur_end:
            prev = n
    wall = t1 - t0
    print(f"== {tag}: {reps} replays, wall {wall / reps / 1e3:.3f} ms/step, busy {busy / reps / 1e3:.3f} ms, "
          f"idle {(wall - busy) / reps / 1e3:.3f} ms, {len(ev) // reps} kernels/step")
    for (a, b), (t, c) in sorted(gaps.items(), key=lambda kv: -kv[1][0])[:18]:
        print(f"   idle {t / reps / 1e3:7.3f} ms/step  {c // reps:4d}x  avg {t / max(c, 1):6.1f} us   {a}  ->  {b}")
    tot = defaultdict(float)
    for s, e, n in ev:
        tot[short(n)] += e - s
    print("   kernel time: " + ", ".join(f"{k} {v / reps / 1e3:.2f}" for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:10]))


gts = GeneratorTrainStep(G)
gts.opt_G = tr.opt_G
side = torch.cuda.Stream()
side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    for _ in range(2):
        gts.step(lr, hr)
torch.cuda.current_stream().wait_stream(side)
torch.cuda.synchronize()
gg = torch.cuda.CUDAGraph()
with torch.cuda.graph(gg):
    gts.step(lr, hr)
analyse("generator-only train step (CUDA graph)", gg.replay)
G.eval()
with torch.no_grad():
    for _ in range(2):
        G(lr)
    torch.cuda.synchronize()
    gi = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gi):
        G(lr)
analyse("inference (CUDA graph)", gi.replay)
with torch.no_grad():
    analyse("inference (eager)", lambda: G(lr))
G.train()
gs = GraphedStep(tr, lr, hr, warmup=2)
analyse("full GAN step (CUDA graph)", lambda: gs.step(lr, hr))
