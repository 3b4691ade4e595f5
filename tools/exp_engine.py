"""A/B timing of the generator engine on one box: per-launch CUDA-event times of the trunk's chained forward launches,
chained input-gradient launches and grouped weight-gradient launches (ops.PROFILE), plus inference images/s, for
engine switches given as env-style pairs.  Usage: python tools/exp_engine.py [batch]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
import bench
from oracle import oracle as O
from srb200 import ops
from srb200.train_step import GeneratorTrainStep

B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
dev = torch.device("cuda:0")
cfg = bench.CONFIGS["gan"]
hr, lr = O.synthetic_jets(B, 80, 4, 1)
hr, lr = hr.to(dev), lr.to(dev)


def run(tag, **switches):
    G, _ = bench.build_models(dev, cfg)
    eng = G._engine()
    for k, v in switches.items():
        setattr(eng, k, v)
    gts = GeneratorTrainStep(G)
    for _ in range(3):
        gts.step(lr, hr)
    torch.cuda.synchronize()
    acc = {}
    for rep in range(3):
        ops.PROFILE = []
        gts.step(lr, hr)
        torch.cuda.synchronize()
        prof, ops.PROFILE = ops.PROFILE, None
        convs = [(fl, e0.elapsed_time(e1)) for kind, fl, e0, e1 in prof if kind == "conv3x3"]
        wg = [(fl, e0.elapsed_time(e1)) for kind, fl, e0, e1 in prof if kind == "wgrad3x3"]
        chain_fl = max(fl for fl, _ in convs[:30])
        fwd = [t for fl, t in convs[:30]]
        rest = convs[34:] if len(convs) > 64 else convs[30:]
        dg = [t for fl, t in rest if abs(fl - chain_fl) < 1e-3 * chain_fl]
        other = [t for fl, t in convs[30:] if abs(fl - chain_fl) >= 1e-3 * chain_fl]
        wg_grp_fl = max(fl for fl, _ in wg)
        wgg = [t for fl, t in wg if abs(fl - wg_grp_fl) < 1e-3 * wg_grp_fl]
        for k, v in (("fwd_chain_us", 1e3 * sum(fwd) / len(fwd)), ("dgrad_chain_us", 1e3 * sum(dg) / max(len(dg), 1)),
                     ("wgrad_group_us", 1e3 * sum(wgg) / len(wgg)), ("other_conv_ms", sum(other)),
                     ("conv_total_ms", sum(t for _, t in convs)), ("wgrad_total_ms", sum(t for _, t in wg))):
            acc.setdefault(k, []).append(v)
    res = {k: sorted(v)[len(v) // 2] for k, v in acc.items()}
    res["fwd_chain_TF"] = chain_fl / (res["fwd_chain_us"] * 1e-6) / 1e12
    res["dgrad_chain_TF"] = chain_fl / (res["dgrad_chain_us"] * 1e-6) / 1e12
    res["wgrad_group_TF"] = wg_grp_fl / (res["wgrad_group_us"] * 1e-6) / 1e12
    G.eval()
    with torch.no_grad():
        ms = bench.timed(lambda: G(lr), 10, 3, lambda: None)
    res["inference_img_s"] = B / (ms * 1e-3)
    ms = bench.timed(lambda: gts.step(lr, hr), 5, 2, lambda: None)
    res["g_train_eager_img_s"] = B / (ms * 1e-3)
    print(tag, {k: round(v, 2) for k, v in res.items()}, flush=True)
    del G, gts
    torch.cuda.empty_cache()


run("fp32_stream=1 (default)", fp32_stream=True)
run("fp32_stream=0          ", fp32_stream=False)
run("fp32_stream=1 again    ", fp32_stream=True)
print("fault", ops.read_fault())
