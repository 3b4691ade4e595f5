#!/usr/bin/env python
"""The bar SURVEY.md section 2a sets: the REFERENCE path under stock PyTorch / cuDNN on the same B200.

The reference is pure torch.nn (models.py) + the tensor math of esrgan.py:399-626; on a CUDA device every convolution
dispatches to cuDNN's sm_100 kernels.  This tool times exactly that - the reference's own modules from baseline/_ref
when the copy is present, else the functional restatement in oracle/oracle.py (same ATen calls) - for

    precision  fp32 eager (stock defaults: cuDNN convolutions may use TF32)
               bf16 autocast + channels_last
    launch     eager, and one CUDA graph per step
    workload   generator inference, generator warm-up train step (esrgan.py:418-427), full GAN step (esrgan.py:457-626
               default terms, 2 patch discriminators, gradient penalty, 3 Adam)
    batch      16 / 64 / 512

on synthetic sparse jet images, weights from the same seeds bench.py uses.  It is a BASELINE leg: nothing of
super-resolution_b200/ is on this path.  Results: one JSON object (also used by bench.py's `gpu_baseline` block).

    python tools/gpu_baseline.py --out gpurun_out/gpu_baseline.json [--quick]
"""
import argparse
import contextlib
import importlib.util
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from oracle import oracle as O  # noqa: E402

NB, NUP, HR, FACTOR, RES_SCALE = 10, 2, 80, 4, 0.1


def _ref_models():
    p = os.path.join(ROOT, "baseline", "_ref", "models.py")
    if not os.path.exists(p):
        return None
    spec = importlib.util.spec_from_file_location("ref_models_baseline", p)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


class _FunctionalG(torch.nn.Module):
    """oracle.generator_forward as a module: the same ATen calls as the reference's GeneratorRRDB.forward, but `power` /
    `multiplier` stay host scalars, so the forward has none of the reference's .item() synchronisations
    (models.py:116,118,133) and CAN be captured in a CUDA graph.  Used for the cuda_graph rows (the reference modules
    themselves cannot be captured) and when baseline/_ref is absent."""

    def __init__(self, sd, nb, nup):
        super().__init__()
        self.names = [k for k in sd if k not in ("power", "multiplier")]
        self.ps = torch.nn.ParameterList([torch.nn.Parameter(sd[k].clone()) for k in self.names])
        self.host = {k: sd[k].clone() for k in ("power", "multiplier")}      # CPU tensors: .item() costs no device sync
        self.thres = 0.0

    def forward(self, x):
        sd = dict(zip(self.names, self.ps))
        sd.update(self.host)
        out, self.srs = O.generator_forward(sd, x, res_scale=RES_SCALE, thres=0.0, training=self.training)
        return out


def build(dev, nb=NB, nup=NUP, channels_last=False, sync_free=False):
    R = _ref_models()
    sd = O.make_generator_sd(1234, num_res_blocks=nb, num_upsample=nup)
    Ds = None
    if R is not None:
        Ds = []
        for k in range(2):
            d = R.Markovian_Discriminator((1, HR, HR), [16, 32, 32, 64])
            d.load_state_dict(O.make_discriminator_sd(1235 + k, gain=2.0))
            Ds.append(d)
    if R is not None and not sync_free:
        G = R.GeneratorRRDB(1, filters=64, num_res_blocks=nb, num_upsample=nup, res_scale=RES_SCALE)
        G.load_state_dict(sd)
        kind = "reference modules (baseline/_ref/models.py)"
    else:
        G = _FunctionalG(sd, nb, nup)
        kind = "functional port of the generator without the reference's .item() syncs (oracle/oracle.py)" + \
               ("" if R is None else " + reference discriminator modules")
    G = G.to(dev)
    if Ds is not None:
        Ds = [d.to(dev) for d in Ds]
    if channels_last:
        G = G.to(memory_format=torch.channels_last)
        if Ds is not None:
            Ds = [d.to(memory_format=torch.channels_last) for d in Ds]
    return G, Ds, kind


def _bce(x, t):
    return F.binary_cross_entropy_with_logits(x, torch.full_like(x, t))


def gan_step_factory(G, Ds, hr, lr, ctx, capturable, gate=True):
    """esrgan.py:457-626 default terms, written with the reference's modules (same ops as oracle.generator_loss)."""
    B = hr.shape[0]
    opt_g = torch.optim.Adam([p for p in G.parameters() if p.requires_grad], lr=2e-4, betas=(0.9, 0.999),
                             capturable=capturable)
    opt_d = [torch.optim.Adam(d.parameters(), lr=2e-4, betas=(0.9, 0.999), capturable=capturable) for d in Ds]
    eps = 1e-7
    lam = (0.2, 1.0)

    def step():
        opt_g.zero_grad(set_to_none=True)
        with ctx():
            out = G(lr)
            gen = (out, G.srs)
            gen_lr = O.sum_pool(out.float(), FACTOR)
            loss_g = 0.0
            for k in range(2):
                l_hr = F.l1_loss(gen[k].float().mean(0)[None], hr.mean(0)[None])
                l_lr = F.l1_loss(gen_lr, lr)
                pr = Ds[k](hr).detach()
                pf = Ds[k](gen[k])
                l_adv = 0.5 * (_bce(eps + pf - pr.mean(0, keepdim=True), 1.0) + _bce(eps + pr - pf.mean(0, keepdim=True), 0.0))
                loss_g = loss_g + lam[k] * (l_hr + 0.01 * l_adv + 0.1 * l_lr)
        loss_g.backward()
        opt_g.step()
        for k in range(2):
            opt_d[k].zero_grad(set_to_none=True)
            fake = gen[k].detach().float()
            with ctx():
                pr, pf = Ds[k](hr), Ds[k](fake)
                loss_d = (_bce(eps + pr - pf.mean(0, keepdim=True), 1.0) + _bce(eps + pf - pr.mean(0, keepdim=True), 0.0)) / 2
                e = torch.rand(B, 1, 1, 1, device=hr.device)
                interp = (e * hr + (1 - e) * fake).requires_grad_(True)
                pi = Ds[k](interp)
                gr = torch.autograd.grad(pi, interp, torch.ones_like(pi), create_graph=True)[0].reshape(B, -1)
                loss_d = loss_d + ((gr.float().norm(2, dim=1) - 1) ** 2).mean() * 0.01 / 2
            loss_d.backward()
            if not gate or loss_d.item() > 0.001:          # esrgan.py:623 (host sync; dropped under a CUDA graph)
                opt_d[k].step()
        return loss_g

    return step


def warmup_step_factory(G, hr, lr, ctx, capturable):
    opt_g = torch.optim.Adam([p for p in G.parameters() if p.requires_grad], lr=2e-4, betas=(0.9, 0.999),
                             capturable=capturable)

    def step():
        opt_g.zero_grad(set_to_none=True)
        with ctx():
            loss = F.l1_loss(G(lr).float(), hr)
        loss.backward()
        opt_g.step()
        return loss

    return step


def timed(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def graphed(fn, warmup=3):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(warmup):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    return g.replay


def measure(dev, batch, precision, workload, use_graph, steps=5, warmup=3, nb=NB, nup=NUP, factor=FACTOR):
    """images/s of one (batch, precision, workload, launch) point; None + reason when it cannot run."""
    cl = precision == "bf16_channels_last"
    G, Ds, kind = build(dev, nb, nup, channels_last=cl, sync_free=use_graph)
    hr, lr = O.synthetic_jets(batch, HR, factor, 1234)
    hr, lr = hr.to(dev), lr.to(dev)
    if cl:
        hr, lr = hr.contiguous(memory_format=torch.channels_last), lr.contiguous(memory_format=torch.channels_last)
    ctx = (lambda: torch.autocast("cuda", dtype=torch.bfloat16)) if cl else contextlib.nullcontext
    try:
        if workload == "inference":
            G.eval()

            def fn():
                with torch.no_grad(), ctx():
                    return G(lr)
        elif workload == "generator_train":
            G.train()
            fn = warmup_step_factory(G, hr, lr, ctx, capturable=use_graph)
        else:
            if Ds is None:
                return None, "needs baseline/_ref"
            G.train()
            fn = gan_step_factory(G, Ds, hr, lr, ctx, capturable=use_graph, gate=not use_graph)
        if use_graph:
            fn = graphed(fn, warmup=2)
        ms = timed(fn, steps, 0 if use_graph else warmup)
        return {"images_per_s": batch / (ms * 1e-3), "ms_per_step": ms, "impl": kind}, None
    except Exception as e:  # noqa: BLE001   (OOM / capture failure: reported, not fatal)
        return None, f"{type(e).__name__}: {str(e)[:200]}"
    finally:
        del G, Ds
        torch.cuda.empty_cache()


def run_point(dev, wl, prec, use_graph, b, steps):
    t0 = time.time()
    res, err = measure(dev, b, prec, wl, use_graph, steps=steps)
    row = {"workload": wl, "precision": prec, "launch": "cuda_graph" if use_graph else "eager", "batch": b,
           "wall_s": round(time.time() - t0, 1)}
    row.update(res if res is not None else {"unavailable": err})
    return row


def run_matrix(dev, batches, precisions, workloads, launches, steps=5):
    """Eager points run in this process.  Every cuda_graph point runs in its OWN subprocess: a capture that fails (or
    runs out of memory) leaves the RNG / allocator of its process unusable for the points after it."""
    import subprocess
    rows = []
    for wl in workloads:
        for prec in precisions:
            for use_graph in launches:
                for b in batches:
                    if not use_graph:
                        row = run_point(dev, wl, prec, False, b, steps)
                    else:
                        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--point",
                                            f"{wl},{prec},{b},{steps}"], capture_output=True, text=True, timeout=600)
                        lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
                        row = json.loads(lines[-1]) if lines else {
                            "workload": wl, "precision": prec, "launch": "cuda_graph", "batch": b,
                            "unavailable": "subprocess failed: " + (r.stderr.strip().splitlines() or ["?"])[-1][:200]}
                    rows.append(row)
                    print(json.dumps(row), flush=True)
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--quick", action="store_true", help="batch 512 only, eager + graph, 3 steps")
    ap.add_argument("--point", default=None, help="internal: workload,precision,batch,steps of one cuda_graph point")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    if a.point:
        wl, prec, b, steps = a.point.split(",")
        print(json.dumps(run_point(dev, wl, prec, True, int(b), int(steps))), flush=True)
        return
    batches = [512] if a.quick else [16, 64, 512]
    rows = run_matrix(dev, batches, ["fp32", "bf16_channels_last"], ["inference", "generator_train", "gan_step"],
                      [False, True], steps=3 if a.quick else 5)
    res = {"torch": torch.__version__, "cudnn": torch.backends.cudnn.version(),
           "cudnn_allow_tf32": bool(torch.backends.cudnn.allow_tf32), "gpu": torch.cuda.get_device_name(0),
           "config": f"GeneratorRRDB nb={NB} F=64 x{FACTOR} (LR {HR // FACTOR} -> HR {HR}), 2 patch discriminators",
           "rows": rows}
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
