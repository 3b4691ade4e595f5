#!/usr/bin/env python
"""Benchmark of the B200-native ESRGAN hot path (BASELINE.json metric: ESRGAN train images/s).

  python bench.py --gpus N --steps K --warmup W            our arm   (torchrun for N > 1, one rank per GPU, NCCL)
  python bench.py --impl reference --gpus N ...            reference arm: the CPU oracle (oracle/oracle.py, a
                                                           restatement of the reference's own torch-CPU path) on the
                                                           host cores; rank 0 only.
  python bench.py --config deep  ...                       BASELINE.json configs[3]: nb = 23, F = 64, x8
  python bench.py --config sweep ...                       BASELINE.json configs[4]: hyper_search shapes, timed per point

A "step" is one full GAN training iteration (esrgan.py:399-626 default path: generator fwd/bwd + Adam, two patch
discriminators with relativistic loss + gradient penalty + gated Adam) on a batch of synthetic sparse jet images
(options/default.json generator: 10 RRDB blocks, 64 filters; factor 4: LR 20x20 -> HR 80x80).  Weak scaling: the
per-GPU batch is fixed, gradients are all-reduced over NCCL.  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200"))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

HR = 80
RES_SCALE = 0.1                      # options/default.json "res_scale"
GFLOP_D_FWD_PER_IMG = 0.0461
# name -> (num_res_blocks, filters, factor, default per-GPU batch, workload text)
CONFIGS = {
    "gan": dict(nb=10, F=64, factor=4, batch=512,
                text="full ESRGAN GAN train step, GeneratorRRDB nb=10 F=64 x4 (LR 20x20 -> HR 80x80), "
                     "2 patch discriminators, RaGAN + gradient penalty, 3 Adam"),
    # per-GPU batch 2496: 2496 * 11 * 11 padded LR rows = 590 tiles of 512 rows = 3.99 waves of 148 CTAs
    "deep": dict(nb=23, F=64, factor=8, batch=2496,
                 text="full ESRGAN GAN train step, deep GeneratorRRDB nb=23 F=64 x8 (LR 10x10 -> HR 80x80), "
                      "2 patch discriminators, RaGAN + gradient penalty, 3 Adam"),
}


def gflop_fwd_per_img(nb, F, factor, ch=1, hr=HR):
    """BASELINE.md section 3 / SURVEY.md section 8d: algorithmic forward FLOPs of GeneratorRRDB per image."""
    U = factor.bit_length() - 1
    hw = (hr // factor) ** 2
    mac = hw * (9 * ch * F + nb * 3 * 135 * F * F + 9 * F * F)
    mac += sum(4 ** i * hw * 36 * F * F for i in range(U))
    mac += 4 ** U * hw * (9 * F * F + 9 * F * ch)
    return 2.0 * mac / 1e9


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="gan", choices=["gan", "deep", "sweep"])
    ap.add_argument("--batch", type=int, default=int(os.environ.get("SRB_BENCH_BATCH", 0)), help="per-GPU batch")
    ap.add_argument("--cpu-batch", type=int, default=16)
    ap.add_argument("--no-graph", action="store_true", help="launch the step eagerly instead of replaying one CUDA graph")
    ap.add_argument("--no-extras", action="store_true", help="skip inference / generator-only / cpu_baseline legs")
    ap.add_argument("--no-gpu-baseline", action="store_true", help="skip the stock PyTorch/cuDNN leg")
    ap.add_argument("--no-parity-check", action="store_true", help="skip the step-0 loss check against the oracle")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------- helpers
def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(bf16_burst=d.get("bf16_tflops"), bf16_sustained=d.get("bf16_tflops_sustained"),
                    hbm=d.get("hbm_gbs"), source="measured (MEASURED_PEAKS.json)")
    return dict(bf16_burst=1590.0, bf16_sustained=1400.0, hbm=6650.0, source="fallback (B200_PROFILING.md)")


def traffic_from_profile():
    """dram__bytes_read.sum + dram__bytes_write.sum per chained launch, read from the kept ncu summary (never a
    constant in this file).  The newest profiles/r*_ncu_full_conv3x3_chain*.txt that carries both counters wins."""
    import glob
    import re
    best = None
    def order(f):       # r<round>_s<session><letter>_...: numeric order (s17 after s9c)
        m = re.match(r"r(\d+)_s(\d+)([a-z]*)_", os.path.basename(f))
        return (int(m.group(1)), int(m.group(2)), m.group(3)) if m else (0, 0, "")
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_full_conv3x3_chain*.txt")), key=order):
        txt = open(f).read()
        m = re.search(r"traffic_bytes_per_launch\s*[:=]\s*([0-9.eE+]+)", txt)
        if m:
            best = (float(m.group(1)), os.path.relpath(f, ROOT))
            continue
        unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        vals = {}
        for key in ("read", "write"):       # "<name> [unit]: v | v ..." (r01 files) or "<name>   v unit" (ncu raw page rows)
            a = re.search(r"dram__bytes_%s\.sum\s*\[(\w+)\]\s*:\s*([0-9.,]+)" % key, txt)
            b = re.search(r"dram__bytes_%s\.sum\s+([0-9.,]+)\s+(\w+)" % key, txt)
            if a:
                vals[key] = float(a.group(2).replace(",", "")) * unit.get(a.group(1), 1.0)
            elif b:
                vals[key] = float(b.group(1).replace(",", "")) * unit.get(b.group(2), 1.0)
        if len(vals) == 2:
            best = (vals["read"] + vals["write"], os.path.relpath(f, ROOT))
    return best


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        pw = sorted(float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace(".", "", 1).isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i] == "Active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm), "power_w": pw[len(pw) // 2] if pw else None}


def build_models(dev, cfg):
    import models
    from oracle import oracle as O
    nup = cfg["factor"].bit_length() - 1
    G = models.GeneratorRRDB(1, filters=cfg["F"], num_res_blocks=cfg["nb"], num_upsample=nup, res_scale=RES_SCALE)
    G.load_state_dict(O.make_generator_sd(1234, filters=cfg["F"], num_res_blocks=cfg["nb"], num_upsample=nup))
    Ds = []
    for k in range(2):
        d = models.Markovian_Discriminator((1, HR, HR), [16, 32, 32, 64])
        d.load_state_dict(O.make_discriminator_sd(1235 + k, gain=2.0))
        Ds.append(d.to(dev))
    return G.to(dev).train(), Ds


def make_trainer(G, Ds, cfg, pg):
    from srb200.train_step import GANTrainStep
    return GANTrainStep(G, Ds, factor=cfg["factor"], lambdas=(0.2, 1.0), lambda_hr=1.0, lambda_lr=0.1, lambda_adv=0.01,
                        lambda_reg=0.01, d_threshold=0.001, lr=2e-4, process_group=pg)


def timed(fn, steps, warmup, barrier):
    """CUDA-event timing of `steps` calls after `warmup`, barrier + synchronize on both sides. Returns ms/step."""
    for _ in range(warmup):
        fn()
    barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    barrier()
    return e0.elapsed_time(e1) / steps


# ----------------------------------------------------------------------------------------------- reference arm
def cpu_gan_step_factory(batch, cfg, seed=7):
    """The reference's CPU path for one GAN iteration, restated by the oracle (torch CPU fp32, all host threads)."""
    from oracle import oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    nup = cfg["factor"].bit_length() - 1
    sg = {k: (v.clone().requires_grad_(True) if k not in ("power", "multiplier") else v)
          for k, v in O.make_generator_sd(1234, filters=cfg["F"], num_res_blocks=cfg["nb"], num_upsample=nup).items()}
    sds = [{k: v.clone().requires_grad_(True) for k, v in O.make_discriminator_sd(1235 + k, gain=2.0).items()}
           for k in range(2)]
    opt_g = torch.optim.Adam([v for v in sg.values() if v.requires_grad], lr=2e-4, betas=(0.9, 0.999))
    opt_d = [torch.optim.Adam(list(sd.values()), lr=2e-4, betas=(0.9, 0.999)) for sd in sds]
    hr, lr = O.synthetic_jets(batch, HR, cfg["factor"], seed)
    gen = torch.Generator().manual_seed(seed)

    def step():
        opt_g.zero_grad()
        out, srs = O.generator_forward(sg, lr, res_scale=RES_SCALE, thres=0.0, training=True)
        loss_g, _ = O.generator_loss((out, srs), hr, lr, sds, factor=cfg["factor"])
        loss_g.backward()
        opt_g.step()
        for k in range(2):
            opt_d[k].zero_grad()
            eps = torch.rand(batch, 1, 1, 1, generator=gen)
            loss_d, _ = O.discriminator_loss(sds[k], hr, (out, srs)[k].detach(), eps, lambda_reg=0.01)
            loss_d.backward()
            if loss_d.item() > 0.001:
                opt_d[k].step()
        return float(loss_g.detach())
    return step


def time_cpu(batch, steps, warmup, cfg):
    step = cpu_gan_step_factory(batch, cfg)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    return batch / dt, dt


CPU_SAMPLE_STEPS = 20      # in-run cpu_baseline sample: ~10 s of host work at 0.5 s per batch-16 step


def time_reference_train(batch, steps, warmup, cfg):
    """The UNMODIFIED reference (baseline/_ref, a copy of the reference tree made by __graft_entry__.build()) through
    its own esrgan.train() on the host cores: tests/ref_harness.py supplies opt, an in-memory synthetic dataset and
    stubs for the plotting imports; nothing of this repo is on that path.  Step time = spacing of consecutive
    iterations (first gradient-penalty draw of each, esrgan.py:598), median over the timed steps."""
    import tempfile
    out = os.path.join(tempfile.mkdtemp(prefix="srb_refarm_"), "trace.json")
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")     # esrgan.py:173 takes cuda:0 when it sees one: this arm is the CPU path
    cmd = [sys.executable, os.path.join(ROOT, "tests", "ref_harness.py"), "--ref", os.path.join(ROOT, "baseline", "_ref"),
           "--out", out, "--n-batches", str(warmup + steps + 1), "--batch-size", str(batch), "--seed", "4321",
           "--factor", str(cfg["factor"]), "--residual-blocks", str(cfg["nb"]), "--preload",
           "--threads", str(os.cpu_count() or 1)]
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=1500)
    if r.returncode != 0:
        raise RuntimeError("reference run failed: " + r.stderr[-300:])
    res = json.load(open(out))
    if res.get("device") != "cpu":
        raise RuntimeError("reference run did not stay on the CPU")
    t = res["step_t"]
    d = sorted(b - a for a, b in zip(t[warmup:-1], t[warmup + 1:]))
    dt = d[len(d) // 2]
    return batch / dt, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    cfg = CONFIGS["deep" if args.config == "deep" else "gan"]
    steps, warm = max(1, min(args.steps, 20)), max(1, min(args.warmup, 2))
    kind, how = "port", (f"batch {args.cpu_batch} per step on the host cores (oracle/oracle.py, torch CPU fp32; pinned "
                         "against the reference's own esrgan.train trace, tests/test_trace_oracle.py)")
    if os.path.isdir(os.path.join(ROOT, "baseline", "_ref")):
        try:
            ips, dt = time_reference_train(args.cpu_batch, steps, warm, cfg)
            kind, how = "reference", (f"batch {args.cpu_batch} per step: the unmodified reference's own esrgan.train() "
                                      "(baseline/_ref, torch CPU fp32, all host threads) driven by tests/ref_harness.py, "
                                      "dataset preloaded in memory")
        except Exception as e:  # noqa: BLE001  (fall back to the port, say why)
            ips, dt = time_cpu(args.cpu_batch, steps, warm, cfg)
            how += f"; reference run failed: {type(e).__name__}: {str(e)[:120]}"
    else:
        ips, dt = time_cpu(args.cpu_batch, steps, warm, cfg)
    cores = os.cpu_count() or 1
    line = {
        "impl": "reference", "metric": "esrgan_gan_train_images_per_s", "value": ips, "unit": "images/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warm,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic sparse jet images (20-60 hits/image, Exp(5) energies), seeded weights",
        "config": {"workload": cfg["text"], "per_step_batch": args.cpu_batch},
        "cpu_baseline": {"value": ips, "unit": "images/s", "cores": cores, "kind": kind, "sample": how},
        "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------- our arm
def parity_check(dev, cfg):
    """Step-0 loss components of the fused step against the CPU oracle on a small batch (same seeds as the bench
    models), before anything is timed: a bench whose kernels drifted from the reference must not print a number."""
    from oracle import oracle as O
    B = 8
    G, Ds = build_models(dev, cfg)
    tr = make_trainer(G, Ds, cfg, None)
    hr, lr = O.synthetic_jets(B, HR, cfg["factor"], 4242)
    gen = torch.Generator().manual_seed(5)
    eps = [torch.rand(B, 1, 1, 1, generator=gen) for _ in range(2)]
    nup = cfg["factor"].bit_length() - 1
    sg = O.make_generator_sd(1234, filters=cfg["F"], num_res_blocks=cfg["nb"], num_upsample=nup)
    sds = [O.make_discriminator_sd(1235 + k, gain=2.0) for k in range(2)]
    with torch.no_grad():
        out, srs = O.generator_forward(sg, lr, res_scale=RES_SCALE, thres=0.0, training=True)
        ref_g, parts = O.generator_loss((out, srs), hr, lr, sds, factor=cfg["factor"])
    got = tr.step(lr.to(dev), hr.to(dev), epsilons=[e.to(dev) for e in eps])
    worst = 0.0
    rep = {}
    for k in ("hr0", "lr0", "adv0", "hr1", "adv1"):
        rel = abs(float(got[k]) - float(parts[k])) / abs(float(parts[k]))
        rep[k] = rel
        worst = max(worst, rel)
    rel = abs(float(got["g_loss"]) - float(ref_g)) / abs(float(ref_g))
    rep["g_loss"] = rel
    worst = max(worst, rel)
    for k in range(2):
        ref_d, _ = O.discriminator_loss(sds[k], hr, (out, srs)[k], eps[k], lambda_reg=0.01)
        rel = abs(float(got[f"d_loss{k}"]) - float(ref_d)) / abs(float(ref_d))
        rep[f"d_loss{k}"] = rel
        worst = max(worst, rel)
    if not worst <= 2e-2:
        raise RuntimeError(f"CUDA srb200 bench: step-0 losses deviate from the oracle by {worst:.3e} (> 2e-2): {rep}")
    del tr, G, Ds
    torch.cuda.empty_cache()
    return {"batch": B, "tolerance": 2e-2, "worst_rel": worst, "components": rep,
            "what": "step-0 loss components of GANTrainStep vs oracle/oracle.py (CPU fp32), checked before timing"}


def run_ours(args):
    import torch.distributed as dist
    from oracle import oracle as O
    from srb200 import _lib as L, ops
    from srb200.train_step import GeneratorTrainStep, GraphedStep

    world = int(os.environ.get("WORLD_SIZE", 1))
    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    pg = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        pg = dist.group.WORLD

    def barrier():
        if world > 1:
            dist.barrier()

    L.lib()                                   # fail loudly here if the CUDA extension is missing
    if args.config == "sweep":
        return run_sweep(args, dev, world, rank, pg, barrier)
    cfg = CONFIGS[args.config]
    B = args.batch or cfg["batch"]
    gf = gflop_fwd_per_img(cfg["nb"], cfg["F"], cfg["factor"])
    parity = None
    if not args.no_parity_check and rank == 0:
        parity = parity_check(dev, cfg)
    barrier()
    t_build = time.perf_counter()
    G, Ds = build_models(dev, cfg)
    trainer = make_trainer(G, Ds, cfg, pg)
    hr_cpu, lr_cpu = O.synthetic_jets(B, HR, cfg["factor"], 1234 + rank)
    hr, lr = hr_cpu.to(dev), lr_cpu.to(dev)
    torch.cuda.synchronize()
    t_build = time.perf_counter() - t_build

    # ---- headline: device-resident inputs
    sampler = ClockSampler(local)
    state = {}
    graphed = None
    t_capture = None
    if not args.no_graph:
        c0 = L.CALLS[0]
        t0 = time.perf_counter()
        graphed = GraphedStep(trainer, lr, hr, warmup=2)       # 2 eager steps + 1 captured
        _GRAPHS.append(graphed.graph)
        torch.cuda.synchronize()
        t_capture = time.perf_counter() - t0
        calls_per_step = (L.CALLS[0] - c0) // 3
    calls0 = L.CALLS[0]
    if rank == 0:
        sampler.start()

    def step_dev():
        state["losses"] = graphed.step(lr, hr) if graphed is not None else trainer.step(lr, hr)

    ms = timed(step_dev, args.steps, args.warmup, barrier)
    losses_f = {k: float(v) for k, v in state["losses"].items()}
    calls = (L.CALLS[0] - calls0) * args.steps // (args.steps + args.warmup) if graphed is None else calls_per_step * args.steps
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = world * B / (ms * 1e-3)

    # ---- end to end: pinned host inputs, H2D every step, D2H of the loss scalar
    hr_pin, lr_pin = hr_cpu.pin_memory(), lr_cpu.pin_memory()
    loss_host = torch.zeros(1).pin_memory()

    # Input pipeline of a real training loop: the H2D copy of step i+1 is issued on a copy stream while step i runs
    # (double buffered; every timed step still issues exactly one H2D copy of a full batch and one D2H of the loss).
    copy_stream = torch.cuda.Stream()
    dbuf = [(torch.empty_like(hr), torch.empty_like(lr)) for _ in range(2)]
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    e2e_state = {"i": 0}

    def prefetch(i):
        with torch.cuda.stream(copy_stream):
            if i >= 2:
                copy_stream.wait_event(consumed[i % 2])        # the step that read this buffer has finished
            dbuf[i % 2][0].copy_(hr_pin, non_blocking=True)
            dbuf[i % 2][1].copy_(lr_pin, non_blocking=True)
            ready[i % 2].record(copy_stream)

    prefetch(0)

    def step_e2e():
        i = e2e_state["i"]
        main = torch.cuda.current_stream()
        main.wait_event(ready[i % 2])
        prefetch(i + 1)
        h, l_ = dbuf[i % 2]
        losses = graphed.step(l_, h) if graphed is not None else trainer.step(l_, h)
        consumed[i % 2].record(main)
        loss_host.copy_(losses["g_loss"].reshape(1), non_blocking=True)
        e2e_state["i"] = i + 1

    ms_e2e = timed(step_e2e, args.steps, 1, barrier)
    copy_stream.synchronize()
    t = torch.tensor([ms_e2e], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_e2e = float(t.item())
    e2e = {"value": world * B / (ms_e2e * 1e-3), "unit": "images/s",
           "h2d_bytes_per_step": int(hr_pin.numel() * 4 + lr_pin.numel() * 4), "d2h_bytes_per_step": 4,
           "ms_per_step": ms_e2e}

    # ---- roofline of the dominant kernel (conv3x3 implicit GEMM): one instrumented step after the timed region,
    #      CUDA events around every launch on the launching stream
    pk = peaks()
    ops.PROFILE = []
    d_streams, trainer.d_streams = trainer.d_streams, False     # kernels timed alone: no discriminator work beside them
    trainer.step(lr, hr)
    torch.cuda.synchronize()
    trainer.d_streams = d_streams
    prof, ops.PROFILE = ops.PROFILE, None
    agg = {}
    for kind, fl, e0, e1 in prof:
        a = agg.setdefault(kind, [0.0, 0.0, 0])
        a[0] += fl; a[1] += e0.elapsed_time(e1); a[2] += 1
    conv = agg.get("conv3x3", [0.0, 1e-9, 0])
    wg = agg.get("wgrad3x3", [0.0, 1e-9, 0])
    ach = conv[0] / (conv[1] * 1e-3) / 1e12
    chain = G._engine().chain
    pair = os.environ.get("SRB_CONV_PAIR", "1") != "0"
    kname = (("conv3x3_chain_pair_kernel (tcgen05.mma.cta_group::2 implicit GEMM on CTA pairs, %d convolutions per "
              "persistent launch in the RRDB trunk, single launches for the remaining layers); fwd + dgrad launches" % chain)
             if pair else
             ("conv3x3_chain_kernel (tcgen05 implicit GEMM, %d convolutions per persistent launch in the RRDB trunk) + "
              "conv3x3_igemm_kernel (remaining layers); fwd + dgrad launches" % chain)) if chain else \
        "conv3x3_igemm_kernel (tcgen05 implicit GEMM, fwd + dgrad launches)"
    traffic = traffic_from_profile() if (chain == 5 and args.config == "gan" and B == 512) else None
    roofline = {"bound": "tensor", "kernel": kname,
                "achieved": ach, "peak": pk["bf16_sustained"], "unit": "TFLOP/s", "frac": ach / pk["bf16_sustained"],
                "frac_of_burst": ach / pk["bf16_burst"], "peak_source": pk["source"] + ", sustained figure",
                "traffic": traffic[0] if traffic else None,
                "traffic_note": ("dram__bytes_read.sum + dram__bytes_write.sum per chained launch (one dense block, "
                                 "B=512), read from " + traffic[1] + "; bytes") if traffic else None,
                "launches_per_step": conv[2], "kernel_ms_per_step": conv[1],
                "share_of_step": conv[1] / ms,
                "wgrad": {"achieved": wg[0] / (wg[1] * 1e-3) / 1e12, "launches_per_step": wg[2],
                          "kernel_ms_per_step": wg[1], "share_of_step": wg[1] / ms,
                          "frac": wg[0] / (wg[1] * 1e-3) / 1e12 / pk["bf16_sustained"]},
                "ceiling_note": "Cout = 64: a 128x64x16 MMA reads 6 KB of shared-memory operands per 32 tensor clocks "
                                "= 192 B/clk against 128 B/clk per SM (48 clk per MMA measured); on CTA pairs "
                                "(cta_group::2, weights split) 5 KB per CTA = 40 clk on paper, 43 measured "
                                "(profiles/r01_s21_mma_rate_probe.txt): 0.74 of the tensor rate at best",
                "how": "algorithmic FLOPs (2*B*H*W*9*Cin*Cout per launch, summed over a chained launch) / CUDA-event "
                       "time per launch, summed over one instrumented step run right after the timed region"}
    train_flops = world * B * (3 * gf) * 1e9
    extras = {"tensor_frac_step_vs_sustained": train_flops / (ms * 1e-3) / 1e12 / pk["bf16_sustained"] / world,
              "gflop_fwd_per_image": gf,
              "setup_s": {"construct_models_and_trainer": t_build, "warmup_and_graph_capture": t_capture}}

    # ---- inference (BASELINE.json configs[1]) at every N: device-resident and end to end (LR H2D + SR D2H per batch,
    #      what evaluation/eval.py:475 does)
    if not args.no_extras:
        G.eval()
        Bi = 512 if args.config == "gan" else B
        xb_cpu = O.synthetic_jets(Bi, HR, cfg["factor"], 99 + rank)[1]
        xb = xb_cpu.to(dev)
        with torch.no_grad():
            ms_inf = timed(lambda: G(xb), max(3, args.steps), 2, barrier)
            xb_pin = xb_cpu.pin_memory()
            sr_host = torch.empty(Bi, 1, HR, HR).pin_memory()
            xb_dev = torch.empty_like(xb)

            def inf_e2e():
                xb_dev.copy_(xb_pin, non_blocking=True)
                sr_host.copy_(G(xb_dev), non_blocking=True)

            ms_inf_e2e = timed(inf_e2e, max(3, args.steps), 2, barrier)
        G.train()
        t = torch.tensor([ms_inf, ms_inf_e2e], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_inf, ms_inf_e2e = float(t[0]), float(t[1])
        extras["inference"] = {
            "images_per_s": world * Bi / (ms_inf * 1e-3), "per_gpu_batch": Bi, "n_gpus": world,
            "tensor_frac_vs_sustained": Bi * gf / (ms_inf * 1e-3) / 1e3 / pk["bf16_sustained"],
            "e2e_images_per_s": world * Bi / (ms_inf_e2e * 1e-3),
            "e2e_h2d_bytes_per_step": int(xb_pin.numel() * 4), "e2e_d2h_bytes_per_step": int(sr_host.numel() * 4)}
        extras["inference_b512_images_per_s"] = Bi / (ms_inf * 1e-3)
        extras["inference_tensor_frac_vs_sustained"] = extras["inference"]["tensor_frac_vs_sustained"]

    # ---- secondary numbers (rank 0, single GPU legs)
    if not args.no_extras and world == 1:
        gts = GeneratorTrainStep(G, process_group=None)
        gts.opt_G = trainer.opt_G
        g_step = lambda: gts.step(lr, hr)                          # noqa: E731
        if not args.no_graph:                                      # same launch mode as the headline: one CUDA graph
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(2):
                    gts.step(lr, hr)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            g_graph = torch.cuda.CUDAGraph()
            _GRAPHS.append(g_graph)
            with torch.cuda.graph(g_graph):
                gts.step(lr, hr)
            g_step = g_graph.replay
        ms_g = timed(g_step, max(3, args.steps), 2, barrier)
        extras["generator_train_images_per_s"] = B / (ms_g * 1e-3)
        extras["generator_train_tensor_frac_vs_sustained"] = B * 3 * gf / (ms_g * 1e-3) / 1e3 / pk["bf16_sustained"]
        extras["generator_train_tensor_frac_vs_burst"] = B * 3 * gf / (ms_g * 1e-3) / 1e3 / pk["bf16_burst"]
        # data path: constituent lists -> HR/LR batch on the device, HBM-bound on the image write
        from srb200 import data as srb_data
        pairs = O.synthetic_constituents(B, 128, HR, HR, 3).to(dev)
        ms_dp = timed(lambda: srb_data.extract_batch(pairs, HR, HR, cfg["factor"]), 20, 3, barrier)
        dp_bytes = B * (128 * 8 + HR * HR * 4 * (1 + 1.0 / cfg["factor"] ** 2))
        extras["data_path"] = {"images_per_s": B / (ms_dp * 1e-3), "us_per_batch": ms_dp * 1e3,
                               "achieved_GBps": dp_bytes / (ms_dp * 1e-3) / 1e9, "peak_GBps": pk.get("hbm", None),
                               "workload": f"extract_batch, {B} lists of 128 (index, energy) pairs -> {HR}x{HR} + pooled"}
        # evaluation statistics: one pass over a batch of HR images, HBM-bound on the read
        imgs = hr[:B].contiguous()
        ms_st = timed(lambda: srb_data.image_stats(imgs, 0.05, 4), 20, 3, barrier)
        extras["eval_stats"] = {"images_per_s": B / (ms_st * 1e-3), "us_per_batch": ms_st * 1e3,
                                "achieved_GBps": B * HR * HR * 4 / (ms_st * 1e-3) / 1e9, "peak_GBps": pk.get("hbm", None),
                                "workload": f"image_stats (max, min, mean, nnz, sum, meannnz, 4 hardest) of {B} {HR}x{HR} images"}
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_extras:
        kind = "port"
        try:
            if not os.path.isdir(os.path.join(ROOT, "baseline", "_ref")):
                raise FileNotFoundError("baseline/_ref")
            ips, dt = time_reference_train(args.cpu_batch, CPU_SAMPLE_STEPS, 1, cfg)
            kind, what = "reference", "the unmodified reference's own esrgan.train() (baseline/_ref) via tests/ref_harness.py"
        except Exception:  # noqa: BLE001
            ips, dt = time_cpu(args.cpu_batch, CPU_SAMPLE_STEPS, 1, cfg)
            what = "oracle/oracle.py (torch CPU fp32, the reference's own backend)"
        cpu_baseline = {"value": ips, "unit": "images/s", "cores": os.cpu_count() or 1, "kind": kind,
                        "sample": f"{CPU_SAMPLE_STEPS} GAN steps of batch {args.cpu_batch} ({dt:.2f} s/step, median) on the host cores, " + what}
    gpu_baseline = None
    if rank == 0 and world == 1 and not args.no_extras and not args.no_gpu_baseline and args.config == "gan":
        # the bar SURVEY.md section 2a sets: the reference's modules under stock PyTorch / cuDNN on this GPU
        # (tools/gpu_baseline.py has the full matrix; this is the bounded in-run sample).  Our buffers are released first.
        del graphed, trainer, G, Ds
        torch.cuda.empty_cache()
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import gpu_baseline as GB
        rows = []
        for wl in ("gan_step", "inference"):
            for prec in ("bf16_channels_last", "fp32"):
                res, err = GB.measure(dev, B, prec, wl, False, steps=3, warmup=2)
                row = {"workload": wl, "precision": prec, "launch": "eager", "batch": B}
                row.update(res if res is not None else {"unavailable": err})
                rows.append(row)
        best = {}
        for r in rows:
            if "images_per_s" in r:
                best[r["workload"]] = max(best.get(r["workload"], 0.0), r["images_per_s"])
        gpu_baseline = {"what": "the reference's modules under stock PyTorch/cuDNN on the same GPU (not the CPU "
                                "reference arm, not our kernels); full matrix: tools/gpu_baseline.py, BASELINE.md section 5",
                        "torch": torch.__version__, "cudnn": torch.backends.cudnn.version(), "rows": rows,
                        "best_gan_step_images_per_s": best.get("gan_step"),
                        "best_inference_images_per_s": best.get("inference"),
                        "ours_over_best_gan_step": value / best["gan_step"] if best.get("gan_step") else None}
    if rank == 0:
        line = {
            "metric": "esrgan_gan_train_images_per_s", "value": value, "unit": "images/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic sparse jet images (20-60 hits/image, Exp(5) energies), seeded weights",
            "config": {"workload": cfg["text"],
                       "per_gpu_batch": B, "global_batch": world * B, "parallelism": f"dp{world}",
                       "launch": "eager" if args.no_graph else "one CUDA graph per step (C-ABI launches captured; the two "
                                 "discriminators on their own streams = parallel graph branches)",
                       "l2": "per-step working set (activations + gradients, > 1 GB) exceeds the 126 MB L2",
                       "dp_semantics": "batch statistics of the losses (esrgan.py:487,501-508) are per rank, as DDP "
                                       "around the reference would compute them"},
            "e2e": e2e, "gpu_launches": int(calls), "roofline": roofline, "cpu_baseline": cpu_baseline,
            "gpu_baseline": gpu_baseline, "parity": parity,
            "clocks": clocks, "extras": extras,
            "losses": losses_f,
        }
        print(json.dumps(line), flush=True)
    finish(world)


_GRAPHS = []        # every CUDA graph this process captured (they hold NCCL kernels when world > 1)


def finish(world):
    """Orderly multi-rank teardown.  destroy_process_group() hung in round 1 while captured graphs that contain NCCL
    kernels were still alive: release the graphs' device resources first (CUDAGraph.reset()), synchronise, meet at a
    barrier, then destroy the communicator.  A watchdog keeps the old hard exit as the fallback, so that a stalled
    teardown can never hold the node (everything has been printed and flushed by then)."""
    if world <= 1:
        return
    import threading
    import torch.distributed as dist
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    sys.stderr.flush()
    if os.environ.get("SRB_BENCH_HARD_EXIT", "0") == "1":
        os._exit(0)
    timer = threading.Timer(float(os.environ.get("SRB_BENCH_TEARDOWN_S", "20")), lambda: os._exit(0))
    timer.daemon = True
    timer.start()
    for g in _GRAPHS:
        g.reset()
    del _GRAPHS[:]
    import gc
    gc.collect()
    torch.cuda.synchronize()
    dist.destroy_process_group()
    timer.cancel()
    print("teardown: process group destroyed", file=sys.stderr, flush=True)


# ----------------------------------------------------------------------------------------------- hyper-search sweep
def run_sweep(args, dev, world, rank, pg, barrier):
    """BASELINE.json configs[4]: hyper_search.grid_search (hyper_search.py:25-113) over options/hyper_ranges.json calls
    train() afresh for each of 27 points = 3 generator depths (10 / 15 / 23 RRDBs, batch 5 / 3 / 2) x 9 loss-weight
    points at the literal default factor 2 (LR 40x40).  The loss weights do not change the work, so one point per depth
    is timed - at the reference's batch sizes and x64 scaled - plus the width axis F = 128 that the harness adds through
    the `filters` constructor argument.  Per point: seconds to construct models + optimizers, to warm up and capture the
    CUDA graph (what every train() call pays), and ms per GAN step."""
    from oracle import oracle as O
    from srb200 import _lib as L
    from srb200.train_step import GraphedStep
    pk = peaks()
    points = []
    for F in (64, 128):
        for nb, b_ref in ((10, 5), (15, 3), (23, 2)):
            for scale in (1, 64):
                if F == 128 and scale == 1:
                    continue
                points.append(dict(nb=nb, F=F, factor=2, batch=b_ref * scale))
    rows = []
    t_all = time.perf_counter()
    for pt in points:
        cfg = dict(pt, text="")
        B = pt["batch"]
        try:
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            G, Ds = build_models(dev, cfg)
            trainer = make_trainer(G, Ds, cfg, pg)
            hr, lr = O.synthetic_jets(B, HR, 2, 1234 + rank)
            hr, lr = hr.to(dev), lr.to(dev)
            torch.cuda.synchronize()
            t_build = time.perf_counter() - t0
            t0 = time.perf_counter()
            graphed = GraphedStep(trainer, lr, hr, warmup=2)
            _GRAPHS.append(graphed.graph)
            torch.cuda.synchronize()
            t_cap = time.perf_counter() - t0
            ms = timed(lambda: graphed.step(lr, hr), max(3, min(args.steps, 10)), 2, barrier)
            t = torch.tensor([ms], device=dev)
            if world > 1:
                torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            ms = float(t.item())
            gf = gflop_fwd_per_img(pt["nb"], pt["F"], 2)
            rows.append(dict(pt, per_gpu_batch=B, ms_per_step=ms, images_per_s=world * B / (ms * 1e-3),
                             construct_s=t_build, warmup_capture_s=t_cap,
                             tensor_frac_vs_sustained=B * 3 * gf / (ms * 1e-3) / 1e3 / pk["bf16_sustained"]))
            del graphed, trainer, G, Ds
        except RuntimeError as e:           # hyper_search.py:108-110 catches exactly this and moves on
            rows.append(dict(pt, per_gpu_batch=B, unavailable=str(e)[:200]))
        torch.cuda.empty_cache()
    total_s = time.perf_counter() - t_all
    if rank == 0:
        big = [r for r in rows if "images_per_s" in r and r["per_gpu_batch"] >= 128 and r["F"] == 64]
        value = sum(r["images_per_s"] for r in big) / max(len(big), 1)
        line = {"metric": "esrgan_gan_train_images_per_s", "value": value, "unit": "images/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": None, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
                "data": "synthetic sparse jet images (20-60 hits/image, Exp(5) energies), seeded weights",
                "config": {"workload": "hyper_search sweep (hyper_search.py:25-113, options/hyper_ranges.json): GAN train "
                                       "step per generator shape, factor 2 (LR 40x40 -> HR 80x80); value = mean images/s "
                                       "of the F=64 points at the x64-scaled batch sizes",
                           "parallelism": f"dp{world}", "points": len(rows)},
                "extras": {"sweep": rows, "sweep_wall_s": total_s,
                           "note": "27 grid points of the reference = these 3 depths x 9 loss-weight points (same work "
                                   "per depth); construct_s + warmup_capture_s is what every train() call pays"},
                "gpu_launches": int(L.CALLS[0])}
        print(json.dumps(line), flush=True)
    finish(world)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
