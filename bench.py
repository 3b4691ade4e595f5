#!/usr/bin/env python
"""Benchmark of the B200-native ESRGAN hot path (BASELINE.json metric: ESRGAN train images/s).

  python bench.py --gpus N --steps K --warmup W            our arm   (torchrun for N > 1, one rank per GPU, NCCL)
  python bench.py --impl reference --gpus N ...            reference arm: the CPU oracle (oracle/oracle.py, a
                                                           restatement of the reference's own torch-CPU path) on the
                                                           host cores; rank 0 only.

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

NB, FILTERS, NUP, HR, FACTOR = 10, 64, 2, 80, 4
RES_SCALE = 0.1                      # options/default.json "res_scale"
GFLOP_FWD_PER_IMG = 14.370           # BASELINE.md section 3 (nb=10, F=64, x4, HR 80), algorithmic
GFLOP_D_FWD_PER_IMG = 0.0461


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=int(os.environ.get("SRB_BENCH_BATCH", 512)), help="per-GPU batch")
    ap.add_argument("--cpu-batch", type=int, default=16)
    ap.add_argument("--no-graph", action="store_true", help="launch the step eagerly instead of replaying one CUDA graph")
    ap.add_argument("--no-extras", action="store_true", help="skip inference / generator-only / cpu_baseline legs")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------- helpers
def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(bf16_burst=d.get("bf16_tflops"), bf16_sustained=d.get("bf16_tflops_sustained"),
                    hbm=d.get("hbm_gbs"), source="measured (MEASURED_PEAKS.json)")
    return dict(bf16_burst=1590.0, bf16_sustained=1400.0, hbm=6650.0, source="fallback (B200_PROFILING.md)")


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
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i] == "Active"})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def build_models(dev):
    import models
    from oracle import oracle as O
    G = models.GeneratorRRDB(1, filters=FILTERS, num_res_blocks=NB, num_upsample=NUP, res_scale=RES_SCALE)
    G.load_state_dict(O.make_generator_sd(1234, num_res_blocks=NB, num_upsample=NUP))
    Ds = []
    for k in range(2):
        d = models.Markovian_Discriminator((1, HR, HR), [16, 32, 32, 64])
        d.load_state_dict(O.make_discriminator_sd(1235 + k, gain=2.0))
        Ds.append(d.to(dev))
    return G.to(dev).train(), Ds


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
def cpu_gan_step_factory(batch, seed=7):
    """The reference's CPU path for one GAN iteration, restated by the oracle (torch CPU fp32, all host threads)."""
    from oracle import oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    sg = {k: (v.clone().requires_grad_(True) if k not in ("power", "multiplier") else v)
          for k, v in O.make_generator_sd(1234, num_res_blocks=NB, num_upsample=NUP).items()}
    sds = [{k: v.clone().requires_grad_(True) for k, v in O.make_discriminator_sd(1235 + k, gain=2.0).items()}
           for k in range(2)]
    opt_g = torch.optim.Adam([v for v in sg.values() if v.requires_grad], lr=2e-4, betas=(0.9, 0.999))
    opt_d = [torch.optim.Adam(list(sd.values()), lr=2e-4, betas=(0.9, 0.999)) for sd in sds]
    hr, lr = O.synthetic_jets(batch, HR, FACTOR, seed)
    gen = torch.Generator().manual_seed(seed)

    def step():
        opt_g.zero_grad()
        out, srs = O.generator_forward(sg, lr, res_scale=RES_SCALE, thres=0.0, training=True)
        loss_g, _ = O.generator_loss((out, srs), hr, lr, sds, factor=FACTOR)
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


def time_cpu(batch, steps, warmup):
    step = cpu_gan_step_factory(batch)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    return batch / dt, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    ips, dt = time_cpu(args.cpu_batch, max(1, min(args.steps, 5)), max(1, min(args.warmup, 1)))
    cores = os.cpu_count() or 1
    line = {
        "impl": "reference", "metric": "esrgan_gan_train_images_per_s", "value": ips, "unit": "images/s",
        "n_gpus": args.gpus, "steps": max(1, min(args.steps, 5)), "warmup": max(1, min(args.warmup, 1)),
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic sparse jet images (20-60 hits/image, Exp(5) energies), seeded weights",
        "config": {"workload": "full ESRGAN GAN train step, GeneratorRRDB nb=10 F=64 x4 (LR 20x20 -> HR 80x80), "
                               "2 patch discriminators, RaGAN + gradient penalty, 3 Adam",
                   "per_step_batch": args.cpu_batch},
        "cpu_baseline": {"value": ips, "unit": "images/s", "cores": cores, "kind": "port",
                         "sample": f"batch {args.cpu_batch} per step on the host cores (oracle/oracle.py, torch CPU fp32)"},
        "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch.distributed as dist
    from oracle import oracle as O
    from srb200 import _lib as L, ops
    from srb200.train_step import GANTrainStep, GeneratorTrainStep, GraphedStep

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
    B = args.batch
    G, Ds = build_models(dev)
    trainer = GANTrainStep(G, Ds, factor=FACTOR, lambdas=(0.2, 1.0), lambda_hr=1.0, lambda_lr=0.1, lambda_adv=0.01,
                           lambda_reg=0.01, d_threshold=0.001, lr=2e-4, process_group=pg)
    hr_cpu, lr_cpu = O.synthetic_jets(B, HR, FACTOR, 1234 + rank)
    hr, lr = hr_cpu.to(dev), lr_cpu.to(dev)

    # ---- headline: device-resident inputs
    sampler = ClockSampler(local)
    state = {}
    graphed = None
    if not args.no_graph:
        c0 = L.CALLS[0]
        graphed = GraphedStep(trainer, lr, hr, warmup=2)       # 2 eager steps + 1 captured
        calls_per_step = (L.CALLS[0] - c0) // 3
    calls0 = L.CALLS[0]
    if rank == 0:
        sampler.start()

    def step_dev():
        state["losses"] = graphed.step(lr, hr) if graphed is not None else trainer.step(lr, hr)

    ms = timed(step_dev, args.steps, args.warmup, barrier)
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
    trainer.step(lr, hr)
    torch.cuda.synchronize()
    prof, ops.PROFILE = ops.PROFILE, None
    agg = {}
    for kind, fl, e0, e1 in prof:
        a = agg.setdefault(kind, [0.0, 0.0, 0])
        a[0] += fl; a[1] += e0.elapsed_time(e1); a[2] += 1
    conv = agg.get("conv3x3", [0.0, 1e-9, 0])
    wg = agg.get("wgrad3x3", [0.0, 1e-9, 0])
    ach = conv[0] / (conv[1] * 1e-3) / 1e12
    chain = G._engine().chain
    kname = ("conv3x3_chain_kernel (tcgen05 implicit GEMM, %d convolutions per persistent launch in the RRDB trunk) + "
             "conv3x3_igemm_kernel (remaining layers); fwd + dgrad launches" % chain) if chain else \
        "conv3x3_igemm_kernel (tcgen05 implicit GEMM, fwd + dgrad launches)"
    roofline = {"bound": "tensor", "kernel": kname,
                "achieved": ach, "peak": pk["bf16_sustained"], "unit": "TFLOP/s", "frac": ach / pk["bf16_sustained"],
                "frac_of_burst": ach / pk["bf16_burst"], "peak_source": pk["source"] + ", sustained figure",
                "traffic": 450.2e6 if chain == 5 else None,
                "traffic_note": "dram__bytes_read.sum + dram__bytes_write.sum per chained launch (one dense block, "
                                "B=512) from profiles/r01_s29b_ncu_full_conv3x3_chain.txt; bytes" if chain == 5 else None,
                "launches_per_step": conv[2], "kernel_ms_per_step": conv[1],
                "share_of_step": conv[1] / ms,
                "wgrad": {"achieved": wg[0] / (wg[1] * 1e-3) / 1e12, "launches_per_step": wg[2],
                          "kernel_ms_per_step": wg[1], "share_of_step": wg[1] / ms},
                "ceiling_note": "Cout = 64 makes every MMA 128x64x16: 6 KB of shared-memory operands per 32 tensor "
                                "clocks = 192 B/clk against 128 B/clk per SM, i.e. 2/3 of the tensor rate at best "
                                "(48 clk per MMA measured, profiles/r01_s21_mma_rate_probe.txt)",
                "how": "algorithmic FLOPs (2*B*H*W*9*Cin*Cout per launch, summed over a chained launch) / CUDA-event "
                       "time per launch, summed over one instrumented step run right after the timed region"}
    train_flops = world * B * (3 * GFLOP_FWD_PER_IMG) * 1e9
    extras = {"tensor_frac_step_vs_sustained": train_flops / (ms * 1e-3) / 1e12 / pk["bf16_sustained"] / world}

    # ---- secondary numbers (rank 0, single GPU legs)
    if not args.no_extras and world == 1:
        G.eval()
        with torch.no_grad():
            xb = O.synthetic_jets(512, HR, FACTOR, 99)[1].to(dev)
            ms_inf = timed(lambda: G(xb), max(3, args.steps), 2, barrier)
        G.train()
        extras["inference_b512_images_per_s"] = 512 / (ms_inf * 1e-3)
        extras["inference_tensor_frac_vs_sustained"] = 512 * GFLOP_FWD_PER_IMG / (ms_inf * 1e-3) / 1e3 / pk["bf16_sustained"]
        gts = GeneratorTrainStep(G, process_group=None)
        gts.opt_G = trainer.opt_G
        ms_g = timed(lambda: gts.step(lr, hr), max(3, args.steps), 2, barrier)
        extras["generator_train_images_per_s"] = B / (ms_g * 1e-3)
        extras["generator_train_tensor_frac_vs_sustained"] = B * 3 * GFLOP_FWD_PER_IMG / (ms_g * 1e-3) / 1e3 / pk["bf16_sustained"]
        # data path: constituent lists -> HR/LR batch on the device, HBM-bound on the image write
        from srb200 import data as srb_data
        pairs = O.synthetic_constituents(B, 128, HR, HR, 3).to(dev)
        ms_dp = timed(lambda: srb_data.extract_batch(pairs, HR, HR, FACTOR), 20, 3, barrier)
        dp_bytes = B * (128 * 8 + HR * HR * 4 * (1 + 1.0 / FACTOR ** 2))
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
        ips, dt = time_cpu(args.cpu_batch, 3, 1)
        cpu_baseline = {"value": ips, "unit": "images/s", "cores": os.cpu_count() or 1, "kind": "port",
                        "sample": f"3 GAN steps of batch {args.cpu_batch} ({dt:.2f} s/step) on the host cores, "
                                  "oracle/oracle.py (torch CPU fp32, the reference's own backend)"}
    if rank == 0:
        line = {
            "metric": "esrgan_gan_train_images_per_s", "value": value, "unit": "images/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic sparse jet images (20-60 hits/image, Exp(5) energies), seeded weights",
            "config": {"workload": "full ESRGAN GAN train step, GeneratorRRDB nb=10 F=64 x4 (LR 20x20 -> HR 80x80), "
                                   "2 patch discriminators, RaGAN + gradient penalty, 3 Adam",
                       "per_gpu_batch": B, "global_batch": world * B, "parallelism": f"dp{world}",
                       "launch": "eager" if graphed is None else "one CUDA graph per step (C-ABI launches captured)",
                       "l2": "per-step working set (activations + gradients, > 1 GB) exceeds the 126 MB L2"},
            "e2e": e2e, "gpu_launches": int(calls), "roofline": roofline, "cpu_baseline": cpu_baseline,
            "clocks": clocks, "extras": extras,
            "losses": {k: float(v) for k, v in state["losses"].items()},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        # The NCCL all-reduces live inside the captured CUDA graph; tearing the communicator down with
        # destroy_process_group() after graph replays was seen to hang (session 27, 2 GPUs) although every rank had
        # finished.  All results are printed by now: agree that everyone is done, then leave without the teardown.
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
