"""GPU bring-up probe: correctness of the tcgen05 conv/wgrad kernels against the naive CUDA kernels and torch,
kernel timings at benchmark scale, and the raw tcgen05.mma issue-rate microbenchmark.
Every section runs in its own subprocess under a timeout so a trapped kernel cannot take the others down.
Usage: python tools/probe1.py [section ...]   (no args = all sections, each in a subprocess)."""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200"))
sys.path.insert(0, ROOT)


def _setup():
    import torch
    from srb200 import _lib as L, ops
    return torch, L, ops


def rand_pnhwc(torch, B, H, W, C, seed, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    buf = torch.zeros(B, H + 1, W + 1, C, dtype=torch.bfloat16, device="cuda")
    buf[:, :H, :W, :] = (torch.randn(B, H, W, C, generator=g, device="cuda") * scale).to(torch.bfloat16)
    return buf


def sec_conv(flags):
    torch, L, ops = _setup()
    import torch.nn.functional as F
    cases = [  # B,H,W,cin,cout, variant
        (2, 8, 8, 64, 64, "plain"), (3, 20, 20, 128, 64, "lrelu"), (2, 20, 20, 320, 64, "res"),
        (5, 7, 13, 192, 128, "mask"), (2, 10, 10, 64, 256, "shuffle"), (128, 20, 20, 128, 64, "lrelu"),
        (3, 80, 80, 64, 64, "plain"), (40, 40, 40, 64, 64, "res"),
    ]
    out = []
    for (B, H, W, cin, cout, var) in cases:
        Ctot = cin + 64
        x = rand_pnhwc(torch, B, H, W, Ctot, 1)
        g = torch.Generator(device="cuda").manual_seed(2)
        w = torch.randn(cout, cin, 3, 3, generator=g, device="cuda") / (3.0 * cin ** 0.5)
        b = torch.randn(cout, generator=g, device="cuda") * 0.1
        shuffle = var == "shuffle"
        wp = ops.pack_fwd(w, shuffle=shuffle)
        bp = b
        if shuffle:
            C4 = cout // 4
            n = torch.arange(cout, device="cuda")
            bp = b[4 * (n % C4) + n // C4].contiguous()
        kw = dict(alpha=1.0)
        res1 = res2 = mask = None
        if var == "lrelu":
            kw.update(act=L.ACT_LRELU, slope=0.01)
        if var == "res":
            res1 = rand_pnhwc(torch, B, H, W, 64, 3); res2 = rand_pnhwc(torch, B, H, W, 128, 4)
            kw.update(alpha=0.2, res1=(res1, 0, 0.5), res2=(res2, 64, 1.0))
        if var == "mask":
            mask = rand_pnhwc(torch, B, H, W, cout, 5)
            kw.update(act=L.ACT_LRELU_MASK, slope=0.01, mask=(mask, 0))
        ys = []
        for impl in (L.IMPL_NAIVE, L.IMPL_TCGEN05):
            if shuffle:
                y = torch.zeros(B, 2 * H + 1, 2 * W + 1, 64, dtype=torch.bfloat16, device="cuda")
                if var == "shuffle":
                    kw.update(act=L.ACT_LRELU, slope=0.01)
            else:
                y = torch.zeros(B, H + 1, W + 1, cout + 64, dtype=torch.bfloat16, device="cuda")
            ops.conv3x3(x, 64, cin, wp, bp, y, 0 if shuffle else 64, cout, shuffle=shuffle, impl=impl,
                        flags=flags if impl == L.IMPL_TCGEN05 else 0, **kw)
            torch.cuda.synchronize()
            ys.append(y)
        # torch reference on the bf16-rounded operands
        xin = ops.from_pnhwc(x, 64, cin)
        ref = F.conv2d(xin, w.to(torch.bfloat16).float(), b, 1, 1)
        if var in ("lrelu", "shuffle"):
            ref = F.leaky_relu(ref, 0.01)
        if var == "res":
            ref = ref * 0.2 + 0.5 * ops.from_pnhwc(res1, 0, 64)[:, :cout] + ops.from_pnhwc(res2, 64, 64)[:, :cout]
        if var == "mask":
            ref = ref * torch.where(ops.from_pnhwc(mask, 0, cout) > 0, 1.0, 0.01)
        if shuffle:
            ref = F.pixel_shuffle(ref, 2)
            got = [ops.from_pnhwc(y, 0, 64) for y in ys]
        else:
            got = [ops.from_pnhwc(y, 64, cout) for y in ys]
        scale = ref.abs().max().item()
        pad_ok = all(float(y[:, -1].abs().max()) == 0 and float(y[:, :, -1].abs().max()) == 0 for y in ys)
        rec = dict(case=[B, H, W, cin, cout, var], flags=flags, ref_absmax=scale,
                   naive_vs_torch=(got[0] - ref).abs().max().item(), tc_vs_torch=(got[1] - ref).abs().max().item(),
                   tc_vs_naive=(got[1] - got[0]).abs().max().item(), pad_zero=pad_ok, fault=ops.read_fault())
        print(json.dumps(rec), flush=True)
        out.append(rec)
    return out


def sec_wgrad(flags):
    torch, L, ops = _setup()
    import torch.nn.functional as F
    cases = [(2, 8, 8, 64, 64, False), (3, 20, 20, 192, 64, False), (16, 20, 20, 320, 64, False),
             (2, 10, 10, 64, 256, True), (4, 40, 40, 128, 64, False)]
    for (B, H, W, cin, cout, unsh) in cases:
        x = rand_pnhwc(torch, B, H, W, cin + 64, 11)
        gy = rand_pnhwc(torch, B, H, W, cout, 12, 0.1)
        res = []
        for impl in (L.IMPL_NAIVE, L.IMPL_TCGEN05):
            dw = torch.zeros(cout, cin, 3, 3, device="cuda")
            db = torch.zeros(cout, device="cuda")
            ops.wgrad3x3(x, 64, cin, gy, 0, cout, dw, db, alpha=0.5, unshuffle=unsh, impl=impl,
                         flags=flags if impl == L.IMPL_TCGEN05 else 0)
            torch.cuda.synchronize()
            res.append((dw, db))
        xin = ops.from_pnhwc(x, 64, cin).requires_grad_(False)
        gyn = ops.from_pnhwc(gy, 0, cout)
        if unsh:
            C4 = cout // 4
            n = torch.arange(cout, device="cuda")
            src = torch.empty_like(n); src[4 * (n % C4) + n // C4] = n      # original channel o reads packed src[o]
            gyn = gyn[:, src]
        wref = torch.zeros(cout, cin, 3, 3, device="cuda", requires_grad=True)
        out = F.conv2d(xin, wref, None, 1, 1)
        (dwref,) = torch.autograd.grad(out, wref, gyn)
        dwref = 0.5 * dwref
        dbref = 0.5 * gyn.sum((0, 2, 3))
        s = dwref.abs().max().item()
        print(json.dumps(dict(case=[B, H, W, cin, cout, unsh], flags=flags, ref_absmax=s,
                              naive_vs_torch=(res[0][0] - dwref).abs().max().item(),
                              tc_vs_torch=(res[1][0] - dwref).abs().max().item(),
                              db_naive=(res[0][1] - dbref).abs().max().item(),
                              db_tc=(res[1][1] - dbref).abs().max().item(), fault=ops.read_fault())), flush=True)


def _time(torch, fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def sec_time(flags):
    torch, L, ops = _setup()
    for (B, H, W) in [(512, 20, 20), (128, 20, 20), (64, 80, 80)]:
        for cin in (64, 192, 320):
            x = rand_pnhwc(torch, B, H, W, 320, 1)
            w = torch.randn(64, cin, 3, 3, device="cuda") * 0.02
            wp = ops.pack_fwd(w)
            y = torch.zeros(B, H + 1, W + 1, 64, dtype=torch.bfloat16, device="cuda")
            ms = _time(torch, lambda: ops.conv3x3(x, 0, cin, wp, None, y, 0, 64, act=L.ACT_LRELU, flags=flags))
            flop = 2.0 * B * H * W * 9 * cin * 64
            gy = rand_pnhwc(torch, B, H, W, 64, 2)
            dw = torch.zeros(64, cin, 3, 3, device="cuda")
            ws = torch.empty(ops.wgrad_workspace_bytes(B, H, W, cin, 64), dtype=torch.uint8, device="cuda")
            ms2 = _time(torch, lambda: ops.wgrad3x3(x, 0, cin, gy, 0, 64, dw, None, workspace=ws, flags=flags))
            print(json.dumps(dict(shape=[B, H, W, cin], conv_ms=ms, conv_tflops=flop / ms / 1e9,
                                  wgrad_ms=ms2, wgrad_tflops=flop / ms2 / 1e9, fault=ops.read_fault())), flush=True)


def sec_rate(_):
    torch, L, ops = _setup()
    lib = L.lib()
    cyc = torch.zeros(148, dtype=torch.int64, device="cuda")
    iters = 4096
    for (cg, N, shift, mn, lbo) in [(1, 64, -1, 0, 0), (1, 64, -2, 0, 0), (1, 64, -3, 0, 0), (1, 64, 0, 0, 0), (1, 64, 3, 0, 0), (1, 128, 0, 0, 0), (1, 256, 0, 0, 0),
                                    (2, 64, 0, 0, 0), (2, 128, 0, 0, 0), (2, 256, 0, 0, 0),
                                    (1, 192, 0, 1, 128), (1, 192, 0, 1, 8192), (1, 64, 0, 1, 8192)]:
        for n_ctas in (2, 148):
            cyc.zero_()
            L.check(lib.srb_probe_mma_rate(cg, N, iters, shift, mn, lbo, n_ctas, cyc.data_ptr(),
                                           torch.cuda.current_stream().cuda_stream))
            torch.cuda.synchronize()
            c = cyc[:n_ctas][cyc[:n_ctas] > 0].float()
            M = 256 if cg == 2 else 128
            per = c.mean().item() / iters
            # flops per clk per SM: a cta_group::2 instruction spans two SMs
            fl = 2.0 * M * N * 16 / per / cg
            print(json.dumps(dict(cg=cg, N=N, a_shift=shift, mn_major=mn, b_lbo=lbo, ctas=n_ctas, clk_per_mma=per,
                                  flop_per_clk_sm=fl, frac_of_8192=fl / 8192.0)), flush=True)


SECTIONS = {"conv0": (sec_conv, 0), "conv1": (sec_conv, 1), "wgrad0": (sec_wgrad, 0), "wgrad1": (sec_wgrad, 1),
            "rate": (sec_rate, 0), "time0": (sec_time, 0), "time1": (sec_time, 1)}

if __name__ == "__main__":
    if len(sys.argv) == 2 and sys.argv[1].startswith("--run="):
        name = sys.argv[1][6:]
        fn, arg = SECTIONS[name]
        fn(arg)
        sys.exit(0)
    names = sys.argv[1:] or list(SECTIONS)
    for n in names:
        print(f"=== {n}", flush=True)
        t = time.time()
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), f"--run={n}"], timeout=240,
                               capture_output=True, text=True)
            print(r.stdout[-6000:], flush=True)
            if r.returncode != 0:
                print(f"--- exit {r.returncode}\n{r.stderr[-3000:]}", flush=True)
        except subprocess.TimeoutExpired as e:
            print(f"--- TIMEOUT {n}: {(e.stdout or b'')[-2000:]}", flush=True)
        print(f"=== {n} done in {time.time() - t:.1f}s", flush=True)
