"""Timing experiments on the grouped tcgen05 weight-gradient kernel at the benchmark's dense-block shape."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
from srb200 import _lib as L, ops
dev = torch.device("cuda:0")
def t(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
B, H, W, F = int(os.environ.get("B", 512)), 20, 20, 64
x = torch.zeros(B, H + 1, W + 1, 5 * F, dtype=torch.bfloat16, device=dev)
x[:, :H, :W] = torch.randn(B, H, W, 5 * F, device=dev).to(torch.bfloat16)
g = torch.zeros(B, H + 1, W + 1, 5 * F, dtype=torch.bfloat16, device=dev)
g[:, :H, :W] = torch.randn(B, H, W, 5 * F, device=dev).to(torch.bfloat16)
def run(ks, bias):
    probs, fl = [], 0.0
    cins = [(k + 1) * F for k in ks]
    for k in ks:
        cin = (k + 1) * F
        dw = torch.zeros(F, cin, 3, 3, device=dev); db = torch.zeros(F, device=dev) if bias else None
        probs.append((0, cin, k * F, F, dw, db, 1.0))
        fl += 2.0 * B * H * W * 9 * cin * F
    nbytes = ops.wgrad_group_workspace_bytes(B, H, W, cins, [F] * len(ks))
    ws = torch.empty(nbytes // 4 + 16, device=dev)
    ms = t(lambda: ops.wgrad3x3_group(x, g, probs, ws))
    return ms, fl
for ks, bias in [((0, 1, 2, 3, 4), True), ((0, 1, 2, 3, 4), False), ((4,), False), ((4,), True), ((0,), False), ((0,), True)]:
    ms, fl = run(ks, bias)
    print(f"B={B} convs {ks} bias={bias}: {ms*1e3:.1f} us  {fl/ms/1e9:.0f} TF (useful)", flush=True)
