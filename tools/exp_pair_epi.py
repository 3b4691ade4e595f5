"""Which part of the conv epilogue costs the time on the K = 64 layers?  Chains of five independent K = 64 (and 128)
convolutions on the CTA-pair kernel with parts of the epilogue left out through the PUBLIC arguments (no bias, no sign
words, no activation).  us per launch."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
from srb200 import _lib as L, ops
dev = torch.device("cuda:0")
B, H, W, F_ = int(os.environ.get("B", 512)), 20, 20, 64


def timeit(fn, it=30):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it * 1e3


src = ops.pnhwc_zeros(B, H, W, 5 * F_, dev)
src[:, :H, :W] = (torch.randn(B, H, W, 5 * F_, device=dev) * 0.5).to(torch.bfloat16)
outs = [ops.pnhwc_zeros(B, H, W, F_, dev) for _ in range(5)]
wide = ops.pnhwc_zeros(B, H, W, 5 * F_, dev)
signs = torch.zeros(8, B * (H + 1) * (W + 1), dtype=torch.int32, device=dev)
bias = torch.zeros(F_, device=dev)
for k in (1, 2):
    pack = ops.pack_fwd(torch.randn(F_, k * F_, 3, 3, device=dev) * (1.0 / (9 * k * F_)) ** 0.5)
    for name, kw in (("full", dict(bias=bias, sign=True, act=True)), ("no bias", dict(bias=None, sign=True, act=True)),
                     ("no sign", dict(bias=bias, sign=False, act=True)),
                     ("no bias, no sign", dict(bias=None, sign=False, act=True)),
                     ("no bias, no sign, no act", dict(bias=None, sign=False, act=False)),
                     ("full, 320-wide output rows", dict(bias=bias, sign=True, act=True, wide=True))):
        layers = []
        for i in range(5):
            y, yc = (wide, i * F_) if kw.get("wide") else (outs[i], 0)
            layers.append(ops.conv3x3_args(src, 0, k * F_, pack, kw["bias"], y, yc, F_,
                                           act=L.ACT_LRELU if kw["act"] else L.ACT_NONE, slope=0.01,
                                           sign_out=(signs, 0) if kw["sign"] else None, flags=0x400000))
        flags = torch.zeros(ops.chain_flag_ints(B, H, W, 5), dtype=torch.int32, device=dev)

        def run():
            flags.zero_()
            ops.conv3x3_chain(layers, flags)
        print(f"K={64 * k:3d} {name:28s} {timeit(run):8.1f} us", flush=True)
print("fault", ops.read_fault())
