"""Where does a chained launch spend its time?  Chains of five convolutions with the SAME K (cin = 64 k, k = 1..5)
at the bench geometry, CTA-pair kernel (flags bit 22) against the cta_group::1 kernel (bit 23): the slope over k is the
cost of a k-block of MMAs per layer, the intercept is what a layer costs besides them (epilogue, flag waits) plus the
launch's fill / drain.  Also a single-layer launch per K (fill / drain + one layer).  us per launch, TFLOP/s."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
from srb200 import _lib as L, ops
dev = torch.device("cuda:0")
B, H, W, F_ = int(os.environ.get("B", 512)), 20, 20, 64
EXTRA = int(os.environ.get("KFLAGS", "0"), 0)


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
signs = torch.zeros(8, B * (H + 1) * (W + 1), dtype=torch.int32, device=dev)
bias = torch.zeros(F_, device=dev)
rows = []
for k in range(1, 6):
    pack = ops.pack_fwd(torch.randn(F_, k * F_, 3, 3, device=dev) * (1.0 / (9 * k * F_)) ** 0.5)
    for name, fl in (("pair", 0x400000), ("single", 0x800000)):
        for n in (5, 1):
            layers = [ops.conv3x3_args(src, 0, k * F_, pack, bias, outs[i], 0, F_, act=L.ACT_LRELU, slope=0.01,
                                       sign_out=(signs, 0), flags=fl | EXTRA) for i in range(n)]
            flags = torch.zeros(ops.chain_flag_ints(B, H, W, n), dtype=torch.int32, device=dev)

            def run():
                flags.zero_()
                ops.conv3x3_chain(layers, flags)
            us = timeit(run)
            fl_total = sum(f for _, f in layers)
            rows.append((k, name, n, us, fl_total / us / 1e6))
            print(f"K={64 * k:3d} {name:6s} layers={n}  {us:8.1f} us  {fl_total / us / 1e6:7.1f} TFLOP/s", flush=True)
# timing-experiment switches of the pair kernel (flags = dbg << 1): 0x8 no epilogue I/O, 0x20 no flag signal / wait
# (+ no __threadfence), 0x40 no accumulator read-out.  Results are wrong by design; the layers here are independent.
if os.environ.get("VARIANTS", "1") == "1":
    for k in (1, 2, 5):
        pack = ops.pack_fwd(torch.randn(F_, k * F_, 3, 3, device=dev) * (1.0 / (9 * k * F_)) ** 0.5)
        for extra in (0, 0x8, 0x20, 0x28, 0x68):
            layers = [ops.conv3x3_args(src, 0, k * F_, pack, bias, outs[i], 0, F_, act=L.ACT_LRELU, slope=0.01,
                                       sign_out=(signs, 0), flags=0x400000 | extra) for i in range(5)]
            flags = torch.zeros(ops.chain_flag_ints(B, H, W, 5), dtype=torch.int32, device=dev)

            def run():
                flags.zero_()
                ops.conv3x3_chain(layers, flags)
            us = timeit(run)
            print(f"K={64 * k:3d} pair layers=5 dbg-flags={extra:#04x}  {us:8.1f} us", flush=True)
print("fault", ops.read_fault())
