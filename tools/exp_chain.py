"""Chained vs per-convolution launches of a DenseResidualBlock's ten trunk convolutions (5 forward + 5 input
gradient) at the bench geometry.  Prints us per block and TFLOP/s; results are not checked here (tests do that)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import ctypes as C
import torch
from srb200 import _lib as L, ops
dev = torch.device("cuda:0")
B, H, W, F_ = int(os.environ.get("B", 512)), 20, 20, 64
NBLK = 6


def timeit(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it


bufs = [ops.pnhwc_zeros(B, H, W, 5 * F_, dev) for _ in range(NBLK + 1)]
for b in bufs:
    b[:, :H, :W] = (torch.randn(B, H, W, 5 * F_, device=dev) * 0.5).to(torch.bfloat16)
packs = [ops.pack_fwd(torch.randn(F_, k * F_, 3, 3, device=dev) * (1.0 / (9 * k * F_)) ** 0.5) for k in range(1, 6)]
bias = [torch.zeros(F_, device=dev) for _ in range(5)]
st = lambda: torch.cuda.current_stream().cuda_stream


SIGN = os.environ.get("SIGN", "0")      # 0: no sign words; row: [pixel][8]; plane: [8][pixel]
signs = None
if SIGN == "row":
    signs = [torch.zeros(B, H + 1, W + 1, 8, dtype=torch.int32, device=dev) for _ in range(NBLK)]
elif SIGN == "plane":
    signs = [torch.zeros(8, B * (H + 1) * (W + 1), dtype=torch.int32, device=dev) for _ in range(NBLK)]


def fwd_layers(i):
    cur, nxt = bufs[i], bufs[i + 1]
    ls = [ops.conv3x3_args(cur, 0, k * F_, packs[k - 1], bias[k - 1], cur, k * F_, F_, act=L.ACT_LRELU, slope=0.01,
                           sign_out=(signs[i], 2 * (k - 1)) if signs is not None else None)
          for k in range(1, 5)]
    ls.append(ops.conv3x3_args(cur, 0, 5 * F_, packs[4], bias[4], nxt, 0, F_, alpha=0.2, res1=(cur, 0, 1.0)))
    return ls


def bwd_layers(i):
    G, nxt, gb = bufs[i], bufs[i + 1], bufs[(i + 2) % (NBLK + 1)]
    ls = [ops.conv3x3_args(G, 0, (5 - jj) * F_, packs[4 - jj], None, G, (5 - jj) * F_, F_, act=L.ACT_LRELU_MASK,
                           slope=0.01, mask=(gb, jj * F_)) for jj in (4, 3, 2, 1)]
    ls.append(ops.conv3x3_args(G, 0, 5 * F_, packs[4], None, nxt, 0, F_, res1=(G, 0, 1.0)))
    return ls


flops_blk = sum(f for _, f in fwd_layers(0))
for name, mk in (("fwd", fwd_layers), ("dgrad", bwd_layers)):
    layers = [mk(i) for i in range(NBLK)]

    def seq():
        for ls in layers:
            for a, _ in ls:
                L.check(L.lib().srb_conv3x3(C.byref(a), st()))

    def chained(per):
        flat = [l for ls in layers for l in ls]
        groups = [flat[i:i + per] for i in range(0, len(flat), per)]
        fl = [torch.zeros(ops.chain_flag_ints(B, H, W, len(g)), dtype=torch.int32, device=dev) for g in groups]

        def run():
            for f in fl:
                f.zero_()
            for g, f in zip(groups, fl):
                ops.conv3x3_chain(g, f)
        return run

    ms = timeit(seq)
    print(f"{name}: per-conv launches {ms * 1e3 / NBLK:7.1f} us/block  {flops_blk * NBLK / ms / 1e9:6.0f} TF", flush=True)
    for per in (5, 10, 15):
        if name == "dgrad" and per > 5:
            continue
        ms = timeit(chained(per))
        print(f"{name}: chain of {per:2d}        {ms * 1e3 / NBLK:7.1f} us/block  {flops_blk * NBLK / ms / 1e9:6.0f} TF", flush=True)
print("fault", ops.read_fault())
