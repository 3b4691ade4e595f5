"""Timing experiments on the conv3x3 kernel (debug flags make results numerically wrong on purpose)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
from srb200 import _lib as L, ops
dev = torch.device("cuda:0")
def t(fn, it=30):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
for (B, H, W) in [(512, 20, 20), (64, 80, 80)]:
    x = torch.zeros(B, H + 1, W + 1, 320, dtype=torch.bfloat16, device=dev)
    x[:, :H, :W] = torch.randn(B, H, W, 320, device=dev).to(torch.bfloat16)
    y = torch.zeros(B, H + 1, W + 1, 64, dtype=torch.bfloat16, device=dev)
    for cin in (64, 320):
        wp = ops.pack_fwd(torch.randn(64, cin, 3, 3, device=dev) * 0.02)
        fl = 2.0 * B * H * W * 9 * cin * 64
        out = []
        for dbg in (0, 1, 2, 3, 4, 7):
            ms = t(lambda: ops.conv3x3(x, 0, cin, wp, None, y, 0, 64, act=L.ACT_LRELU, flags=dbg << 1))
            out.append(f"dbg{dbg}: {ms*1e3:.1f}us {fl/ms/1e9:.0f}TF")
        print(B, H, cin, " | ".join(out), flush=True)
print("fault", ops.read_fault())
# pipeline stamps of CTA 0 (flags bit 4 -> dbg 8)
import ctypes as C
for (B, H, W, cin) in [(256, 20, 20, 64), (256, 20, 20, 320), (512, 20, 20, 192)]:
    x = torch.zeros(B, H + 1, W + 1, 320, dtype=torch.bfloat16, device=dev)
    y = torch.zeros(B, H + 1, W + 1, 64, dtype=torch.bfloat16, device=dev)
    wp = ops.pack_fwd(torch.randn(64, cin, 3, 3, device=dev) * 0.02)
    for _ in range(3):
        ops.conv3x3(x, 0, cin, wp, None, y, 0, 64, act=L.ACT_LRELU, flags=8 << 1)
    torch.cuda.synchronize()
    st = (C.c_longlong * 16)()
    L.check(L.lib().srb_igemm_read_stamps(C.cast(st, C.c_void_p)))
    t0 = st[0]
    names = ["entry", "setup done", "first A landed", "tile0 MMAs issued", "tile0 epilogue done", "tile1 A ready", "tile1 MMAs issued", "tile1 epilogue done", "exit"]
    print(B, cin, " | ".join(f"{n}: {st[i]-t0}" for i, n in enumerate(names)))
    ms = t(lambda: ops.conv3x3(x, 0, cin, wp, None, y, 0, 64, act=L.ACT_LRELU))
    print("   kernel", round(ms * 1e3, 1), "us")
