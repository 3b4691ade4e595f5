"""One line of conv3x3 timings (LR 20x20, B=512, cin 64/192/320 -> 64, fwd) for the current SRB_ROW_DBG / SRB_ROW_MT /
SRB_CONV_ROW environment (read once per process).  Results under SRB_ROW_DBG != 0 are numerically wrong on purpose."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
from srb200 import _lib as L, ops
dev = torch.device("cuda:0")
B, H, W = 512, 20, 20
def t(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
x = torch.zeros(B, H + 1, W + 1, 320, dtype=torch.bfloat16, device=dev)
x[:, :H, :W] = torch.randn(B, H, W, 320, device=dev).to(torch.bfloat16)
y = torch.zeros(B, H + 1, W + 1, 64, dtype=torch.bfloat16, device=dev)
out = []
for cin in (64, 192, 320):
    wp = ops.pack_fwd(torch.randn(64, cin, 3, 3, device=dev) * 0.02)
    bias = torch.zeros(64, device=dev)
    ms = t(lambda: ops.conv3x3(x, 0, cin, wp, bias, y, 0, 64, act=L.ACT_LRELU))
    out.append(f"cin{cin}: {ms*1e3:6.1f}us")
env = " ".join(f"{k}={os.environ[k]}" for k in ("SRB_CONV_ROW", "SRB_ROW_DBG", "SRB_ROW_MT") if k in os.environ)
print(f"{env:40s}", " | ".join(out), "fault", ops.read_fault(), flush=True)
