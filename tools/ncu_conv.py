"""Small ncu target: a handful of launches of the generator's hot kernels at bench scale (B=256, LR 20x20):
conv3x3 implicit GEMM (cin 64 / 192 / 320 -> 64) and the weight-gradient kernel.  Usage: python tools/ncu_conv.py [B]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
from srb200 import _lib as L, ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
H = W = 20
dev = torch.device("cuda:0")
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.zeros(B, H + 1, W + 1, 320, dtype=torch.bfloat16, device=dev)
x[:, :H, :W] = torch.randn(B, H, W, 320, generator=g, device=dev).to(torch.bfloat16)
gy = torch.zeros(B, H + 1, W + 1, 64, dtype=torch.bfloat16, device=dev)
gy[:, :H, :W] = (torch.randn(B, H, W, 64, generator=g, device=dev) * 0.1).to(torch.bfloat16)
y = torch.zeros(B, H + 1, W + 1, 64, dtype=torch.bfloat16, device=dev)
for rep in range(2):
    for cin in (64, 192, 320):
        w = torch.randn(64, cin, 3, 3, device=dev) * 0.02
        wp = ops.pack_fwd(w)
        bias = torch.zeros(64, device=dev)
        ops.conv3x3(x, 0, cin, wp, bias, y, 0, 64, act=L.ACT_LRELU, slope=0.01)
        dw = torch.zeros(64, cin, 3, 3, device=dev)
        db = torch.zeros(64, device=dev)
        ops.wgrad3x3(x, 0, cin, gy, 0, 64, dw, db)
torch.cuda.synchronize()
print("ok", ops.read_fault())
