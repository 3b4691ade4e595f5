"""Image-end layer of the discriminator through the C ABI (srb_dl1_dgrad / srb_dl1_fwd): timing at the benchmark geometry and
the error against a float64 torch evaluation of the same bf16 operands.  SRB_DL1_GENERAL=1 selects the CUDA-core kernels.
Usage: python tools/exp_dl1.py [B] [C]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F
from srb200 import _lib as L
B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
C = int(sys.argv[2]) if len(sys.argv) > 2 else 16
H = W = int(os.environ.get("HW", 80))
dev = torch.device("cuda:0")
lib = L.lib()
g = torch.Generator(device="cpu").manual_seed(5)
st = lambda: torch.cuda.current_stream().cuda_stream
def t(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it * 1e3
w = (torch.randn(C, 1, 3, 3, generator=g) * 0.3).to(dev)
bias = (torch.randn(C, generator=g) * 0.1).to(dev)
gz = (torch.randn(B, H, W, C, generator=g) * 0.1).to(dev).to(torch.bfloat16)
gx = torch.empty(B, 1, H, W, device=dev)
def dgrad():
    L.check(lib.srb_dl1_dgrad(gz.data_ptr(), w.data_ptr(), gx.data_ptr(), C, B, H, W, st()), "dl1_dgrad")
dgrad(); torch.cuda.synchronize()
nb = min(B, 6)
ref = torch.nn.grad.conv2d_input((nb, 1, H, W), w.double(), gz[:nb].double().permute(0, 3, 1, 2), padding=1)
e_d = max(((gx[:nb].double() - ref).abs().max() / ref.abs().max()).item(),
          ((gx[B - 1:].double() - torch.nn.grad.conv2d_input((1, 1, H, W), w.double(), gz[B - 1:].double().permute(0, 3, 1, 2), padding=1)).abs().max() / ref.abs().max()).item())
img = (torch.randn(B, 1, H, W, generator=g) * (torch.rand(B, 1, H, W, generator=g) < 0.05) * 20).to(dev)
out = torch.empty(B, H, W, C, dtype=torch.bfloat16, device=dev)
refa = (torch.randn(B, H, W, C, generator=g)).to(dev).to(torch.bfloat16)
def fwd(act=1):
    L.check(lib.srb_dl1_fwd(img.data_ptr(), w.data_ptr(), bias.data_ptr() if act == 1 else None, out.data_ptr(),
                            refa.data_ptr() if act == 2 else None, C, B, H, W, act, 0.2, st()), "dl1_fwd")
errs = []
for act in (1, 2):
    fwd(act); torch.cuda.synchronize()
    z = F.conv2d(img[:nb].double(), w.double(), bias.double() if act == 1 else None, 1, 1)
    r = F.leaky_relu(z, 0.2) if act == 1 else z * torch.where(refa[:nb].double().permute(0, 3, 1, 2) > 0, 1.0, 0.2)
    errs.append(((out[:nb].double().permute(0, 3, 1, 2) - r).abs().max() / r.abs().max()).item())
print(f"general={os.environ.get('SRB_DL1_GENERAL', '0')} B={B} C={C} HW={H}: dgrad {t(dgrad):.1f} us (rel err {e_d:.2e}) | "
      f"fwd lrelu {t(lambda: fwd(1)):.1f} us (err {errs[0]:.2e})  fwd mask {t(lambda: fwd(2)):.1f} us (err {errs[1]:.2e})", flush=True)
