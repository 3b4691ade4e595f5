"""conv3.2 (+ head) forward and backward through the C ABI at the benchmark geometry: timing, and a dump of the results
so that the taps-in-the-GEMM kernels (csrc/conv_out_tc.cu) can be compared with the general warp-MMA path
(SRB_CONV_OUT_GENERAL=1) across two processes and with a float64 torch evaluation in this one.
Usage: python tools/exp_conv_out.py [B] [out.pt]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F
from srb200 import _lib as L
B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
H = W = int(os.environ.get("HW", 80))
dev = torch.device("cuda:0")
lib = L.lib()
g = torch.Generator(device="cpu").manual_seed(3)
act = torch.zeros(B, H + 1, W + 1, 64, dtype=torch.bfloat16, device=dev)
a = torch.randn(B, H, W, 64, generator=g) * 0.5
a = torch.where(a > 0, a, 0.01 * a)
act[:, :H, :W] = a.to(dev).to(torch.bfloat16)
w = (torch.randn(1, 64, 3, 3, generator=g) * 0.05).to(dev)
bias = torch.tensor([0.03], device=dev)
out = torch.empty(B, 1, H, W, device=dev); srs = torch.empty_like(out); z = torch.empty_like(out)
mult, power, thres = 1.0, 1.0, 0.0
st = lambda: torch.cuda.current_stream().cuda_stream
def fwd():
    L.check(lib.srb_conv_out_fwd(act.data_ptr(), 64, 0, B, H, W, 64, w.data_ptr(), bias.data_ptr(), 1, mult, power, thres, 1,
                                 out.data_ptr(), srs.data_ptr(), z.data_ptr(), st()), "fwd")
go = (torch.randn(B, 1, H, W, generator=g) * (torch.rand(B, 1, H, W, generator=g) < 0.3)).to(dev)
gs = torch.zeros_like(go)
gzs = torch.empty(B, 1, H, W, device=dev)
gy = torch.zeros(B, H + 1, W + 1, 64, dtype=torch.bfloat16, device=dev)
dw = torch.zeros(1, 64, 3, 3, device=dev); db = torch.zeros(1, device=dev)
def bwd():
    L.check(lib.srb_conv_out_bwd(z.data_ptr(), go.data_ptr(), gs.data_ptr(), B, 1, H, W, mult, power, thres, 1, w.data_ptr(), 64,
                                 act.data_ptr(), 64, 0, 0.01, gzs.data_ptr(), gy.data_ptr(), 64, 0, dw.data_ptr(), db.data_ptr(),
                                 st()), "bwd")
def t(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it * 1e3
fwd(); bwd(); torch.cuda.synchronize()
# float64 evaluation of the same bf16 operands (batch slice: enough to see a wrong tap or a wrong mask)
nb = min(B, 8)
x64 = act[:nb, :H, :W].double().permute(0, 3, 1, 2)
ref = F.conv2d(x64, w.double(), bias.double(), 1, 1)
e_f = ((out[:nb].double() - ref).abs().max() / ref.abs().max()).item()
gz64 = go[:nb].double()            # power 1, thres 0, training: d out / d z = 1
gin = torch.nn.grad.conv2d_input(x64.shape, w.double(), gz64, padding=1) * torch.where(x64 > 0, 1.0, 0.01)
got = gy[:nb, :H, :W].double().permute(0, 3, 1, 2)
e_g = ((got - gin).abs().max() / gin.abs().max()).item()
pad = max(gy[:, H].abs().max().item(), gy[:, :, W].abs().max().item())
dw64 = torch.nn.grad.conv2d_weight(act[:, :H, :W].double().permute(0, 3, 1, 2), w.shape, go.double(), padding=1)
e_w = ((dw.double() - dw64).abs().max() / dw64.abs().max()).item()
e_b = abs(db.item() - go.double().sum().item()) / abs(go.double().sum().item())
print(f"general_path={os.environ.get('SRB_CONV_OUT_GENERAL', '0')} B={B}: fwd {t(fwd):.1f} us  bwd {t(bwd):.1f} us | "
      f"rel err fwd {e_f:.2e} dgrad {e_g:.2e} (pad rows {pad:.1e}) wgrad {e_w:.2e} bias {e_b:.2e}", flush=True)
