"""A/B timing of the conv3x3 kernels at bench geometry (run once per SRB_CONV_ROW=0/1; the switch is read once per
process).  Prints us and TFLOP/s per launch for the dense-block shapes (LR 20x20, Cin 64..320 -> 64), the input-gradient
form (mask epilogue) and the HR layers."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
from srb200 import _lib as L, ops
dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
def t(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it
print("SRB_CONV_ROW =", os.environ.get("SRB_CONV_ROW", "(default 1)"), "B =", B)
tot = 0.0
for (b, H, W, cins, cout, shuffle) in [(B, 20, 20, (64, 128, 192, 256, 320), 64, False), (B, 20, 20, (64,), 256, True),
                                        (B, 40, 40, (64,), 256, True), (B, 80, 80, (64,), 64, False)]:
    x = torch.zeros(b, H + 1, W + 1, 320, dtype=torch.bfloat16, device=dev)
    x[:, :H, :W] = torch.randn(b, H, W, 320, device=dev).to(torch.bfloat16)
    y = torch.zeros(b, (2 * H if shuffle else H) + 1, (2 * W if shuffle else W) + 1, 64, dtype=torch.bfloat16, device=dev)
    for cin in cins:
        wp = ops.pack_fwd(torch.randn(cout, cin, 3, 3, device=dev) * 0.02, shuffle=shuffle)
        bias = torch.zeros(cout, device=dev)
        fl = 2.0 * b * H * W * 9 * cin * cout
        ms = t(lambda: ops.conv3x3(x, 0, cin, wp, bias, y, 0, cout, act=L.ACT_LRELU, shuffle=shuffle))
        line = f"{H}x{W} cin {cin:3d} cout {cout:3d}: fwd {ms*1e3:7.1f} us {fl/ms/1e9:6.0f} TF"
        if not shuffle and cout == 64 and H == 20:
            ms2 = t(lambda: ops.conv3x3(x, 0, cin, wp, None, y, 0, cout, act=L.ACT_LRELU_MASK, mask=(x, 256)))
            line += f" | dgrad(mask) {ms2*1e3:7.1f} us {fl/ms2/1e9:6.0f} TF"
            tot += ms + ms2
        print(line, flush=True)
print(f"dense block fwd+dgrad (5+5 launches): {tot*1e3:.0f} us; x30 blocks = {tot*30:.2f} ms")
print("fault", ops.read_fault())
