"""clock64 timeline of CTA 0 of the filter-row conv kernel (SRB_ROW_DBG must include 32)."""
import os, sys, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "super-resolution_b200")); sys.path.insert(0, ROOT)
import torch
from srb200 import _lib as L, ops
dev = torch.device("cuda:0")
B, H, W = 512, 20, 20
names = ["entry", "setup", "pdl_wait", "A0 landed", "tile0 issued", "epi t_full0", "epi tmem done", "epi tile0 done",
         "last tile issued", "epi last done", "exit", "ld d2d1", "shfl+xch", "bar", "ld d0"]
x = torch.zeros(B, H + 1, W + 1, 320, dtype=torch.bfloat16, device=dev)
y = torch.zeros(B, H + 1, W + 1, 64, dtype=torch.bfloat16, device=dev)
for cin in (64, 320):
    wp = ops.pack_fwd(torch.randn(64, cin, 3, 3, device=dev) * 0.02)
    bias = torch.zeros(64, device=dev)
    for _ in range(4):
        ops.conv3x3(x, 0, cin, wp, bias, y, 0, 64, act=L.ACT_LRELU)
    torch.cuda.synchronize()
    st = (C.c_longlong * 16)()
    L.check(L.lib().srb_igemm_read_stamps(C.cast(st, C.c_void_p)))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): ops.conv3x3(x, 0, cin, wp, bias, y, 0, 64, act=L.ACT_LRELU)
    e1.record(); torch.cuda.synchronize()
    print(f"DBG={os.environ.get('SRB_ROW_DBG')} cin {cin}: {e0.elapsed_time(e1)/20*1e3:.1f} us |",
          " | ".join(f"{n}: {st[i]-st[0]}" for i, n in enumerate(names)), flush=True)
