"""srb_wgrad3x3_static: the allocation-free variant of the 3x3 weight gradient (split-K scratch in a static device array
of the library, include/srb200.h).  Called through the C ABI and compared with the existing entry point srb_wgrad3x3 on
the same operands, on two calls in a row (the second on a different problem, so the array still holds the first call's
partial tiles), plus a torch fp32 weight gradient of the same bf16-rounded operands.

The file sorts last on purpose: a failure here must not hide the other tests.
Tolerance: the two entry points run the same kernels with the same split plan -> 1e-6 of the gradient's range; against
torch fp32: 2e-3 of the range (fp32 accumulation of bf16 products in a different order)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
SAME_TOL, TORCH_TOL = 1e-6, 2e-3


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda:0")


def _args(x, cin, gy, cout, dw, db, *, accumulate=False, alpha=1.0):
    from srb200 import _lib as L
    a = L.Wgrad3x3Args()
    B, Hp, Wp, Cx = x.shape
    a.batch, a.H, a.W = B, Hp - 1, Wp - 1
    a.x, a.x_cstride, a.x_coff, a.cin = x.data_ptr(), Cx, 0, cin
    a.gy, a.gy_cstride, a.gy_coff, a.cout = gy.data_ptr(), gy.shape[-1], 0, cout
    a.dw, a.dbias, a.alpha = dw.data_ptr(), db.data_ptr() if db is not None else None, alpha
    a.accumulate, a.unshuffle = int(accumulate), 0
    a.impl, a.flags = L.IMPL_TCGEN05, 0
    return a


def _problem(B, H, W, cin, cout, seed, dev):
    from srb200 import ops
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(B, cin, H, W, generator=g, device=dev)
    gy = torch.randn(B, cout, H, W, generator=g, device=dev) * 0.1
    return ops.to_pnhwc(x), ops.to_pnhwc(gy), x.to(torch.bfloat16).float(), gy.to(torch.bfloat16).float()


def _run_both(B, H, W, cin, cout, seed, dev, *, accumulate=False, alpha=1.0):
    from srb200 import _lib as L
    lib = L.lib()
    st = torch.cuda.current_stream().cuda_stream
    xp, gp, xf, gf = _problem(B, H, W, cin, cout, seed, dev)
    base = torch.full((cout, cin, 3, 3), 0.25, device=dev) if accumulate else None
    outs = []
    # the existing entry point, caller-owned workspace
    need = int(lib.srb_wgrad3x3_workspace_bytes(B, H, W, cin, cout))
    assert need <= int(lib.srb_wgrad3x3_static_max_bytes())
    ws = torch.empty(need, dtype=torch.uint8, device=dev)
    dw = base.clone() if accumulate else torch.full((cout, cin, 3, 3), float("nan"), device=dev)
    db = torch.zeros(cout, device=dev) if accumulate else torch.full((cout,), float("nan"), device=dev)
    a = _args(xp, cin, gp, cout, dw, db, accumulate=accumulate, alpha=alpha)
    a.workspace, a.workspace_bytes = ws.data_ptr(), need
    L.check(lib.srb_wgrad3x3(C.byref(a), st), "wgrad3x3")
    outs.append((dw, db))
    # the allocation-free one: no workspace given
    dw2 = base.clone() if accumulate else torch.full((cout, cin, 3, 3), float("nan"), device=dev)
    db2 = torch.zeros(cout, device=dev) if accumulate else torch.full((cout,), float("nan"), device=dev)
    a2 = _args(xp, cin, gp, cout, dw2, db2, accumulate=accumulate, alpha=alpha)
    a2.workspace, a2.workspace_bytes = None, 0
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    L.check(lib.srb_wgrad3x3_static(C.byref(a2), st), "wgrad3x3_static")
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    outs.append((dw2, db2))
    ref_w = torch.nn.grad.conv2d_weight(xf, (cout, cin, 3, 3), gf, padding=1) * alpha
    ref_b = gf.sum((0, 2, 3)) * alpha
    if accumulate:
        ref_w = ref_w + 0.25
    return outs, (ref_w, ref_b), (free0, free1)


def _check(outs, ref):
    (dw, db), (dw2, db2) = outs
    assert torch.isfinite(dw2).all() and torch.isfinite(db2).all()
    sw, sb = ref[0].abs().max().item(), ref[1].abs().max().item()
    assert (dw2 - dw).abs().max().item() <= SAME_TOL * sw, (dw2 - dw).abs().max().item()
    assert (db2 - db).abs().max().item() <= SAME_TOL * sb
    assert (dw2 - ref[0]).abs().max().item() <= TORCH_TOL * sw, ((dw2 - ref[0]).abs().max().item(), sw)
    assert (db2 - ref[1]).abs().max().item() <= TORCH_TOL * sb


def test_static_scratch_wgrad_matches_the_workspace_entry_point_twice_in_a_row():
    dev = _dev()
    from srb200 import ops
    # call 1: a dense block's widest convolution (cin 320): 18 K splits, 13.3 MB of the 16 MiB array
    outs, ref, _ = _run_both(48, 20, 20, 320, 64, 11, dev)
    _check(outs, ref)
    # call 2, straight after, on a smaller problem with accumulate + alpha: the array still holds call 1's tiles
    outs, ref, (free0, free1) = _run_both(7, 13, 17, 64, 128, 12, dev, accumulate=True, alpha=0.5)
    _check(outs, ref)
    assert free1 == free0, "the allocation-free entry point changed the device's free memory"
    # call 3: the first problem again gives the first answer again
    outs3, ref3, (free0, free1) = _run_both(48, 20, 20, 320, 64, 11, dev)
    _check(outs3, ref3)
    assert free1 == free0
    assert ops.read_fault() == 0


def test_static_scratch_wgrad_rejects_problems_beyond_the_array():
    dev = _dev()
    from srb200 import _lib as L
    lib = L.lib()
    cin, cout = 1024, 512
    assert int(lib.srb_wgrad3x3_workspace_bytes(2, 20, 20, cin, cout)) > int(lib.srb_wgrad3x3_static_max_bytes())
    x = torch.zeros(2, 21, 21, cin, dtype=torch.bfloat16, device=dev)
    gy = torch.zeros(2, 21, 21, cout, dtype=torch.bfloat16, device=dev)
    dw = torch.full((cout, cin, 3, 3), 7.0, device=dev)
    a = _args(x, cin, gy, cout, dw, None)
    rc = lib.srb_wgrad3x3_static(C.byref(a), torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.srb_last_error().decode()
    assert "wgrad3x3_static" in msg and str(int(lib.srb_wgrad3x3_static_max_bytes())) in msg, msg
    torch.cuda.synchronize()
    assert (dw == 7.0).all()            # nothing was launched
