"""conv3.2 + output head of the generator (models.py:99,114-118,131-135) and its backward through the C ABI
(srb_conv_out_fwd / srb_conv_out_bwd -> csrc/conv_out_tc.cu, taps folded into the GEMM) against a float64 torch
evaluation.  The kernels multiply bf16 operands (activations, bf16-rounded weights, the bf16-rounded 1-channel gradient -
as the general warp-MMA path did) and accumulate in fp32: against a reference built from the SAME rounded operands the
fp32 outputs agree to 2e-5 of range, the weight gradient to 1e-4, the bf16 input gradient to 5e-3 (its own output rounding);
against the unrounded fp32 weights / gradient the budget is 4e-3 (8e-3 for the bf16 input gradient).  Geometries cover
partial row bands (H not a multiple of 16), partial 16-pixel groups, widths down to the kernel's minimum and the pad
row / column of PNHWC staying zero."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda:0")


def _st():
    return torch.cuda.current_stream().cuda_stream


def _setup(B, H, W, seed, dev):
    g = torch.Generator().manual_seed(seed)
    act = torch.zeros(B, H + 1, W + 1, 64, dtype=torch.bfloat16, device=dev)
    a = torch.randn(B, H, W, 64, generator=g) * 0.5
    a = torch.where(a > 0, a, 0.01 * a)                     # what conv3.0 + LeakyReLU(0.01) leaves
    act[:, :H, :W] = a.to(dev).to(torch.bfloat16)
    w = (torch.randn(1, 64, 3, 3, generator=g) * 0.05).to(dev)
    bias = torch.tensor([0.03], device=dev)
    return g, act, w, bias


@pytest.mark.parametrize("B,H,W", [(3, 80, 80), (5, 24, 24), (2, 20, 20), (2, 17, 19), (1, 9, 8), (4, 33, 12)])
@pytest.mark.parametrize("power,thres,training", [(1.0, 0.0, 1), (0.5, 0.05, 1), (1.0, 0.02, 0)])
def test_conv_out_forward_and_head(B, H, W, power, thres, training):
    dev = _dev()
    from srb200 import _lib as L
    lib = L.lib()
    _, act, w, bias = _setup(B, H, W, 7 * H + W, dev)
    mult = 2.5
    out = torch.empty(B, 1, H, W, device=dev); srs = torch.empty_like(out); z = torch.empty_like(out)
    L.check(lib.srb_conv_out_fwd(act.data_ptr(), 64, 0, B, H, W, 64, w.data_ptr(), bias.data_ptr(), 1, mult, power, thres,
                                 training, out.data_ptr(), srs.data_ptr(), z.data_ptr(), _st()), "conv_out_fwd")
    torch.cuda.synchronize()
    x64 = act[:, :H, :W].double().permute(0, 3, 1, 2)
    zr = F.conv2d(x64, w.double(), bias.double(), 1, 1) / mult
    scale = zr.abs().max().item()
    zb = F.conv2d(x64, w.to(torch.bfloat16).double(), bias.double(), 1, 1) / mult
    assert (z.double() - zb).abs().max().item() <= 2e-5 * scale
    assert (z.double() - zr).abs().max().item() <= 4e-3 * scale
    # the head on the kernel's own z (models.py:114-118,131-135): hardshrink / relu / ** (1 / power)
    zz = z.double()
    lam_p = thres ** power
    s_ref = F.hardshrink(zz if training else zz.clamp(min=0), lam_p)
    o2 = zz if power == 1.0 else zz.clamp(min=0) ** (1.0 / power)
    if not training:
        o2 = o2.clamp(min=0)
    o_ref = F.hardshrink(o2, thres)
    # values within fp32 rounding of a shrink threshold may land on either side of it
    for got, ref, lam, pre in ((srs, s_ref, lam_p, zz), (out, o_ref, thres, o2)):
        away = (pre.abs() - lam).abs() > 1e-5 * max(scale, 1.0)
        assert (got.double() - ref)[away].abs().max().item() <= 1e-5 * max(scale, 1.0)


@pytest.mark.parametrize("B,H,W", [(3, 80, 80), (5, 24, 24), (2, 20, 20), (2, 17, 19), (1, 9, 8), (4, 33, 12)])
def test_conv_out_backward_input_weight_and_bias_gradients(B, H, W):
    dev = _dev()
    from srb200 import _lib as L
    lib = L.lib()
    g, act, w, bias = _setup(B, H, W, 11 * H + W, dev)
    z = torch.randn(B, 1, H, W, generator=g).to(dev)
    go = (torch.randn(B, 1, H, W, generator=g) * (torch.rand(B, 1, H, W, generator=g) < 0.4)).to(dev)
    gs = (torch.randn(B, 1, H, W, generator=g) * 0.3).to(dev)
    gzs = torch.empty(B, 1, H, W, device=dev)
    gy = torch.full((B, H + 1, W + 1, 64), 7.0, dtype=torch.bfloat16, device=dev)
    gy[:, H] = 0
    gy[:, :, W] = 0
    dw = torch.full((1, 64, 3, 3), 5.0, device=dev); db = torch.full((1,), 5.0, device=dev)      # overwritten, not added to
    mult = 2.0
    L.check(lib.srb_conv_out_bwd(z.data_ptr(), go.data_ptr(), gs.data_ptr(), B, 1, H, W, mult, 1.0, 0.0, 1, w.data_ptr(), 64,
                                 act.data_ptr(), 64, 0, 0.01, gzs.data_ptr(), gy.data_ptr(), 64, 0, dw.data_ptr(),
                                 db.data_ptr(), _st()), "conv_out_bwd")
    torch.cuda.synchronize()
    # training head with power 1, thres 0: out = srs = z, z = raw / multiplier
    gz64 = (go.double() + gs.double()) / mult
    assert (gzs.double() - gz64).abs().max().item() <= 1e-6 * gz64.abs().max().item()
    x64 = act[:, :H, :W].double().permute(0, 3, 1, 2)
    mask = torch.where(x64 > 0, 1.0, 0.01)
    gzb, wb = gzs.to(torch.bfloat16).double(), w.to(torch.bfloat16).double()               # what enters the MMAs
    got = gy[:, :H, :W].double().permute(0, 3, 1, 2)
    gin_b = torch.nn.grad.conv2d_input(x64.shape, wb, gzb, padding=1) * mask
    gin = torch.nn.grad.conv2d_input(x64.shape, w.double(), gz64, padding=1) * mask
    assert (got - gin_b).abs().max().item() <= 5e-3 * gin_b.abs().max().item()
    assert (got - gin).abs().max().item() <= 8e-3 * gin.abs().max().item()
    assert gy[:, H].abs().max().item() == 0.0 and gy[:, :, W].abs().max().item() == 0.0          # padding stays zero
    dw_b = torch.nn.grad.conv2d_weight(x64, w.shape, gzb, padding=1)
    dw64 = torch.nn.grad.conv2d_weight(x64, w.shape, gz64, padding=1)
    assert (dw.double() - dw_b).abs().max().item() <= 1e-4 * dw_b.abs().max().item()
    assert (dw.double() - dw64).abs().max().item() <= 4e-3 * dw64.abs().max().item()
    assert abs(db.item() - gz64.sum().item()) <= 1e-5 * gz64.abs().sum().item()


def test_conv_out_is_deterministic_in_its_forward_and_input_gradient():
    dev = _dev()
    from srb200 import _lib as L
    lib = L.lib()
    B, H, W = 4, 40, 40
    g, act, w, bias = _setup(B, H, W, 99, dev)
    outs = []
    for _ in range(2):
        out = torch.empty(B, 1, H, W, device=dev); srs = torch.empty_like(out); z = torch.empty_like(out)
        L.check(lib.srb_conv_out_fwd(act.data_ptr(), 64, 0, B, H, W, 64, w.data_ptr(), bias.data_ptr(), 1, 1.0, 1.0, 0.0, 1,
                                     out.data_ptr(), srs.data_ptr(), z.data_ptr(), _st()), "conv_out_fwd")
        go = torch.ones(B, 1, H, W, device=dev) * 0.25
        gzs = torch.empty(B, 1, H, W, device=dev)
        gy = torch.zeros(B, H + 1, W + 1, 64, dtype=torch.bfloat16, device=dev)
        dw = torch.zeros(1, 64, 3, 3, device=dev); db = torch.zeros(1, device=dev)
        L.check(lib.srb_conv_out_bwd(z.data_ptr(), go.data_ptr(), None, B, 1, H, W, 1.0, 1.0, 0.0, 1, w.data_ptr(), 64,
                                     act.data_ptr(), 64, 0, 0.01, gzs.data_ptr(), gy.data_ptr(), 64, 0, dw.data_ptr(),
                                     db.data_ptr(), _st()), "conv_out_bwd")
        torch.cuda.synchronize()
        outs.append((out.clone(), gy.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
