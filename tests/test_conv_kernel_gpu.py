"""GPU parity of the tcgen05 3x3 convolution kernel in isolation (C ABI srb_conv3x3) against torch fp32 conv2d on the
same bf16-rounded operands: every epilogue option (bias, alpha, two residual streams, LeakyReLU, LeakyReLU' mask,
PixelShuffle store, channel offsets) and geometries that hit the tile-schedule edges of the filter-row kernel
(sub-tiles of 126 rows: partial last sub-tile, one sub-tile per CTA, odd pairs, ranges that straddle an n-tile).
Tolerance: bf16 output rounding, 6e-3 of the output range."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
TOL = 6e-3
KERNELS = {"one_tap": 0x20000, "filter_row": 0x10000}      # srb_conv3x3_args.flags bits 17 / 16 (conv_igemm.cu)


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda:0")


def _rand_pnhwc(B, H, W, C, dev, g, scale=1.0):
    from srb200 import ops
    x = torch.randn(B, C, H, W, generator=g, device=dev) * scale
    return ops.to_pnhwc(x), x.to(torch.bfloat16).float()


@pytest.mark.parametrize("B,H,W,cin,cout", [
    (1, 20, 20, 64, 64),        # 4 sub-tiles, one per CTA
    (3, 20, 20, 128, 64),       # partial last sub-tile
    (40, 20, 20, 320, 64),      # 140 sub-tiles < 148 CTAs
    (90, 20, 20, 192, 64),      # 315 sub-tiles: pairs + odd singles
    (130, 20, 20, 64, 64),      # > 2 per CTA
    (2, 13, 17, 64, 64),        # odd width / height
    (2, 80, 80, 64, 64),        # HR geometry (two TMA boxes per window)
    (1, 160, 160, 64, 64),      # wide image: one sub-tile per tile (window would not fit twice)
    (5, 20, 20, 64, 256),       # four n-tiles, ranges straddle n-tile boundaries
])
@pytest.mark.parametrize("kernel", list(KERNELS))
def test_conv3x3_plain_vs_torch(B, H, W, cin, cout, kernel):
    dev = _dev()
    from srb200 import _lib as L, ops
    g = torch.Generator(device="cuda").manual_seed(B * 1000 + H + cin)
    xb, xr = _rand_pnhwc(B, H, W, cin, dev, g)
    w = torch.randn(cout, cin, 3, 3, generator=g, device=dev) * (2.0 / (9 * cin)) ** 0.5
    bias = torch.randn(cout, generator=g, device=dev) * 0.1
    y = ops.pnhwc_zeros(B, H, W, cout, dev)
    ops.conv3x3(xb, 0, cin, ops.pack_fwd(w), bias, y, 0, cout, act=L.ACT_LRELU, slope=0.01, flags=KERNELS[kernel])
    ref = F.leaky_relu(F.conv2d(xr, w.to(torch.bfloat16).float(), bias, padding=1), 0.01)
    got = ops.from_pnhwc(y)
    assert (got - ref).abs().max().item() <= TOL * ref.abs().max().item()
    # padding row / column stay zero (kernels only store valid pixels)
    assert y[:, H].abs().max().item() == 0 and y[:, :, W].abs().max().item() == 0
    assert ops.read_fault() == 0


@pytest.mark.parametrize("kernel", list(KERNELS))
def test_conv3x3_epilogue_options_vs_torch(kernel):
    fl = KERNELS[kernel]
    dev = _dev()
    from srb200 import _lib as L, ops
    g = torch.Generator(device="cuda").manual_seed(11)
    B, H, W = 7, 20, 20
    growth = ops.pnhwc_zeros(B, H, W, 320, dev)
    xs = torch.randn(B, 320, H, W, generator=g, device=dev)
    growth[:, :H, :W] = xs.permute(0, 2, 3, 1).to(torch.bfloat16)
    xr = xs.to(torch.bfloat16).float()
    w = torch.randn(64, 192, 3, 3, generator=g, device=dev) * 0.03
    bias = torch.randn(64, generator=g, device=dev) * 0.1
    wr = w.to(torch.bfloat16).float()
    r2b, r2 = _rand_pnhwc(B, H, W, 64, dev, g)
    conv = F.conv2d(xr[:, 64:256], wr, bias, padding=1)
    # alpha + two residual streams, input at channel offset 64, output into a 128-wide buffer at offset 64
    y = ops.pnhwc_zeros(B, H, W, 128, dev)
    ops.conv3x3(growth, 64, 192, ops.pack_fwd(w), bias, y, 64, 64, alpha=0.2, res1=(growth, 0, 0.5), res2=(r2b, 0, 1.0),
                flags=fl)
    ref = 0.2 * conv + 0.5 * xr[:, :64] + r2
    got = ops.from_pnhwc(y, 64, 64)
    assert (got - ref).abs().max().item() <= TOL * ref.abs().max().item()
    assert y[..., :64].abs().max().item() == 0
    # LeakyReLU' mask taken from another buffer's channel slice (input-gradient convs), written in place into a slot
    y2 = ops.pnhwc_zeros(B, H, W, 320, dev)
    ops.conv3x3(growth, 64, 192, ops.pack_fwd(w), None, y2, 256, 64, act=L.ACT_LRELU_MASK, slope=0.01,
                mask=(growth, 128), flags=fl)
    ref2 = F.conv2d(xr[:, 64:256], wr, None, padding=1) * torch.where(xr[:, 128:192] > 0, 1.0, 0.01)
    got2 = ops.from_pnhwc(y2, 256, 64)
    assert (got2 - ref2).abs().max().item() <= TOL * ref2.abs().max().item()
    assert ops.read_fault() == 0


@pytest.mark.parametrize("B,H,W", [(3, 20, 20), (2, 40, 40)])
@pytest.mark.parametrize("kernel", list(KERNELS))
def test_conv3x3_pixelshuffle_store_vs_torch(B, H, W, kernel):
    dev = _dev()
    from srb200 import _lib as L, ops
    g = torch.Generator(device="cuda").manual_seed(5 + H)
    xb, xr = _rand_pnhwc(B, H, W, 64, dev, g)
    w = torch.randn(256, 64, 3, 3, generator=g, device=dev) * 0.05
    bias = torch.randn(256, generator=g, device=dev) * 0.1
    idx = torch.arange(256, device=dev)
    bias_p = bias[4 * (idx % 64) + idx // 64].contiguous()          # same re-ordering as the weight pack (shuffle=True)
    y = ops.pnhwc_zeros(B, 2 * H, 2 * W, 64, dev)
    ops.conv3x3(xb, 0, 64, ops.pack_fwd(w, shuffle=True), bias_p, y, 0, 256, act=L.ACT_LRELU, slope=0.01, shuffle=True,
                flags=KERNELS[kernel])
    ref = F.pixel_shuffle(F.leaky_relu(F.conv2d(xr, w.to(torch.bfloat16).float(), bias, padding=1), 0.01), 2)
    got = ops.from_pnhwc(y)
    assert (got - ref).abs().max().item() <= TOL * ref.abs().max().item()
    assert ops.read_fault() == 0
