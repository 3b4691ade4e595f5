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
# srb_conv3x3_args.flags (conv_igemm.cu): bit 16 filter-row kernel, bit 17 one-tap kernel, bit 18 / 19 forbid / force its
# 512-row tiles (where the geometry fits them; otherwise the 256- or 128-row tiles run)
# bit 20 / 21 force / forbid the 2+1-tap kernel (conv_chain21.cu)
# bit 22 / 23 force / forbid the CTA-pair kernel (conv_chain2.cu, cta_group::2)
KERNELS = {"one_tap": 0x20000 | 0x40000 | 0x200000 | 0x800000, "one_tap_512": 0x20000 | 0x80000 | 0x200000 | 0x800000,
           "cta_pair": 0x400000 | 0x40000 | 0x200000, "cta_pair_512": 0x400000 | 0x80000 | 0x200000,
           "filter_row": 0x10000 | 0x200000 | 0x800000, "taps_2_1": 0x100000 | 0x800000}


def _experimental_build():
    """the filter-row / 2+1-tap kernels exist only in SRB_EXPERIMENTAL builds of libsrb200 (csrc/srb200_all.cu)"""
    try:
        from srb200 import _lib as L
        return bool(L.lib().srb_has_experimental())
    except Exception:
        return False


EXPERIMENTAL = _experimental_build()
if not EXPERIMENTAL:
    KERNELS = {k: v for k, v in KERNELS.items() if k in ("one_tap", "one_tap_512", "cta_pair", "cta_pair_512")}


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


# ------------------------------------------------------------------ chained launches (csrc/conv_chain.cu)
def _dense_block_layers(ops, L, buf, dst, packs, biases, F_=64, flags=0):
    """the five convolutions of a DenseResidualBlock (models.py:34-41) over a growth buffer, as conv3x3_args"""
    layers = []
    for k in range(1, 5):
        layers.append(ops.conv3x3_args(buf, 0, k * F_, packs[k - 1], biases[k - 1], buf, k * F_, F_, act=L.ACT_LRELU,
                                       slope=0.01, flags=flags))
    layers.append(ops.conv3x3_args(buf, 0, 5 * F_, packs[4], biases[4], dst, 0, F_, alpha=0.2, res1=(buf, 0, 1.0),
                                   flags=flags))
    return layers


@pytest.mark.parametrize("B,H,W,nblocks", [
    (3, 20, 20, 1),        # fewer tiles than CTAs, MT = 1
    (130, 20, 20, 1),      # several tiles per CTA, MT = 2
    (512, 20, 20, 3),      # bench geometry, one RRDB = 15 layers in one launch
    (2, 80, 80, 2),        # HR geometry: halo of 82 rows, two TMA boxes per window
    (5, 13, 17, 3),        # odd sizes
])
@pytest.mark.parametrize("kflags", [0x800000, 0x400000, 0x100000] if EXPERIMENTAL else [0x800000, 0x400000],
                         ids=["one_tap", "cta_pair", "taps_2_1"] if EXPERIMENTAL else ["one_tap", "cta_pair"])
def test_conv3x3_chain_is_bitwise_the_sequential_launches(B, H, W, nblocks, kflags):
    """One chained launch over `nblocks` dense blocks == the same convolutions launched one by one (same kernels, same
    accumulation order: bit-exact), and the sequential result matches torch fp32 on bf16-rounded operands."""
    dev = _dev()
    from srb200 import _lib as L, ops
    g = torch.Generator(device="cuda").manual_seed(B + H + 7 * nblocks)
    F_ = 64
    x0 = torch.randn(B, F_, H, W, generator=g, device=dev)
    ws = [[torch.randn(F_, k * F_, 3, 3, generator=g, device=dev) * (1.0 / (9 * k * F_)) ** 0.5 for k in range(1, 6)]
          for _ in range(nblocks)]
    bs = [[torch.randn(F_, generator=g, device=dev) * 0.05 for _ in range(5)] for _ in range(nblocks)]
    packs = [[ops.pack_fwd(w) for w in wl] for wl in ws]

    def run(chained):
        bufs = [ops.to_pnhwc(x0, 5 * F_) if i == 0 else ops.pnhwc_zeros(B, H, W, 5 * F_, dev) for i in range(nblocks)]
        out = ops.pnhwc_zeros(B, H, W, F_, dev)
        layers = []
        for i in range(nblocks):
            dst = out if i == nblocks - 1 else bufs[i + 1]
            layers += _dense_block_layers(ops, L, bufs[i], dst, packs[i], bs[i],
                                          flags=kflags | (0x20000 if not chained else 0))
        if chained:
            flags = torch.zeros(ops.chain_flag_ints(B, H, W, len(layers)), dtype=torch.int32, device=dev)
            ops.conv3x3_chain(layers, flags)
        else:
            for a, _ in layers:
                import ctypes as C
                L.check(L.lib().srb_conv3x3(C.byref(a), torch.cuda.current_stream().cuda_stream), "conv3x3")
        torch.cuda.synchronize()
        return out, bufs

    out_s, bufs_s = run(False)
    out_c, bufs_c = run(True)
    assert ops.read_fault() == 0
    assert torch.equal(out_c, out_s)
    for a, b in zip(bufs_c, bufs_s):
        assert torch.equal(a, b)
    # and the sequential result is the dense block (first block only: later ones compound bf16 rounding)
    xr = x0.to(torch.bfloat16).float()
    feats = [xr]
    for k in range(5):
        y = F.conv2d(torch.cat(feats, 1), ws[0][k].to(torch.bfloat16).float(), bs[0][k], padding=1)
        if k < 4:
            y = F.leaky_relu(y, 0.01).to(torch.bfloat16).float()
        feats.append(y)
    ref = feats[5] * 0.2 + xr
    got = ops.from_pnhwc(bufs_s[1], 0, F_) if nblocks > 1 else ops.from_pnhwc(out_s)
    assert (got - ref).abs().max().item() <= 1e-2 * ref.abs().max().item()


def test_conv3x3_chain_repeated_launches_are_deterministic():
    """20 back-to-back chained launches (fresh flags each) give the same bits: no stale read across the flag protocol."""
    dev = _dev()
    from srb200 import _lib as L, ops
    g = torch.Generator(device="cuda").manual_seed(3)
    B, H, W, F_ = 512, 20, 20, 64
    x0 = torch.randn(B, F_, H, W, generator=g, device=dev)
    ws = [torch.randn(F_, k * F_, 3, 3, generator=g, device=dev) * (1.0 / (9 * k * F_)) ** 0.5 for k in range(1, 6)]
    bs = [torch.randn(F_, generator=g, device=dev) * 0.05 for _ in range(5)]
    packs = [ops.pack_fwd(w) for w in ws]
    first = None
    for it in range(20):
        buf = ops.to_pnhwc(x0, 5 * F_)
        out = ops.pnhwc_zeros(B, H, W, F_, dev)
        layers = _dense_block_layers(ops, L, buf, out, packs, bs)
        flags = torch.zeros(ops.chain_flag_ints(B, H, W, 5), dtype=torch.int32, device=dev)
        ops.conv3x3_chain(layers, flags)
        torch.cuda.synchronize()
        if first is None:
            first = (out.clone(), buf.clone())
        else:
            assert torch.equal(out, first[0]) and torch.equal(buf, first[1]), it
    assert ops.read_fault() == 0


def test_conv3x3_chain_rejects_bad_arguments():
    dev = _dev()
    from srb200 import _lib as L, ops
    B, H, W, F_ = 2, 20, 20, 64
    buf = ops.pnhwc_zeros(B, H, W, 5 * F_, dev)
    out = ops.pnhwc_zeros(B, H, W, F_, dev)
    w = ops.pack_fwd(torch.zeros(F_, F_, 3, 3, device=dev))
    a = ops.conv3x3_args(buf, 0, F_, w, None, buf, F_, F_)
    b = ops.conv3x3_args(buf, 0, F_, w, None, out, 0, F_)
    small = torch.zeros(1, dtype=torch.int32, device=dev)
    with pytest.raises(RuntimeError, match="flag workspace too small"):
        ops.conv3x3_chain([a, b], small)
    up = ops.pnhwc_zeros(B, 2 * H, 2 * W, F_, dev)
    w4 = ops.pack_fwd(torch.zeros(4 * F_, F_, 3, 3, device=dev), shuffle=True)
    c = ops.conv3x3_args(buf, 0, F_, w4, None, up, 0, 4 * F_, shuffle=True)
    flags = torch.zeros(ops.chain_flag_ints(B, H, W, 2), dtype=torch.int32, device=dev)
    with pytest.raises(RuntimeError, match="share batch"):
        ops.conv3x3_chain([a, c], flags)


@pytest.mark.parametrize("B,H,W,cin", [(3, 20, 20, 64), (130, 20, 20, 128), (512, 20, 20, 64), (2, 80, 80, 64)])
@pytest.mark.parametrize("chained", [False, True])
def test_conv3x3_colsum_is_the_bias_gradient(B, H, W, cin, chained):
    """srb_conv3x3_args.colsum: per-channel sums of the stored outputs over valid pixels (what aten::convolution_backward
    returns as grad_bias of the consumer), accumulated on top of the caller's values; fp32 sums of the pre-rounding
    outputs, compared with the sums of the stored bf16 outputs at 2e-3 of the column's absolute mass."""
    dev = _dev()
    from srb200 import _lib as L, ops
    g = torch.Generator(device="cuda").manual_seed(B + cin)
    xb, _ = _rand_pnhwc(B, H, W, cin, dev, g)
    mask_b, _ = _rand_pnhwc(B, H, W, 64, dev, g)
    w = torch.randn(64, cin, 3, 3, generator=g, device=dev) * (2.0 / (9 * cin)) ** 0.5
    y = ops.pnhwc_zeros(B, H, W, 64, dev)
    y2 = ops.pnhwc_zeros(B, H, W, 64, dev)
    cs = torch.full((128,), 1.5, device=dev)
    wp1, wp2 = ops.pack_fwd(w), ops.pack_fwd(w[:, :64].contiguous())      # the args hold raw pointers: keep these alive
    a1 = ops.conv3x3_args(xb, 0, cin, wp1, None, y, 0, 64, act=L.ACT_LRELU_MASK, slope=0.01,
                          mask=(mask_b, 0), colsum=cs[:64])
    a2 = ops.conv3x3_args(y, 0, 64, wp2, None, y2, 0, 64, alpha=0.5, res1=(y, 0, 1.0), colsum=cs[64:])
    if chained:
        ops.conv3x3_chain([a1, a2], torch.zeros(ops.chain_flag_ints(B, H, W, 2), dtype=torch.int32, device=dev))
    else:
        import ctypes as C
        for a, _ in (a1, a2):
            L.check(L.lib().srb_conv3x3(C.byref(a), torch.cuda.current_stream().cuda_stream), "conv3x3")
    torch.cuda.synchronize()
    assert ops.read_fault() == 0
    for out, got in ((y, cs[:64]), (y2, cs[64:])):
        o = ops.from_pnhwc(out)
        ref = o.sum(dim=(0, 2, 3)) + 1.5
        mass = o.abs().sum(dim=(0, 2, 3)) + 1.0
        assert ((got - ref).abs() <= 2e-3 * mass).all(), ((got - ref).abs() / mass).max().item()


@pytest.mark.parametrize("B,H,W", [(3, 20, 20), (130, 20, 20), (2, 80, 80)])
def test_conv3x3_sign_bits_reproduce_the_activation_mask(B, H, W):
    """sign_out / mask_bits: a LeakyReLU convolution writes (output > 0) bits; an input-gradient convolution masked by
    those bits gives exactly the bits of the one masked by the bf16 activation itself."""
    dev = _dev()
    from srb200 import _lib as L, ops
    g = torch.Generator(device="cuda").manual_seed(B + H)
    xb, _ = _rand_pnhwc(B, H, W, 64, dev, g)
    gb, _ = _rand_pnhwc(B, H, W, 128, dev, g)
    w1 = ops.pack_fwd(torch.randn(64, 64, 3, 3, generator=g, device=dev) * 0.06)
    w2 = ops.pack_fwd(torch.randn(64, 128, 3, 3, generator=g, device=dev) * 0.04)
    bias = torch.randn(64, generator=g, device=dev) * 0.1
    act = ops.pnhwc_zeros(B, H, W, 320, dev)
    sign = torch.zeros(B, H + 1, W + 1, 8, dtype=torch.int32, device=dev)
    ops.conv3x3(xb, 0, 64, w1, bias, act, 128, 64, act=L.ACT_LRELU, slope=0.01, sign_out=(sign, 2))
    y_ref, y_bits = ops.pnhwc_zeros(B, H, W, 64, dev), ops.pnhwc_zeros(B, H, W, 64, dev)
    ops.conv3x3(gb, 0, 128, w2, None, y_ref, 0, 64, act=L.ACT_LRELU_MASK, slope=0.01, mask=(act, 128))
    ops.conv3x3(gb, 0, 128, w2, None, y_bits, 0, 64, act=L.ACT_LRELU_MASK, slope=0.01, mask_bits=(sign, 2))
    torch.cuda.synchronize()
    assert ops.read_fault() == 0
    assert torch.equal(y_ref, y_bits)
    a = ops.from_pnhwc(act, 128, 64)
    words = sign[:, :H, :W, 2:4].reshape(B, H, W, 2, 1)
    bits = ((words >> torch.arange(32, device=dev, dtype=torch.int32)) & 1).reshape(B, H, W, 64).permute(0, 3, 1, 2)
    assert torch.equal(bits.bool(), a > 0)
    assert int(sign[:, :, :, :2].abs().sum()) == 0 and int(sign[:, :, :, 4:].abs().sum()) == 0    # other words untouched
