"""Optional generator loss terms (esrgan.py:530-550): nnz, mask, hit and the differentiable
histogram + KLD.  Pinned by the reference's own esrgan.train traces with those lambdas switched on
(tests/golden/golden_trace_variants_v1.json: "optional_losses" and "hist", the latter after two warm-up steps and with
the bin edges the reference derived by k-means).  CPU: oracle == trace.  GPU: each fused kernel (value + gradient) against
the oracle under autograd, and GANTrainStep against the traces."""
import json
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import oracle as O
from test_trace_variants import batch_of, build, _sd

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def traces():
    with open(os.path.join(ROOT, "tests", "golden", "golden_trace_variants_v1.json")) as f:
        return json.load(f)


def test_oracle_optional_losses_equal_step0(traces):
    tr = traces["optional_losses"]
    ov = tr["cfg"]["overrides"]
    g, ds, _ = build(tr["cfg"])
    hr, lr = batch_of(tr, 0)
    with torch.no_grad():
        out, srs = O.generator_forward(_sd(g), lr, res_scale=0.1, thres=0.0, training=True)
        loss_g, parts = O.generator_loss((out, srs), hr, lr, [_sd(d) for d in ds], factor=4, lambda_nnz=ov["lambda_nnz"],
                                         lambda_mask=ov["lambda_mask"], lambda_hit=ov["lambda_hit"], sigma=500,
                                         hit_threshold=0.5)
    L = tr["loss"]
    for key, name in (("nnz0", "nnz_loss"), ("nnz1", "nnz_loss_pow"), ("mask0", "mask_loss"), ("hit0", "hit_loss"),
                      ("hit1", "hit_loss_pow")):
        assert abs(parts[key].item() - L[name][0]) <= 1e-4 * abs(L[name][0]), (key, parts[key].item(), L[name][0])
    assert abs(loss_g.item() - L["g_loss"][0]) <= 1e-4 * abs(L["g_loss"][0])


def _cpu_warmup(sd, tr, steps):
    """esrgan.py:418-427 on the CPU oracle: `steps` warm-up iterations (L1 against HR, Adam)."""
    params = {k: (v.clone().requires_grad_(True) if k not in ("power", "multiplier") else v) for k, v in sd.items()}
    opt = torch.optim.Adam([v for v in params.values() if v.requires_grad], lr=2e-4, betas=(0.9, 0.999))
    losses = []
    for s in range(steps):
        hr, lr = batch_of(tr, s)
        opt.zero_grad()
        out, _ = O.generator_forward(params, lr, res_scale=0.1, thres=0.0, training=True)
        loss = F.l1_loss(out, hr)
        loss.backward()
        opt.step()
        losses.append(loss.item())
    return {k: v.detach() for k, v in params.items()}, losses


def test_oracle_histogram_loss_equals_the_first_gan_step_after_warmup(traces):
    tr = traces["hist"]
    ov = tr["cfg"]["overrides"]
    g, ds, _ = build(tr["cfg"])
    sd, wl = _cpu_warmup(_sd(g), tr, ov["warmup_batches"])
    L = tr["loss"]
    for a, b in zip(wl, L["g_loss"][:2]):
        assert abs(a - b) <= 1e-5 * abs(b)                       # the two warm-up pixel losses
    hr, lr = batch_of(tr, ov["warmup_batches"])
    edges = [torch.tensor(tr["info_extra"][f"binedges{k}"], dtype=torch.float32) for k in range(2)]
    with torch.no_grad():
        out, srs = O.generator_forward(sd, lr, res_scale=0.1, thres=0.0, training=True)
        loss_g, parts = O.generator_loss((out, srs), hr, lr, [_sd(d) for d in ds], factor=4,
                                         lambda_hist=ov["lambda_hist"], binedges=edges, sigma=500)
    assert abs(parts["hist0"].item() - L["hist_loss"][0]) <= 2e-4 * L["hist_loss"][0], (parts["hist0"].item(), L["hist_loss"][0])
    assert abs(loss_g.item() - L["g_loss"][2]) <= 2e-4 * abs(L["g_loss"][2])


# ------------------------------------------------------------------------------------------------ GPU
def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda:0")


def _pair(seed, B=6):
    hr, _ = O.synthetic_jets(B, 80, 4, seed)
    g = torch.Generator().manual_seed(seed + 1)
    gen = hr * (1 + 0.1 * torch.randn(hr.shape, generator=g)) + 0.002 * torch.randn(hr.shape, generator=g)
    return gen, hr


@pytest.mark.gpu
def test_optional_loss_kernels_match_oracle_autograd():
    dev = _dev()
    from srb200 import losses as LS
    gen, hr = _pair(3)
    B, _, H, W = gen.shape
    work = LS.OptionalLossWork(B, H, W, 4, dev, binedges=[[0.0, 0.9, 2.1, 3.7, 5.9, 8.4, 10.8]] * 2)
    edges = torch.tensor([0.0, 0.9, 2.1, 3.7, 5.9, 8.4, 10.8])
    cases = {
        # sigma chosen so that the sigmoids are not all saturated (the production values 5e4 / 500 are checked below)
        "nnz": (lambda x: F.mse_loss(O.softgreater(x, 0, 300.0).sum((1, 2, 3)), (hr > 0).sum((1, 2, 3)).float()),
                lambda xd, gr: LS.nnz_loss(xd, hr.to(dev), 0.7, gr, work, sigma=300.0)),
        "mask": (lambda x: F.l1_loss(O.nnz_mask(x, 200.0), O.nnz_mask(hr, 200.0)),
                 lambda xd, gr: LS.mask_loss(xd, hr.to(dev), 0.7, gr, sigma=200.0)),
        "hit": (lambda x: F.mse_loss(O.get_hitogram(x, 4, 0.5, 3.0), O.get_hitogram(hr, 4, 0.5, 3.0)),
                lambda xd, gr: LS.hit_loss(xd, hr.to(dev), 4, 0.5, 3.0, 0.7, gr, work)),
        "hist": (lambda x: O.kld_hist(O.diffable_histogram(x[x > 0], edges, 4.0),
                                      O.diffable_histogram(hr[hr > 0], edges, 4.0), edges),
                 lambda xd, gr: LS.hist_loss(xd, hr.to(dev), 0, 4.0, 0.7, gr, work)),
    }
    for name, (ref_fn, dev_fn) in cases.items():
        x = gen.clone().requires_grad_(True)
        ref = ref_fn(x)
        ref.backward()
        scale = x.grad.abs().max().item()
        assert scale > 0, name
        base = torch.full((B, 1, H, W), 8.0 * scale)               # kernels ADD w * d loss / d gen to what is there
        grad = base.to(dev)
        got = dev_fn(gen.to(dev), grad)
        assert abs(got.item() - ref.item()) <= 2e-4 * abs(ref.item()) + 1e-9, (name, got.item(), ref.item())
        gd = (grad.cpu() - base) / 0.7
        assert (gd - x.grad).abs().max().item() <= 3e-3 * scale, (name, (gd - x.grad).abs().max().item(), scale)
    # production sigmas: values only (gradients are delta-like there)
    x = gen
    grad = torch.zeros(B, 1, H, W, device=dev)
    for got, ref in ((LS.nnz_loss(x.to(dev), hr.to(dev), 0.0, grad, work),
                      F.mse_loss(O.softgreater(x, 0, 50000).sum((1, 2, 3)), (hr > 0).sum((1, 2, 3)).float())),
                     (LS.mask_loss(x.to(dev), hr.to(dev), 0.0, grad),
                      F.l1_loss(O.nnz_mask(x), O.nnz_mask(hr))),
                     (LS.hit_loss(x.to(dev), hr.to(dev), 4, 0.5, 500.0, 0.0, grad, work),
                      F.mse_loss(O.get_hitogram(x, 4, 0.5, 500), O.get_hitogram(hr, 4, 0.5, 500))),
                     (LS.hist_loss(x.to(dev), hr.to(dev), 1, 500.0, 0.0, grad, work),
                      O.kld_hist(O.diffable_histogram(x[x > 0], edges, 500), O.diffable_histogram(hr[hr > 0], edges, 500),
                                 edges))):
        assert abs(got.item() - ref.item()) <= 1e-3 * abs(ref.item()) + 1e-9, (got.item(), ref.item())


@pytest.mark.gpu
def test_fused_step_with_optional_losses_matches_the_reference_trace_at_step0(traces):
    dev = _dev()
    from srb200.train_step import GANTrainStep
    tr = traces["optional_losses"]
    ov = tr["cfg"]["overrides"]
    G, Ds, _ = build(tr["cfg"])
    G, Ds = G.to(dev).train(), [d.to(dev) for d in Ds]
    st = GANTrainStep(G, Ds, factor=4, lambdas=(0.2, 1.0), lambda_hr=1.0, lambda_lr=0.1, lambda_adv=0.01,
                      lambda_reg=0.01, d_threshold=0.001, lr=2e-4, lambda_nnz=ov["lambda_nnz"],
                      lambda_mask=ov["lambda_mask"], lambda_hit=ov["lambda_hit"], sigma=500.0, hit_threshold=0.5)
    B = tr["cfg"]["batch_size"]
    hr, lr = batch_of(tr, 0)
    eps = [torch.tensor(tr["epsilons"][k]).view(B, 1, 1, 1).to(dev) for k in range(2)]
    Lv = {k: float(v) for k, v in st.step(lr.to(dev), hr.to(dev), epsilons=eps).items()}
    L = tr["loss"]
    # the soft counters sit on sigmoid(5e4 x): every generated pixel within ~1e-4 of zero moves them, so the bf16
    # generator agrees with the fp32 trace to a few per cent there, not to 2e-2 like the pixel terms
    for key, name, tol in (("nnz0", "nnz_loss", 0.1), ("mask0", "mask_loss", 0.03), ("hit0", "hit_loss", 0.1),
                           ("hr0", "pixel_loss", 0.02), ("lr0", "lr_loss", 0.02), ("g_loss", "g_loss", 0.1)):
        assert abs(Lv[key] - L[name][0]) <= tol * abs(L[name][0]), (key, Lv[key], L[name][0])
    for _ in range(3):                                     # and the terms keep training: nnz collapses as in the trace
        Lv2 = {k: float(v) for k, v in st.step(lr.to(dev), hr.to(dev), epsilons=eps).items()}
    assert np.isfinite(list(Lv2.values())).all() and Lv2["nnz0"] < 0.2 * Lv["nnz0"]


@pytest.mark.gpu
def test_fused_step_histogram_loss_follows_the_reference_trace(traces):
    dev = _dev()
    from srb200.train_step import GANTrainStep, GeneratorTrainStep
    tr = traces["hist"]
    ov = tr["cfg"]["overrides"]
    G, Ds, _ = build(tr["cfg"])
    G, Ds = G.to(dev).train(), [d.to(dev) for d in Ds]
    edges = [tr["info_extra"][f"binedges{k}"] for k in range(2)]
    st = GANTrainStep(G, Ds, factor=4, lambdas=(0.2, 1.0), lambda_hr=1.0, lambda_lr=0.1, lambda_adv=0.01,
                      lambda_reg=0.01, d_threshold=0.001, lr=2e-4, lambda_hist=ov["lambda_hist"], binedges=edges,
                      sigma=500.0)
    warm = GeneratorTrainStep(G)
    warm.opt_G = st.opt_G                                  # esrgan.py uses ONE optimizer_G for warm-up and GAN steps
    L = tr["loss"]
    B = tr["cfg"]["batch_size"]
    for s in range(ov["warmup_batches"]):
        hr, lr = batch_of(tr, s)
        l_ = float(warm.step(lr.to(dev), hr.to(dev)))
        assert abs(l_ - L["g_loss"][s]) <= 0.02 * L["g_loss"][s]
    n_gan = tr["cfg"]["n_batches"] - ov["warmup_batches"]
    for s in range(n_gan):
        hr, lr = batch_of(tr, ov["warmup_batches"] + s)
        eps = [torch.tensor(tr["epsilons"][2 * s + k]).view(B, 1, 1, 1).to(dev) for k in range(2)]
        Lv = {k: float(v) for k, v in st.step(lr.to(dev), hr.to(dev), epsilons=eps).items()}
        # sigma = 500 makes the bins near-hard counters: after a couple of updates single pixels that sit on a bin edge
        # fall on different sides in bf16 and fp32 (the batch is 4 images), so the first GAN step is held to 5 % and the
        # later ones to 30 % (measured: 0.3 %, 2 %, 8 %, 17 %)
        tol = 0.05 if s == 0 else 0.30
        assert abs(Lv["hist0"] - L["hist_loss"][s]) <= tol * L["hist_loss"][s], (s, Lv["hist0"], L["hist_loss"][s])
        assert abs(Lv["g_loss"] - L["g_loss"][ov["warmup_batches"] + s]) <= tol * L["g_loss"][ov["warmup_batches"] + s]
