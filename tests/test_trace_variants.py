"""The non-default training recipes of esrgan.train that the fused step covers:
opt.scaling_power != 1 (esrgan.py:470-473) and the second-discriminator ensemble (esrgan.py:221-237,495-505,789-795).
tests/golden/golden_trace_variants_v1.json = info['loss'] of the UNMODIFIED reference for 6 steps of each
(tests/golden/make_golden_trace.py --variants).  CPU: the oracle's loss assembly equals step 0.  GPU: GANTrainStep
replays the same batches / interpolation draws / initial weights and follows the traces within 5 %."""
import json
import os

import numpy as np
import pytest
import torch

import ref_harness as H
from oracle import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SERIES = ["g_loss", "pixel_loss", "pixel_loss_pow", "lr_loss", "lr_loss_pow", "adv_loss", "adv_loss_pow", "d_loss_def",
          "d_loss_pow"]


@pytest.fixture(scope="module")
def traces():
    with open(os.path.join(ROOT, "tests", "golden", "golden_trace_variants_v1.json")) as f:
        return json.load(f)


def build(cfg, power=1.0, second=False):
    """modules in the reference's construction order (esrgan.py:193-237): G, D_def, D_pow, then the second ones"""
    import models
    torch.manual_seed(cfg["seed"])
    np.random.seed(cfg["seed"])
    g = models.GeneratorRRDB(1, filters=64, num_res_blocks=cfg["residual_blocks"], num_upsample=2, power=power,
                             res_scale=0.1)
    ds = [models.Markovian_Discriminator((1, 80, 80), [16, 32, 32, 64]) for _ in range(2)]
    d2 = [models.Markovian_Discriminator((1, 80, 80), [16, 32, 32, 64]) for _ in range(2)] if second else None
    return g, ds, d2


def batch_of(tr, step):
    cfg = tr["cfg"]
    hits = H.synthetic_hits(cfg["n_batches"] * cfg["batch_size"], 80, cfg["seed"])
    idx = tr["order"][step * cfg["batch_size"]:(step + 1) * cfg["batch_size"]]
    return H.hits_to_batch([hits[i] for i in idx], 80, cfg["factor"])


def _sd(m):
    return {k: v.detach().clone() for k, v in m.state_dict().items()}


def test_oracle_scaling_power_equals_step0(traces):
    tr = traces["scaling_power_0.5"]
    g, ds, _ = build(tr["cfg"], power=0.5)
    hr, lr = batch_of(tr, 0)
    with torch.no_grad():
        out, srs = O.generator_forward(_sd(g), lr, res_scale=0.1, thres=0.0, training=True)
        loss_g, parts = O.generator_loss((out, srs), hr, lr, [_sd(d) for d in ds], factor=4, scaling_power=0.5)
    L = tr["loss"]
    for key, name in (("hr0", "pixel_loss"), ("hr1", "pixel_loss_pow"), ("lr0", "lr_loss"), ("lr1", "lr_loss_pow"),
                      ("adv0", "adv_loss"), ("adv1", "adv_loss_pow")):
        assert abs(parts[key].item() - L[name][0]) <= 2e-5 * abs(L[name][0]) + 1e-7, (key, parts[key].item(), L[name][0])
    assert abs(loss_g.item() - L["g_loss"][0]) <= 2e-5 * abs(L["g_loss"][0])
    for k, name in enumerate(["d_loss_def", "d_loss_pow"]):
        eps = torch.tensor(tr["epsilons"][k]).view(-1, 1, 1, 1)
        real = hr if k == 0 else hr ** 0.5
        loss_d, _ = O.discriminator_loss(_sd(ds[k]), real, (out, srs)[k], eps, lambda_reg=0.01)
        assert abs(loss_d.item() - L[name][0]) <= 2e-5 * abs(L[name][0]), name


def test_oracle_second_discriminators_equal_step0(traces):
    tr = traces["second_discriminators"]
    g, ds, d2 = build(tr["cfg"], second=True)
    hr, lr = batch_of(tr, 0)
    with torch.no_grad():
        out, srs = O.generator_forward(_sd(g), lr, res_scale=0.1, thres=0.0, training=True)
        _, p1 = O.generator_loss((out, srs), hr, lr, [_sd(d) for d in ds], factor=4)
        _, p2 = O.generator_loss((out, srs), hr, lr, [_sd(d) for d in d2], factor=4)
    L = tr["loss"]
    for k, name in enumerate(["adv_loss", "adv_loss_pow"]):
        adv = 0.5 * (p1[f"adv{k}"] + p2[f"adv{k}"]).item()                                  # esrgan.py:498-502
        assert abs(adv - L[name][0]) <= 2e-5 * abs(L[name][0]), (name, adv, L[name][0])
    for k in range(2):
        eps = torch.tensor(tr["epsilons"][k]).view(-1, 1, 1, 1)
        for mods, name in ((ds, ["d_loss_def", "d_loss_pow"][k]), (d2, ["d2_loss_def", "d2_loss_pow"][k])):
            loss_d, _ = O.discriminator_loss(_sd(mods[k]), hr, (out, srs)[k], eps, lambda_reg=0.01)
            assert abs(loss_d.item() - L[name][0]) <= 2e-5 * abs(L[name][0]), name


def _run_fused(tr, dev, power=1.0, second=False, interval=0):
    from srb200.train_step import GANTrainStep
    cfg = tr["cfg"]
    G, Ds, D2 = build(cfg, power=power, second=second)
    G, Ds = G.to(dev).train(), [d.to(dev) for d in Ds]
    D2 = [d.to(dev) for d in D2] if D2 else None
    st = GANTrainStep(G, Ds, factor=4, lambdas=(0.2, 1.0), lambda_hr=1.0, lambda_lr=0.1, lambda_adv=0.01,
                      lambda_reg=0.01, d_threshold=0.001, lr=2e-4, scaling_power=power, second_discriminators=D2,
                      second_discr_reset_interval=interval)
    B = cfg["batch_size"]
    got = {}
    for step in range(cfg["n_batches"]):
        hr, lr = batch_of(tr, step)
        eps = [torch.tensor(tr["epsilons"][2 * step + k]).view(B, 1, 1, 1).to(dev) for k in range(2)]
        Lv = {k: float(v) for k, v in st.step(lr.to(dev), hr.to(dev), epsilons=eps).items()}
        for name, key in (("g_loss", "g_loss"), ("pixel_loss", "hr0"), ("pixel_loss_pow", "hr1"), ("lr_loss", "lr0"),
                          ("lr_loss_pow", "lr1"), ("adv_loss", "adv0"), ("adv_loss_pow", "adv1"),
                          ("d_loss_def", "d_loss0"), ("d_loss_pow", "d_loss1"), ("d2_loss_def", "d2_loss0"),
                          ("d2_loss_pow", "d2_loss1")):
            if key in Lv:
                got.setdefault(name, []).append(Lv[key])
    return got, st


def _follow(got, tr, names, upto=None):
    worst = (0.0, "", -1)
    for n in names:
        ref = tr["loss"][n][:upto]
        for i, (a, b) in enumerate(zip(got[n][:upto], ref)):
            worst = max(worst, (abs(a - b) / abs(b), n, i))
    assert worst[0] <= 0.05, worst


@pytest.mark.gpu
def test_fused_step_scaling_power_follows_the_reference_trace(traces):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    got, _ = _run_fused(traces["scaling_power_0.5"], torch.device("cuda:0"), power=0.5)
    _follow(got, traces["scaling_power_0.5"], SERIES)


@pytest.mark.gpu
def test_fused_step_second_discriminators_follow_the_reference_trace(traces):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    tr = traces["second_discriminators"]
    interval = tr["cfg"]["overrides"]["second_discr_reset_interval"]
    dev = torch.device("cuda:0")
    got, st = _run_fused(tr, dev, second=True, interval=interval)
    _follow(got, tr, SERIES)
    # the second discriminators are re-initialised after iteration `interval` (random draw: the traces part there)
    _follow(got, tr, ["d2_loss_def", "d2_loss_pow"], upto=interval)
    assert st.batches_done == tr["cfg"]["n_batches"] - 1
    # and a reset really re-draws their weights
    w0 = st.D2s[0].model[0].weight.detach().clone()
    st.batches_done = interval - 2
    st.end_of_batch()
    assert not torch.equal(w0, st.D2s[0].model[0].weight.detach())
