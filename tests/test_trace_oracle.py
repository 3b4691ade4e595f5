"""Pins the oracle's loss assembly (rows a8 / a9 of SURVEY.md section 8) against the reference's OWN esrgan.train run:
tests/golden/golden_trace_v1.json is info['loss'] of the unmodified reference (tests/golden/make_golden_trace.py, harness
in tests/ref_harness.py) for 20 GAN steps at batch 16.  CPU only."""
import json
import os

import numpy as np
import pytest
import torch

import ref_harness as H
from oracle import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def trace():
    with open(os.path.join(ROOT, "tests", "golden", "golden_trace_v1.json")) as f:
        return json.load(f)


def seeded_modules(cfg):
    """The drop-in modules constructed as esrgan.py:177-220 does (seed, then G, D_def, D_pow): the constructors draw from
    the CPU generator in the reference's order, so the weights are the reference run's initial weights."""
    import models
    torch.manual_seed(cfg["seed"])
    np.random.seed(cfg["seed"])
    g = models.GeneratorRRDB(1, filters=64, num_res_blocks=cfg["residual_blocks"], num_upsample=2, res_scale=0.1)
    ds = [models.Markovian_Discriminator((1, 80, 80), [16, 32, 32, 64]) for _ in range(2)]
    return g, ds


def batch_of(trace, step):
    cfg = trace["cfg"]
    hits = H.synthetic_hits(cfg["n_batches"] * cfg["batch_size"], 80, cfg["seed"])
    idx = trace["order"][step * cfg["batch_size"]:(step + 1) * cfg["batch_size"]]
    return H.hits_to_batch([hits[i] for i in idx], 80, cfg["factor"])


def test_trace_fixture_shape(trace):
    cfg = trace["cfg"]
    assert len(trace["loss"]) == 23 and all(len(v) == cfg["n_batches"] for v in trace["loss"].values())
    assert sorted(trace["order"]) == list(range(cfg["n_batches"] * cfg["batch_size"]))
    assert len(trace["epsilons"]) == 2 * cfg["n_batches"]


def test_drop_in_constructors_reproduce_the_reference_initial_weights(trace):
    g, ds = seeded_modules(trace["cfg"])
    for m, (s, a) in zip([g] + ds, trace["init_checksums"]):
        ps = list(m.parameters())
        assert abs(sum(p.detach().double().sum().item() for p in ps) - s) <= 1e-9 * max(1.0, abs(s))
        assert abs(sum(p.detach().double().abs().sum().item() for p in ps) - a) <= 1e-9 * a


def test_oracle_losses_equal_step0_of_the_reference_train(trace):
    g, ds = seeded_modules(trace["cfg"])
    sd_g = {k: v.detach().clone() for k, v in g.state_dict().items()}
    sd_d = [{k: v.detach().clone() for k, v in d.state_dict().items()} for d in ds]
    hr, lr = batch_of(trace, 0)
    with torch.no_grad():
        out, srs = O.generator_forward(sd_g, lr, res_scale=0.1, thres=0.0, training=True)
        loss_g, parts = O.generator_loss((out, srs), hr, lr, sd_d, factor=4)
    L = trace["loss"]
    ref = {"g_loss": loss_g.item(), "pixel_loss": parts["hr0"].item(), "pixel_loss_pow": parts["hr1"].item(),
           "lr_loss": parts["lr0"].item(), "lr_loss_pow": parts["lr1"].item(), "adv_loss": parts["adv0"].item(),
           "adv_loss_pow": parts["adv1"].item(),
           "def_loss": (parts["hr0"] + 0.01 * parts["adv0"] + 0.1 * parts["lr0"]).item(),      # esrgan.py:549
           "pow_loss": (parts["hr1"] + 0.01 * parts["adv1"] + 0.1 * parts["lr1"]).item()}
    for k, v in ref.items():
        assert abs(v - L[k][0]) <= 2e-5 * abs(L[k][0]) + 1e-7, (k, v, L[k][0])
    for k, name in enumerate(["d_loss_def", "d_loss_pow"]):
        eps = torch.tensor(trace["epsilons"][k]).view(-1, 1, 1, 1)
        loss_d, _ = O.discriminator_loss(sd_d[k], hr, (out, srs)[k], eps, lambda_reg=0.01)
        assert abs(loss_d.item() - L[name][0]) <= 2e-5 * abs(L[name][0]), (name, loss_d.item(), L[name][0])
