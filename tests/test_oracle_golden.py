"""Pins oracle/oracle.py against outputs of the reference itself
(tests/golden/golden_v1.pt, written by tests/golden/make_golden.py)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import oracle as O

G_CASES = ["g_fwd_train_x4", "g_fwd_eval_x4", "g_fwd_eval_pow_x2", "g_fwd_train_pow_x2", "g_fwd_nb2_x8"]


def _inputs(cfg):
    sd = O.make_generator_sd(cfg["seed"], num_res_blocks=cfg["nb"], num_upsample=cfg["nup"], power=cfg["power"],
                             multiplier=cfg["multiplier"], gain=cfg["gain"])
    hr, lr = O.synthetic_jets(cfg["batch"], cfg["lr_size"] * 2 ** cfg["nup"], 2 ** cfg["nup"], cfg["seed"] + 7)
    return sd, hr, lr


@pytest.mark.parametrize("name", G_CASES)
def test_generator_forward_matches_reference(golden, name):
    case = golden[name]
    cfg = case["cfg"]
    sd, hr, lr = _inputs(cfg)
    out, srs = O.generator_forward(sd, lr, res_scale=cfg["res_scale"], thres=cfg["thres"], training=cfg["training"])
    assert out.shape == case["out"].shape
    torch.testing.assert_close(out, case["out"], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(srs, case["srs"], rtol=1e-5, atol=1e-6)


def test_generator_grads_match_reference(golden):
    case = golden["g_train_x4"]
    cfg = case["cfg"]
    sd, hr, lr = _inputs(cfg)
    sd = {k: (v.clone().requires_grad_(True) if k not in ("power", "multiplier") else v) for k, v in sd.items()}
    out, _ = O.generator_forward(sd, lr, res_scale=cfg["res_scale"], thres=cfg["thres"], training=True)
    loss = F.l1_loss(out, hr)
    loss.backward()
    assert abs(loss.item() - case["loss"]) < 1e-6 * max(1.0, abs(case["loss"]))
    assert set(case["grads"]) == {k for k, v in sd.items() if v.requires_grad}
    for k, s in case["grads"].items():
        g = sd[k].grad.reshape(-1)
        assert g.numel() == s["numel"]
        torch.testing.assert_close(g[:64], s["head"], rtol=1e-4, atol=1e-7)
        assert abs(g.double().abs().sum().item() - s["abssum"]) <= 1e-4 * s["abssum"] + 1e-9


def test_discriminator_matches_reference(golden):
    case = golden["d_patch"]
    seed, batch = case["cfg"]["seed"], case["cfg"]["batch"]
    sd = O.make_discriminator_sd(seed, gain=2.0)
    sd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    hr, _ = O.synthetic_jets(batch, 80, 4, seed + 3)
    fake, _ = O.synthetic_jets(batch, 80, 4, seed + 4)
    fake = fake * 0.7 + 0.01
    pred_real = O.discriminator_forward(sd, hr)
    assert tuple(pred_real.shape[1:]) == tuple(case["out_shape"])
    torch.testing.assert_close(pred_real, case["pred_real"], rtol=1e-5, atol=1e-6)
    loss, parts = O.discriminator_loss(sd, hr, fake, case["epsilon"], lambda_reg=0.01)
    assert abs(parts["ragan"].item() - case["loss_ragan"]) < 1e-6
    assert abs(parts["gp"].item() - case["gp"]) < 1e-6 * max(1.0, case["gp"])
    loss.backward()
    for k, g in case["grads"].items():
        torch.testing.assert_close(sd[k].grad, g, rtol=1e-4, atol=1e-7)


def test_sumpool(golden):
    c = golden["sumpool4"]
    torch.testing.assert_close(O.sum_pool(c["x"], 4), c["y"])


@pytest.mark.parametrize("stride", [1, 2])
def test_numpy_conv_pins_torch_conv(stride):
    g = torch.Generator().manual_seed(5)
    x = torch.randn(2, 3, 9, 7, generator=g)
    w = torch.randn(4, 3, 3, 3, generator=g)
    b = torch.randn(4, generator=g)
    ref = F.conv2d(x, w, b, stride, 1).numpy()
    got = O.conv3x3_numpy(x.numpy(), w.numpy(), b.numpy(), stride)
    np.testing.assert_allclose(got, ref, rtol=1e-4, atol=1e-5)


def test_adam_matches_torch():
    g = torch.Generator().manual_seed(1)
    p0 = torch.randn(100, generator=g)
    p = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([p], lr=2e-4, betas=(0.9, 0.999), weight_decay=0.01)
    q, m, v = p0.clone(), torch.zeros(100), torch.zeros(100)
    for step in range(1, 4):
        grad = torch.randn(100, generator=g)
        p.grad = grad.clone()
        opt.step()
        O.adam_step(q, grad, m, v, step, weight_decay=0.01)
        torch.testing.assert_close(q, p.detach(), rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("name", ["nb10_x4", "nb23_x8"])
def test_oracle_matches_reference_at_benchmarked_depth(name):
    """golden_v2.pt (tests/golden/make_golden_v2.py): the reference modules at nb = 10 / x4 and nb = 23 / x8.  Inference
    is per-image, so the oracle is checked on the stored images only (seconds on the CPU)."""
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    case = torch.load(os.path.join(root, "tests", "golden", "golden_v2.pt"), weights_only=False)[name]
    cfg = case["cfg"]
    sd = O.make_generator_sd(cfg["seed"], num_res_blocks=cfg["nb"], num_upsample=cfg["nup"], gain=cfg["gain"])
    batch = cfg.get("batch", cfg.get("batch_eval"))
    _, lr = O.synthetic_jets(batch, 80, 2 ** cfg["nup"], cfg["data_seed"])
    with torch.no_grad():
        out, srs = O.generator_forward(sd, lr[case["sel"]], res_scale=cfg["res_scale"], thres=0.0, training=False)
    torch.testing.assert_close(out, case["eval_out"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(srs, case["eval_srs"], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(out.sum((1, 2, 3)), case["eval_energy"][case["sel"]], rtol=1e-4, atol=1e-4)
