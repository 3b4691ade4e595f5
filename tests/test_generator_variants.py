"""GeneratorRRDB constructor variants (reference models.py:58-106): transposed-conv upsampling,
num_final_layer_res and Dropout2d.  tests/golden/golden_variants_v1.pt holds the reference modules' own train-mode output,
eval output and gradients (tests/golden/make_golden_variants.py).
  CPU: the oracle reproduces them (incl. the Dropout2d masks nn.Dropout2d draws under the same seed).
  GPU: the engine (same tcgen05 kernels as the default generator) against the golden vectors and the oracle's full
       gradients, with the masks the reference drew fed in explicitly."""
import os

import pytest
import torch

from oracle import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NAMES = ["tconv_alternating", "tconv_full", "final_res_1", "final_res_2_tconv", "dropout", "dropout_final"]


@pytest.fixture(scope="module")
def gv():
    return torch.load(os.path.join(ROOT, "tests", "golden", "golden_variants_v1.pt"), weights_only=False)


def _unpack(g):
    return g["q"].float() * g["scale"]


def _oracle_run(case):
    cfg = case["cfg"]
    kw = {k: v for k, v in cfg["kw"].items() if k != "drop_rate"}
    sd = O.make_generator_sd(cfg["seed"], num_res_blocks=cfg["nb"], num_upsample=cfg["nup"], gain=1.2, **kw)
    _, lr = O.synthetic_jets(cfg["batch"], cfg["lr"] << cfg["nup"], 1 << cfg["nup"], cfg["seed"] + 1)
    sdr = {k: (v.clone().requires_grad_(True) if k not in ("power", "multiplier") else v) for k, v in sd.items()}
    drop = None
    if cfg["drop"] > 0:
        torch.manual_seed(cfg["seed"] + 2)           # the reference's draws start here (make_golden_variants.py)
        drop = O.DropMasks(cfg["drop"])
    out, _ = O.generator_forward(sdr, lr, res_scale=0.1, thres=0.0, training=True, drop=drop)
    (out * out).mean().backward()
    with torch.no_grad():
        ev, _ = O.generator_forward(sd, lr, res_scale=0.1, thres=0.0, training=False)
    return sd, lr, out.detach(), ev, {k: v.grad for k, v in sdr.items() if v.requires_grad}, drop


@pytest.mark.parametrize("name", NAMES)
def test_oracle_variants_match_reference(gv, name):
    case = gv[name]
    sd, lr, out, ev, grads, _ = _oracle_run(case)
    torch.testing.assert_close(out, case["train_out"], rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(ev, case["eval_out"], rtol=1e-4, atol=1e-6)
    assert set(case["grad_norms"]) == set(grads)
    for n, ref in case["grad_norms"].items():
        assert abs(grads[n].norm().item() - ref) <= 1e-3 * ref + 1e-9, n
    for n, packed in case["grads"].items():
        r = _unpack(packed)
        assert (grads[n] - r).abs().max().item() <= 2e-3 * packed["scale"] + 1e-9, n


def test_variant_state_dict_layouts():
    import models
    for kw in (dict(use_transposed_conv=True), dict(fully_tconv_upsample=True), dict(num_final_layer_res=2),
               dict(use_transposed_conv=True, fully_tconv_upsample=True)):
        g = models.GeneratorRRDB(1, filters=64, num_res_blocks=1, num_upsample=3, **kw)
        shapes = O.generator_shapes(1, 64, 1, 3, **kw)
        sd = g.state_dict()
        assert list(sd) == list(shapes), kw
        assert all(tuple(sd[k].shape) == tuple(shapes[k]) for k in shapes), kw


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_engine_variants_vs_reference_and_oracle(gv, name):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import models
    from srb200 import ops
    dev = torch.device("cuda:0")
    case = gv[name]
    cfg = case["cfg"]
    sd, lr, o_out, o_ev, o_grads, drop = _oracle_run(case)
    g = models.GeneratorRRDB(1, filters=64, num_res_blocks=cfg["nb"], num_upsample=cfg["nup"], res_scale=0.1, **cfg["kw"])
    g.load_state_dict(sd, strict=True)
    g = g.to(dev).train()
    if drop is not None:
        ms = [m.reshape(m.shape[0], m.shape[1]).to(dev) for m in drop.masks]
        n_main = 3 * cfg["nb"]
        g._drop_masks = {"main": ms[:n_main], "final": ms[n_main:]}
    out = g(lr.to(dev))
    (out * out).mean().backward()

    def close(a, r, maxrel=3e-2, rell2=2e-2):
        a, r = a.float().cpu(), r.float()
        assert (a - r).abs().max().item() <= maxrel * r.abs().max().item(), (a - r).abs().max().item()
        assert (a - r).norm().item() <= rell2 * r.norm().item()

    close(out.detach(), case["train_out"])
    worst = (2.0, "")
    for n, p in g.named_parameters():
        if not p.requires_grad:
            continue
        assert p.grad is not None, n
        a, b = p.grad.float().cpu().reshape(-1), o_grads[n].reshape(-1)
        cos = torch.dot(a, b).item() / (a.norm().item() * b.norm().item() + 1e-30)
        worst = min(worst, (cos, n))
        floor = 0.998 if n.endswith(".bias") else 0.999      # small batches: see tests/test_generator_gpu.py
        assert cos >= floor, (n, cos)
        ref_norm = case["grad_norms"][n]
        assert abs(p.grad.float().norm().item() - ref_norm) <= 0.05 * ref_norm + 1e-12, (n, ref_norm)
    for n, packed in case["grads"].items():
        a, b = dict(g.named_parameters())[n].grad.float().cpu().reshape(-1), _unpack(packed).reshape(-1)
        assert torch.dot(a, b).item() / (a.norm().item() * b.norm().item() + 1e-30) >= 0.998, n
    g.eval()
    g._drop_masks = None
    with torch.no_grad():
        ev = g(lr.to(dev))
    close(ev, case["eval_out"])
    assert ops.read_fault() == 0


@pytest.mark.gpu
def test_dropout_draws_its_own_masks_when_none_are_given():
    """the module API path: masks drawn on the device with the Dropout2d statistics; backward consistent with them"""
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import models
    dev = torch.device("cuda:0")
    g = models.GeneratorRRDB(1, filters=64, num_res_blocks=1, num_upsample=1, drop_rate=0.5, res_scale=0.1)
    g.load_state_dict(O.make_generator_sd(3, num_res_blocks=1, num_upsample=1))
    g = g.to(dev).train()
    _, lr = O.synthetic_jets(8, 32, 2, 4)
    torch.manual_seed(1)
    a = g(lr.to(dev))
    a.sum().backward()
    m = g._engine()._saved["masks"]["main"]
    assert len(m) == 3 and all(set(x.unique().tolist()) <= {0.0, 2.0} for x in m)
    assert 0.2 < float(sum((x == 0).float().mean() for x in m) / 3) < 0.8
    b = g(lr.to(dev))
    assert not torch.equal(a, b)                      # fresh masks per forward
    g.eval()
    with torch.no_grad():
        c, d = g(lr.to(dev)), g(lr.to(dev))
    assert torch.equal(c, d)                          # no dropout in eval
