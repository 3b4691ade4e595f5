"""The drop-in under the reference's OWN driver, and the 20-step loss trace ("loss-trace agreement
within 5 % over 20 steps").

tests/golden/golden_trace_v1.json = info['loss'] of the unmodified reference esrgan.train on CPU fp32 (B = 16, nb = 10,
x4, 20 GAN steps; tests/golden/make_golden_trace.py).  Here, on the GPU:
  * baseline/_ref/esrgan.py (an unmodified copy of the reference tree that __graft_entry__.build() makes; git-ignored)
    is run with super-resolution_b200/ first on sys.path, so `from models import *` resolves to the drop-in modules:
    generator, discriminators, SumPool2d inside the dataset, torch.optim.Adam, autograd double backward of the
    gradient penalty - all through the module API;
  * the fused srb200.train_step.GANTrainStep replays the same batches, interpolation draws and initial weights.
Both must follow the reference trace within 5 % at every step."""
import importlib.util
import json
import os
import subprocess
import sys

import pytest
import torch

import ref_harness as H
from oracle import oracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref")
TRACE_TOL = 0.05
SERIES = ["g_loss", "def_loss", "pow_loss", "pixel_loss", "pixel_loss_pow", "lr_loss", "lr_loss_pow", "adv_loss",
          "adv_loss_pow", "d_loss_def", "d_loss_pow"]


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def trace():
    with open(os.path.join(ROOT, "tests", "golden", "golden_trace_v1.json")) as f:
        return json.load(f)


def _compare(got, trace, what):
    worst = (0.0, "", -1)
    for name in SERIES:
        ref = trace["loss"][name]
        assert len(got[name]) == len(ref), (what, name, len(got[name]))
        for i, (a, b) in enumerate(zip(got[name], ref)):
            worst = max(worst, (abs(a - b) / abs(b), name, i))
    assert worst[0] <= TRACE_TOL, (what, worst)
    return worst


def test_reference_esrgan_train_runs_on_the_drop_in_and_follows_the_reference_trace(trace, tmp_path):
    _dev()
    if not os.path.isdir(REF):
        pytest.skip("baseline/_ref (copy of the reference tree, made by __graft_entry__.build()) is not present")
    out = tmp_path / "trace_dropin.json"
    cfg = trace["cfg"]
    cmd = [sys.executable, os.path.join(ROOT, "tests", "ref_harness.py"), "--ref", REF, "--models",
           os.path.join(ROOT, "super-resolution_b200"), "--out", str(out), "--n-batches", str(cfg["n_batches"]),
           "--batch-size", str(cfg["batch_size"]), "--seed", str(cfg["seed"]), "--factor", str(cfg["factor"]),
           "--residual-blocks", str(cfg["residual_blocks"])]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    got = json.load(open(out))
    assert got["models_file"].startswith(os.path.join(ROOT, "super-resolution_b200")), got["models_file"]
    assert got["device"] == "cuda"
    assert got["order"] == trace["order"]                       # same shuffles: the constructors drew the same numbers
    for a, b in zip(got["init_checksums"], trace["init_checksums"]):
        assert abs(a[1] - b[1]) <= 1e-6 * b[1]
    worst = _compare(got["loss"], trace, "drop-in under esrgan.train")
    print("drop-in under esrgan.train: worst deviation from the reference trace", worst)


def test_fused_gan_step_follows_the_reference_trace(trace):
    dev = _dev()
    import models
    import numpy as np
    from srb200.train_step import GANTrainStep
    cfg = trace["cfg"]
    torch.manual_seed(cfg["seed"])
    np.random.seed(cfg["seed"])
    G = models.GeneratorRRDB(1, filters=64, num_res_blocks=cfg["residual_blocks"], num_upsample=2, res_scale=0.1)
    Ds = [models.Markovian_Discriminator((1, 80, 80), [16, 32, 32, 64]) for _ in range(2)]
    G, Ds = G.to(dev).train(), [d.to(dev) for d in Ds]
    tr = GANTrainStep(G, Ds, factor=4, lambdas=(0.2, 1.0), lambda_hr=1.0, lambda_lr=0.1, lambda_adv=0.01,
                      lambda_reg=0.01, d_threshold=0.001, lr=2e-4)
    hits = H.synthetic_hits(cfg["n_batches"] * cfg["batch_size"], 80, cfg["seed"])
    B = cfg["batch_size"]
    got = {k: [] for k in SERIES}
    for step in range(cfg["n_batches"]):
        idx = trace["order"][step * B:(step + 1) * B]
        hr, lr = H.hits_to_batch([hits[i] for i in idx], 80, 4)
        eps = [torch.tensor(trace["epsilons"][2 * step + k]).view(B, 1, 1, 1).to(dev) for k in range(2)]
        L = {k: float(v) for k, v in tr.step(lr.to(dev), hr.to(dev), epsilons=eps).items()}
        got["g_loss"].append(L["g_loss"])
        got["pixel_loss"].append(L["hr0"]); got["pixel_loss_pow"].append(L["hr1"])
        got["lr_loss"].append(L["lr0"]); got["lr_loss_pow"].append(L["lr1"])
        got["adv_loss"].append(L["adv0"]); got["adv_loss_pow"].append(L["adv1"])
        got["def_loss"].append(L["hr0"] + 0.01 * L["adv0"] + 0.1 * L["lr0"])       # esrgan.py:549
        got["pow_loss"].append(L["hr1"] + 0.01 * L["adv1"] + 0.1 * L["lr1"])
        got["d_loss_def"].append(L["d_loss0"]); got["d_loss_pow"].append(L["d_loss1"])
    worst = _compare(got, trace, "fused GANTrainStep")
    print("fused step: worst deviation from the reference trace", worst)


def test_reference_checkpoints_round_trip_strict():
    """esrgan.py:253,258,380-388: a .pth written by the reference loads into the drop-in with strict=True and back, and
    the drop-in then reproduces the reference module's output (reference on the host CPU, fp32)."""
    dev = _dev()
    if not os.path.isdir(REF):
        pytest.skip("baseline/_ref is not present")
    import models
    spec = importlib.util.spec_from_file_location("ref_models_for_test", os.path.join(REF, "models.py"))
    R = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(R)
    torch.manual_seed(3)
    ref = R.GeneratorRRDB(1, filters=64, num_res_blocks=2, num_upsample=2, res_scale=0.1)
    ours = models.GeneratorRRDB(1, filters=64, num_res_blocks=2, num_upsample=2, res_scale=0.1)
    ours.load_state_dict(ref.state_dict(), strict=True)
    ours = ours.to(dev).eval()
    ref.eval()
    _, lr = O.synthetic_jets(3, 80, 4, 5)
    with torch.no_grad():
        a, b = ours(lr.to(dev)).cpu(), ref(lr)
    assert (a - b).abs().max().item() <= 3e-2 * b.abs().max().item()
    ref2 = R.GeneratorRRDB(1, filters=64, num_res_blocks=2, num_upsample=2, res_scale=0.1)
    ref2.load_state_dict({k: v.cpu() for k, v in ours.state_dict().items()}, strict=True)
    assert all(torch.equal(p, q) for p, q in zip(ref.state_dict().values(), ref2.state_dict().values()))
    rd = R.Markovian_Discriminator((1, 80, 80), [16, 32, 32, 64])
    od = models.Markovian_Discriminator((1, 80, 80), [16, 32, 32, 64])
    od.load_state_dict(rd.state_dict(), strict=True)
    hr, _ = O.synthetic_jets(3, 80, 4, 6)
    with torch.no_grad():
        pa, pb = od.to(dev)(hr.to(dev)).cpu(), rd(hr)
    assert (pa - pb).abs().max().item() <= 1e-3 * pb.abs().max().item() + 1e-5
