"""Evaluation post-processing: evaluation/eval.py:98-105 (get_nth_hardest) and 323-339 (collector modes)
as one device kernel.  CPU: the oracle restatement against vectors produced by the reference's own get_nth_hardest
(tests/golden/make_golden_evalstats.py).  GPU: srb_image_stats against those vectors (exact) and the oracle (exact for
max / min / nnz / n-th hardest; 2e-6 relative for the fp32 sums, which numpy adds pairwise and the kernel in fp64)."""
import os
import sys

import numpy as np
import pytest
import torch

from oracle import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))


@pytest.fixture(scope="module")
def golden_stats():
    return torch.load(os.path.join(ROOT, "tests", "golden", "golden_evalstats_v1.pt"), weights_only=False)["cases"]


def _images(seed, B=24):      # as tests/golden/make_golden_evalstats.py
    hr, _ = O.synthetic_jets(B, 80, 4, seed)
    g = torch.Generator().manual_seed(seed + 1)
    x = hr + 0.01 * torch.randn(hr.shape, generator=g) * (torch.rand(hr.shape, generator=g) < 0.05)
    x[3] = 0.0
    x[4, 0, 5, 5] = x[4, 0, 9, 9] = x[4].max() + 1
    return x


@pytest.mark.parametrize("name", ["seed21", "seed22"])
def test_oracle_nth_hardest_matches_reference(golden_stats, name):
    c = golden_stats[name]
    x = _images(c["seed"]).numpy()
    for n, ref in c["hardest"].items():
        assert np.array_equal(np.asarray(O.nth_hardest_numpy(x, n), dtype=np.float32), ref.numpy())


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["seed21", "seed22"])
def test_image_stats_kernel(golden_stats, name):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from srb200 import data
    c = golden_stats[name]
    x = _images(c["seed"])
    thres = 0.05
    got = data.image_stats(x.cuda(), thres=thres, nmax=8)
    for n, ref in c["hardest"].items():
        assert torch.equal(got["hardest"][:, n - 1].cpu(), ref), n
    ref = O.image_stats_numpy(x.numpy(), thres)
    for k in ("max", "min", "nnz"):
        assert np.array_equal(got[k].cpu().numpy(), ref[k].astype(np.float32)), k
    for k in ("mean", "sum", "meannnz"):
        r = ref[k].astype(np.float64)
        assert np.all(np.abs(got[k].cpu().numpy() - r) <= 2e-6 * np.abs(r) + 1e-9), k
    assert float(got["meannnz"][3]) == 0.0 and float(got["nnz"][3]) == 0.0         # the empty image
    with pytest.raises(RuntimeError, match="CUDA"):
        data.image_stats(x, thres=thres)
    with pytest.raises(RuntimeError, match="nmax"):
        data.image_stats(x.cuda(), nmax=33)


# ------------------------------------------------------------------ SumRaster / MeanImage / L1 metrics / E_30
@pytest.fixture(scope="module")
def golden_metrics():
    return torch.load(os.path.join(ROOT, "tests", "golden", "golden_evalmetrics_v1.pt"), weights_only=False)["cases"]


def _pairs(seed, B=24):       # as tests/golden/make_golden_evalmetrics.py
    hr, lr = O.synthetic_jets(B, 80, 4, seed)
    g = torch.Generator().manual_seed(seed + 1)
    sr = hr * (1 + 0.1 * torch.randn(hr.shape, generator=g)) + \
        0.3 * torch.randn(hr.shape, generator=g) * (torch.rand(hr.shape, generator=g) < 0.05)
    return sr, hr, lr


VARIANTS = [(0.7, False), (0.7, True), (0.0, False)]


@pytest.mark.parametrize("name", ["seed31", "seed32"])
def test_oracle_rasters_and_l1_match_reference(golden_metrics, name):
    c = golden_metrics[name]
    sr, hr, lr = _pairs(c["seed"])
    for thr, energy in VARIANTS:
        ref = c[f"thr{thr}_e{int(energy)}"]
        ras = [sum(O.sum_raster_numpy(t[sl].numpy(), 4, thr) for sl in (slice(0, 10), slice(10, 24))) for t in (sr, hr)]
        mi = [sum(O.mean_image_numpy(t[sl].numpy(), thr, energy) for sl in (slice(0, 10), slice(10, 24))) for t in (sr, hr)]
        if ref["energy_mode"]:
            mi = [m / ref["N"] for m in mi]
        assert np.array_equal(ras[0], ref["raster_sr"].numpy()) and np.array_equal(ras[1], ref["raster_hr"].numpy())
        np.testing.assert_allclose(mi[0], ref["mean_sr"].numpy(), rtol=1e-6, atol=1e-9)
        np.testing.assert_allclose(mi[1], ref["mean_hr"].numpy(), rtol=1e-6, atol=1e-9)
    a, b = O.l1_metrics_numpy(sr, hr, lr, 4)
    np.testing.assert_allclose(a, c["lr_l1"].numpy(), rtol=1e-6)
    np.testing.assert_allclose(b, c["hr_l1"].numpy(), rtol=1e-6)
    for n, ref in c["hardest"].items():
        assert np.array_equal(np.asarray(O.nth_hardest_numpy(sr.numpy(), n), dtype=np.float32), ref.numpy())


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["seed31", "seed32"])
def test_raster_l1_and_e30_kernels(golden_metrics, name):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from srb200 import data
    c = golden_metrics[name]
    sr, hr, lr = _pairs(c["seed"])
    srd, hrd, lrd = sr.cuda(), hr.cuda(), lr.cuda()
    for thr, energy in VARIANTS:
        ref = c[f"thr{thr}_e{int(energy)}"]
        ras = data.SumRaster(4, threshold=thr)
        mi = data.MeanImage(4, threshold=thr, energy=energy)
        for sl in (slice(0, 10), slice(10, 24)):
            ras.add(srd[sl], hrd[sl])
            mi.add(srd[sl], hrd[sl])
        got_sr, got_hr = ras.get_hist()
        assert np.array_equal(got_sr, ref["raster_sr"].numpy()) and np.array_equal(got_hr, ref["raster_hr"].numpy())
        m_sr, m_hr = mi.get_hist()
        np.testing.assert_allclose(m_sr, ref["mean_sr"].numpy(), rtol=2e-6, atol=1e-9)
        np.testing.assert_allclose(m_hr, ref["mean_hr"].numpy(), rtol=2e-6, atol=1e-9)
        assert mi.N == ref["N"]
    lr_l1, hr_l1 = data.l1_metrics(srd, hrd, lrd, 4)
    np.testing.assert_allclose(lr_l1.cpu().numpy(), c["lr_l1"].numpy(), rtol=2e-6)
    np.testing.assert_allclose(hr_l1.cpu().numpy(), c["hr_l1"].numpy(), rtol=2e-6)
    got = data.image_stats(srd, thres=0.05, nmax=30)            # the default eval_modes reach E_30
    for n, ref in c["hardest"].items():
        assert torch.equal(got["hardest"][:, n - 1].cpu(), ref), n
    assert float(got["has_nan"].sum()) == 0.0
    bad = srd.clone()
    bad[2, 0, 3, 3] = float("nan")
    assert data.image_stats(bad)["has_nan"].cpu().tolist() == [0.0, 0.0, 1.0] + [0.0] * 21
