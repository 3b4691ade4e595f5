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
        data.image_stats(x.cuda(), nmax=9)
