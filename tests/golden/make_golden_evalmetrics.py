"""Golden vectors for the rest of the evaluation post-processing, produced by the reference's
own code: ``SumRaster`` / ``MeanImage`` (utils.py:275-328), ``get_nth_hardest`` for the default eval_modes up to E_30
(evaluation/eval.py:98-105, options/default.json) and the HR / LR L1 lines of ``calculate_metrics`` (eval.py:477-485,
run here with the reference's SumPool2d and nn.L1Loss(reduction='none') exactly as written there).
utils.py / eval.py cannot be imported (matplotlib, energyflow, CWD-relative options), so the UNMODIFIED classes /
functions are taken out of the reference files with ``ast`` and executed as they are.
Build container only:  python tests/golden/make_golden_evalmetrics.py  ->  tests/golden/golden_evalmetrics_v1.pt"""
import ast
import os
import sys

import numpy as np
import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402


def take(path, names, ns):
    tree = ast.parse(open(path).read())
    body = [n for n in tree.body if isinstance(n, (ast.FunctionDef, ast.ClassDef)) and n.name in names]
    assert len(body) == len(names), [n.name for n in body]
    exec(compile(ast.Module(body=body, type_ignores=[]), path, "exec"), ns)
    return ns


ns = take("/root/reference/utils.py", ["SumRaster", "MeanImage"], {"np": np, "torch": torch})
ns = take("/root/reference/evaluation/eval.py", ["get_nth_hardest"], ns)
ns = take("/root/reference/models.py", ["SumPool2d"], dict(ns, nn=nn))
SumRaster, MeanImage, get_nth_hardest, SumPool2d = ns["SumRaster"], ns["MeanImage"], ns["get_nth_hardest"], ns["SumPool2d"]


def images(seed, B=24):
    """(SR-like, HR) pairs: the HR jets and a perturbed copy with small positive / negative residue."""
    hr, lr = O.synthetic_jets(B, 80, 4, seed)
    g = torch.Generator().manual_seed(seed + 1)
    sr = hr * (1 + 0.1 * torch.randn(hr.shape, generator=g)) + \
        0.3 * torch.randn(hr.shape, generator=g) * (torch.rand(hr.shape, generator=g) < 0.05)
    return sr, hr, lr


cases = {}
for seed in (31, 32):
    sr, hr, lr = images(seed)
    c = {"seed": seed}
    # two batches accumulated, as evaluation/eval.py:376-379 does per batch
    for thr, energy in ((0.7, False), (0.7, True), (0.0, False)):
        ras = SumRaster(4, threshold=thr)
        mi = MeanImage(4, threshold=thr, energy=energy)
        for sl in (slice(0, 10), slice(10, 24)):
            ras.add(sr[sl], hr[sl])
            mi.add(sr[sl].numpy(), hr[sl].numpy())
        key = f"thr{thr}_e{int(energy)}"
        c[key] = {"raster_sr": torch.tensor(ras.get_hist()[0]), "raster_hr": torch.tensor(ras.get_hist()[1]),
                  "mean_sr": torch.tensor(mi.get_hist()[0]), "mean_hr": torch.tensor(mi.get_hist()[1]), "N": mi.N,
                  "energy_mode": bool(mi.energy)}
    c["hardest"] = {n: torch.tensor(np.asarray(get_nth_hardest(sr.numpy(), n=n), dtype=np.float32)) for n in (1, 5, 10, 20, 30)}
    # evaluation/eval.py:477-485
    l1_criterion = nn.L1Loss(reduction='none')
    pool = SumPool2d(4)
    gen_lr = pool(sr)
    c["lr_l1"] = torch.tensor(l1_criterion(gen_lr, lr).numpy().mean((1, 2, 3)))
    c["hr_l1"] = torch.tensor(l1_criterion(sr, hr).numpy().mean((1, 2, 3)))
    cases[f"seed{seed}"] = c
torch.save({"cases": cases}, os.path.join(HERE, "golden_evalmetrics_v1.pt"))
print({k: (v["thr0.7_e0"]["raster_sr"].sum().item(), v["lr_l1"][:2]) for k, v in cases.items()})
