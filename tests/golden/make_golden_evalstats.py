"""Golden vectors for the evaluation statistics: runs the reference's own ``get_nth_hardest`` (evaluation/eval.py:98-105).
eval.py cannot be imported here (matplotlib, energyflow, a CWD-relative options file), so the UNMODIFIED function is
taken out of the reference file with ``ast`` and executed as is (nothing is copied into this repo).
Build container only:  python tests/golden/make_golden_evalstats.py  ->  tests/golden/golden_evalstats_v1.pt"""
import ast
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402

src = open("/root/reference/evaluation/eval.py").read()
fn = next(n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == "get_nth_hardest")
ns = {"np": np}
exec(compile(ast.Module(body=[fn], type_ignores=[]), "eval.py", "exec"), ns)
get_nth_hardest = ns["get_nth_hardest"]


def images(seed, B=24):
    hr, _ = O.synthetic_jets(B, 80, 4, seed)
    g = torch.Generator().manual_seed(seed + 1)
    x = hr + 0.01 * torch.randn(hr.shape, generator=g) * (torch.rand(hr.shape, generator=g) < 0.05)   # SR-like residue
    x[3] = 0.0                                       # empty image
    x[4, 0, 5, 5] = x[4, 0, 9, 9] = x[4].max() + 1   # tie for the hardest constituent
    return x


cases = {}
for seed in (21, 22):
    x = images(seed)
    cases[f"seed{seed}"] = {"seed": seed, "hardest": {n: torch.tensor(np.asarray(get_nth_hardest(x.numpy(), n=n), dtype=np.float32))
                                                      for n in (1, 2, 3, 4, 8)}}
torch.save({"cases": cases}, os.path.join(HERE, "golden_evalstats_v1.pt"))
print({k: v["hardest"][1][:4] for k, v in cases.items()})
