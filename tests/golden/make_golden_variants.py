"""Golden vectors for the GeneratorRRDB constructor variants, produced by RUNNING THE REFERENCE
modules (import from /root/reference): transposed-conv upsampling (models.py:70-83), num_final_layer_res
(models.py:93-94,128-130) and Dropout2d (models.py:29-31,39-40).  Train-mode forward, loss = mean(out^2), backward;
plus an eval-mode forward.  Dropout masks come from torch.manual_seed(seed) exactly as nn.Dropout2d draws them; the
oracle redraws them the same way (oracle.DropMasks), which is how the GPU tests get the same masks.
Build container only:  python tests/golden/make_golden_variants.py  ->  tests/golden/golden_variants_v1.pt"""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")
from oracle import oracle as O  # noqa: E402
import models as R  # noqa: E402

CASES = {
    "tconv_alternating": dict(nb=1, nup=2, lr=10, batch=3, kw=dict(use_transposed_conv=True)),
    "tconv_full": dict(nb=1, nup=2, lr=10, batch=3, kw=dict(fully_tconv_upsample=True)),
    "final_res_1": dict(nb=1, nup=1, lr=12, batch=3, kw=dict(num_final_layer_res=1)),
    "final_res_2_tconv": dict(nb=2, nup=2, lr=10, batch=6, kw=dict(num_final_layer_res=2, use_transposed_conv=True)),
    "dropout": dict(nb=2, nup=1, lr=16, batch=10, kw=dict(drop_rate=0.3)),
    "dropout_final": dict(nb=1, nup=1, lr=10, batch=3, kw=dict(drop_rate=0.25, num_final_layer_res=1)),
}


def sel(names):
    keep = [n for n in names if n.startswith(("upsampling.", "conv2.")) or n.endswith(".dense_blocks.2.b5.0.bias")
            or (".0.dense_blocks.0.b1.0." in n)]
    return keep


def pack_grad(g):
    s_ = g.abs().max().item()
    return {"scale": s_, "q": (g / (s_ if s_ > 0 else 1.0)).half()}


def run(name, c, seed):
    kw = dict(c["kw"])
    drop = kw.get("drop_rate", 0)
    sd = O.make_generator_sd(seed, num_res_blocks=c["nb"], num_upsample=c["nup"], gain=1.2,
                             **{k: v for k, v in kw.items() if k != "drop_rate"})
    g = R.GeneratorRRDB(1, filters=64, num_res_blocks=c["nb"], num_upsample=c["nup"], res_scale=0.1, **kw)
    g.load_state_dict(sd, strict=True)
    _, lr = O.synthetic_jets(c["batch"], c["lr"] << c["nup"], 1 << c["nup"], seed + 1)
    g.train()
    torch.manual_seed(seed + 2)
    out = g(lr)
    (out * out).mean().backward()
    grads = {n: p.grad.detach().clone() for n, p in g.named_parameters() if p.grad is not None}
    g.eval()
    with torch.no_grad():
        ev = g(lr)
    return {"cfg": dict(nb=c["nb"], nup=c["nup"], lr=c["lr"], batch=c["batch"], seed=seed, kw=kw, drop=drop),
            "train_out": out.detach().clone(), "eval_out": ev.clone(),
            "grad_norms": {n: v.norm().item() for n, v in grads.items()},
            "grads": {n: pack_grad(grads[n]) for n in sel(grads)}}


if __name__ == "__main__":
    res = {name: run(name, c, 50 + i) for i, (name, c) in enumerate(CASES.items())}
    torch.save(res, os.path.join(HERE, "golden_variants_v1.pt"))
    print({k: (tuple(v["train_out"].shape), len(v["grads"])) for k, v in res.items()},
          os.path.getsize(os.path.join(HERE, "golden_variants_v1.pt")) / 1e6, "MB")
