"""Generate golden vectors by RUNNING THE REFERENCE (import from /root/reference).

Run in the build container only:  python tests/golden/make_golden.py
Writes tests/golden/golden_v1.pt (small).  Weights are not stored: they are
rebuilt anywhere from oracle.seeded_state_dict (deterministic per-tensor seeds)
and loaded into the reference modules with load_state_dict(strict=True), which
also pins the state_dict layout.
"""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from oracle import oracle as O  # noqa: E402
import models as R  # noqa: E402  (the reference)

torch.set_num_threads(8)


def summarize(t: torch.Tensor):
    t = t.detach().float().reshape(-1)
    return {"sum": t.double().sum().item(), "abssum": t.double().abs().sum().item(),
            "head": t[:64].clone(), "numel": t.numel()}


def gen_case(name, *, nb, nup, lr_size, batch, res_scale, thres, power, multiplier, training, seed, gain=1.0,
             with_grads=False):
    sd = O.make_generator_sd(seed, num_res_blocks=nb, num_upsample=nup, power=power, multiplier=multiplier, gain=gain)
    g = R.GeneratorRRDB(1, filters=64, num_res_blocks=nb, num_upsample=nup, power=power, multiplier=multiplier,
                        res_scale=res_scale)
    g.load_state_dict(sd, strict=True)
    g.thres = thres
    g.train(training)
    hr, lr = O.synthetic_jets(batch, lr_size * 2 ** nup, 2 ** nup, seed + 7)
    case = {"cfg": dict(nb=nb, nup=nup, lr_size=lr_size, batch=batch, res_scale=res_scale, thres=thres,
                        power=power, multiplier=multiplier, training=training, seed=seed, gain=gain)}
    if with_grads:
        out = g(lr)
        loss = torch.nn.functional.l1_loss(out, hr)          # the warm-up step loss, esrgan.py:424
        loss.backward()
        case["loss"] = loss.item()
        case["grads"] = {k: summarize(p.grad) for k, p in g.named_parameters() if p.grad is not None}
    else:
        with torch.no_grad():
            out = g(lr)
    case["out"] = out.detach().clone()
    case["srs"] = g.srs.detach().clone()
    return name, case


def disc_case(seed, batch):
    sd = O.make_discriminator_sd(seed, gain=2.0)
    d = R.Markovian_Discriminator((1, 80, 80), [16, 32, 32, 64])
    d.load_state_dict(sd, strict=True)
    hr, _ = O.synthetic_jets(batch, 80, 4, seed + 3)
    fake, _ = O.synthetic_jets(batch, 80, 4, seed + 4)
    fake = fake * 0.7 + 0.01
    pred_real = d(hr)
    pred_fake = d(fake)
    eps = 1e-7
    bce = torch.nn.BCEWithLogitsLoss()
    valid = torch.ones_like(pred_real)
    zeros = torch.zeros_like(pred_real)
    # esrgan.py:578-588
    loss = (bce(eps + pred_real - pred_fake.mean(0, keepdim=True), valid)
            + bce(eps + pred_fake - pred_real.mean(0, keepdim=True), zeros)) / 2
    g = torch.Generator().manual_seed(seed + 5)
    epsilon = torch.rand(batch, 1, 1, 1, generator=g)
    interp = (epsilon * hr + (1 - epsilon) * fake).requires_grad_(True)      # esrgan.py:598-600
    pred_i = d(interp)
    grads = torch.autograd.grad(pred_i, interp, valid, create_graph=True, retain_graph=True, only_inputs=True)[0]
    grads = grads.view(batch, -1)
    gp = ((grads.norm(2, dim=1) - 1) ** 2).mean() * 0.01 / 2
    total = loss + gp
    total.backward()
    return {"cfg": dict(seed=seed, batch=batch), "pred_real": pred_real.detach().clone(),
            "loss_ragan": loss.item(), "gp": gp.item(), "epsilon": epsilon,
            "grads": {k: p.grad.detach().clone() for k, p in d.named_parameters()},
            "out_shape": d.output_shape}


def main():
    cases = dict([
        gen_case("g_fwd_train_x4", nb=1, nup=2, lr_size=8, batch=2, res_scale=0.1, thres=0.05, power=1.0,
                 multiplier=1.0, training=True, seed=11),
        gen_case("g_fwd_eval_x4", nb=1, nup=2, lr_size=8, batch=2, res_scale=0.1, thres=0.05, power=1.0,
                 multiplier=1.0, training=False, seed=12),
        gen_case("g_fwd_eval_pow_x2", nb=1, nup=1, lr_size=10, batch=2, res_scale=0.2, thres=0.1, power=0.5,
                 multiplier=2.0, training=False, seed=13),
        gen_case("g_fwd_train_pow_x2", nb=1, nup=1, lr_size=10, batch=2, res_scale=0.2, thres=0.1, power=0.5,
                 multiplier=2.0, training=True, seed=14),
        gen_case("g_fwd_nb2_x8", nb=2, nup=3, lr_size=5, batch=3, res_scale=0.1, thres=0.0, power=1.0,
                 multiplier=1.0, training=False, seed=15, gain=1.5),
        gen_case("g_train_x4", nb=1, nup=2, lr_size=8, batch=2, res_scale=0.1, thres=0.0, power=1.0,
                 multiplier=1.0, training=True, seed=16, with_grads=True),
    ])
    cases["d_patch"] = disc_case(21, 3)
    x = torch.arange(2 * 1 * 8 * 8, dtype=torch.float32).reshape(2, 1, 8, 8)
    cases["sumpool4"] = {"x": x, "y": R.SumPool2d(4)(x)}
    torch.save(cases, os.path.join(HERE, "golden_v1.pt"))
    print("wrote", {k: (v["out"].shape if "out" in v else "-") for k, v in cases.items()})


if __name__ == "__main__":
    main()
