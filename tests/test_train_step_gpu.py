"""GPU parity of the full GAN iteration (esrgan.py:399-626 default path) and of the fused Adam, against the CPU
oracle.  bf16 generator vs fp32 oracle: loss components within 2e-2 relative, generator gradients cosine >= 0.999,
discriminator gradients (fp32 kernels fed bf16-generated images) cosine >= 0.99."""
import copy

import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda:0")


def _cos(a, b):
    a, b = a.float().cpu().reshape(-1), b.float().cpu().reshape(-1)
    return torch.dot(a, b).item() / (a.norm().item() * b.norm().item() + 1e-30)


def test_fused_adam_matches_torch_and_gate():
    dev = _dev()
    from srb200.optim import FusedAdam
    g = torch.Generator().manual_seed(3)
    shapes = [(64, 3, 3, 3), (7,), (33, 5)]
    ps = [torch.nn.Parameter(torch.randn(s, generator=g).to(dev)) for s in shapes]
    ref = [torch.nn.Parameter(p.detach().cpu().clone()) for p in ps]
    opt = FusedAdam(ps, lr=2e-4, betas=(0.9, 0.999), weight_decay=0.01)
    ropt = torch.optim.Adam(ref, lr=2e-4, betas=(0.9, 0.999), weight_decay=0.01)
    for step in range(4):
        for p, r in zip(ps, ref):
            gr = torch.randn(r.shape, generator=g)
            p.grad, r.grad = gr.to(dev), gr.clone()
        opt.step()
        ropt.step()
    for p, r in zip(ps, ref):
        torch.testing.assert_close(p.detach().cpu(), r.detach(), rtol=1e-5, atol=1e-7)
    # gate: loss below the threshold -> nothing moves, step count frozen
    before = [p.detach().clone() for p in ps]
    opt.step(gate=(torch.tensor(0.0005, device=dev), 0.001))
    assert all(torch.equal(a, p.detach()) for a, p in zip(before, ps))
    assert opt.dev_state[0].item() == 4.0
    opt.step(gate=(torch.tensor(0.5, device=dev), 0.001))
    assert opt.dev_state[0].item() == 5.0 and not torch.equal(before[0], ps[0].detach())


def test_fused_adam_lr_lives_on_the_device_and_schedulers_wrap_it():
    """esrgan.py:308-310,666-667: ReduceLROnPlateau lowers the learning rate of optimizer_G / optimizer_D.  FusedAdam is
    a torch Optimizer (the scheduler wraps it) and reads lr from device state, so a CUDA graph captured around step()
    follows the schedule."""
    dev = _dev()
    from srb200.optim import FusedAdam
    g = torch.Generator().manual_seed(5)
    w0 = torch.randn(257, generator=g)
    grads = [torch.randn(257, generator=g) for _ in range(6)]
    p = torch.nn.Parameter(w0.clone().to(dev))
    r = torch.nn.Parameter(w0.clone())
    opt = FusedAdam([p], lr=1e-2)
    ropt = torch.optim.Adam([r], lr=1e-2)
    sched = torch.optim.lr_scheduler.ReduceLROnPlateau(opt, patience=0, factor=0.1)
    rsched = torch.optim.lr_scheduler.ReduceLROnPlateau(ropt, patience=0, factor=0.1)
    static_g = torch.zeros(257, device=dev)
    p.grad = static_g
    opt.step()                                   # warm-up outside the capture (also the first reference step)
    r.grad = torch.zeros(257); ropt.step()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        opt.step()
    for i, gr in enumerate(grads):
        if i == 3:                               # two non-improving metrics -> lr drops by 10x on both
            for s_ in (sched, rsched):
                s_.step(1.0); s_.step(2.0)
            assert opt.param_groups[0]["lr"] == pytest.approx(1e-3)
        static_g.copy_(gr.to(dev))
        opt.sync_lr()
        graph.replay()
        r.grad = gr.clone(); ropt.step()
    torch.testing.assert_close(p.detach().cpu(), r.detach(), rtol=2e-5, atol=1e-6)
    sd = opt.state_dict()
    opt2 = FusedAdam([torch.nn.Parameter(w0.clone().to(dev))], lr=5.0)
    opt2.load_state_dict(sd)
    assert opt2.param_groups[0]["lr"] == pytest.approx(1e-3) and opt2.dev_state[2].item() == pytest.approx(1e-3)


def test_loss_kernels_match_oracle():
    dev = _dev()
    from srb200 import losses as LS
    g = torch.Generator().manual_seed(4)
    B = 6
    u = torch.randn(B, 1, 5, 5, generator=g); v = torch.randn(B, 1, 5, 5, generator=g) * 2
    ur, vr = u.clone().requires_grad_(True), v.clone().requires_grad_(True)
    ref = O.ragan_loss_g(vr, ur)       # oracle signature (pred_real, pred_fake) with u = fake
    ref.backward()
    ud, vd = u.to(dev).requires_grad_(True), v.to(dev).requires_grad_(True)
    got = LS.RaGAN.apply(ud, vd)
    got.backward()
    assert abs(got.item() - ref.item()) < 1e-6
    torch.testing.assert_close(ud.grad.cpu(), ur.grad, rtol=1e-4, atol=1e-7)
    torch.testing.assert_close(vd.grad.cpu(), vr.grad, rtol=1e-4, atol=1e-7)
    # gradient penalty reduction
    gi = torch.randn(B, 640, generator=g) * 0.05
    gr = gi.clone().requires_grad_(True)
    refp = ((gr.norm(2, dim=1) - 1) ** 2).mean() * 0.005
    refp.backward()
    gd = gi.to(dev).requires_grad_(True)
    gotp = LS.GradientPenalty.apply(gd, 0.005)
    gotp.backward()
    assert abs(gotp.item() - refp.item()) < 1e-6
    torch.testing.assert_close(gd.grad.cpu(), gr.grad, rtol=1e-4, atol=1e-8)
    # pixel losses + gradient
    hr, lr = O.synthetic_jets(B, 32, 4, 5)
    gen = (hr * 0.8 + 0.02 * torch.randn(hr.shape, generator=g)).requires_grad_(True)
    l_hr = torch.nn.functional.l1_loss(gen.mean(0)[None], hr.mean(0)[None])
    l_lr = torch.nn.functional.l1_loss(O.sum_pool(gen, 4), lr)
    (0.3 * l_hr + 0.7 * l_lr).backward()
    work = LS.PixelLossWork(B, 32, 32, 4, dev)
    grad = torch.empty(B, 1, 32, 32, device=dev)
    a, b = LS.pixel_losses(gen.detach().to(dev), hr.to(dev), lr.to(dev), 4, 0.3, 0.7, grad, False, work)
    assert abs(a.item() - l_hr.item()) < 1e-6 and abs(b.item() - l_lr.item()) < 1e-5
    torch.testing.assert_close(grad.cpu(), gen.grad, rtol=1e-4, atol=1e-8)


def test_gan_step_matches_oracle():
    dev = _dev()
    import models
    from srb200.train_step import GANTrainStep
    B, factor = 4, 4
    cfg = dict(nb=1, nup=2, res_scale=0.1)
    sd_g = O.make_generator_sd(41, num_res_blocks=1, num_upsample=2, gain=1.0)
    sd_d = [O.make_discriminator_sd(42, gain=2.0), O.make_discriminator_sd(43, gain=2.0)]
    hr, lr = O.synthetic_jets(B, 80, factor, 44)
    gen_rng = torch.Generator().manual_seed(45)
    epsilons = [torch.rand(B, 1, 1, 1, generator=gen_rng) for _ in range(2)]

    # ---- ours
    G = models.GeneratorRRDB(1, filters=64, num_res_blocks=1, num_upsample=2, res_scale=0.1)
    G.load_state_dict(sd_g); G = G.to(dev).train()
    Ds = []
    for k in range(2):
        d = models.Markovian_Discriminator((1, 80, 80), [16, 32, 32, 64]); d.load_state_dict(sd_d[k]); Ds.append(d.to(dev))
    tr = GANTrainStep(G, Ds, factor=factor, lambdas=(0.2, 1.0), lambda_hr=1.0, lambda_lr=0.1, lambda_adv=0.01,
                      lambda_reg=0.01, d_threshold=0.001, lr=2e-4)
    losses = tr.step(lr.to(dev), hr.to(dev), epsilons=epsilons)
    g_grads = {n: gv.clone() for (n, p), gv in zip(((n, p) for n, p in G.named_parameters() if p.requires_grad),
                                                   tr.opt_G.grad_views)}
    d_grads = [[gv.clone() for gv in tr.opt_D[k].grad_views] for k in range(2)]

    # ---- oracle (CPU, fp32)
    sg = {k: (v.clone().requires_grad_(True) if k not in ("power", "multiplier") else v) for k, v in sd_g.items()}
    out, srs = O.generator_forward(sg, lr, res_scale=0.1, thres=0.0, training=True)
    loss_g, parts = O.generator_loss((out, srs), hr, lr, sd_d, factor=factor)
    loss_g.backward()
    for key in ("hr0", "lr0", "adv0", "hr1", "adv1"):
        ref = parts[key].item()
        assert abs(losses[key].item() - ref) <= 2e-2 * abs(ref) + 1e-6, (key, losses[key].item(), ref)
    assert abs(losses["g_loss"].item() - loss_g.item()) <= 2e-2 * abs(loss_g.item())
    for n, gr in g_grads.items():
        assert _cos(gr, sg[n].grad) >= 0.99, (n, _cos(gr, sg[n].grad))   # L1 terms: sign() gradients, see test_generator_gpu
    for k in range(2):
        sdk = {n: v.clone().requires_grad_(True) for n, v in sd_d[k].items()}
        fake = (out, srs)[k].detach()
        loss_d, dparts = O.discriminator_loss(sdk, hr, fake, epsilons[k], lambda_reg=0.01)
        loss_d.backward()
        assert abs(losses[f"d_loss{k}"].item() - loss_d.item()) <= 2e-2 * abs(loss_d.item()), (k,)
        assert abs(losses[f"gp{k}"].item() - dparts["gp"].item()) <= 5e-2 * abs(dparts["gp"].item()) + 1e-7
        for (n, v), gv in zip(sdk.items(), d_grads[k]):
            if v.grad.norm().item() < 1e-6:      # last-layer bias: cancels analytically in the relativistic loss
                assert gv.norm().item() < 1e-5
                continue
            assert _cos(gv, v.grad) >= 0.99, (k, n, _cos(gv, v.grad))
    # parameters actually moved (Adam applied; loss_D is far above d_threshold here)
    assert not torch.equal(G.conv2.weight.detach().cpu(), sd_g["conv2.weight"])
    assert not torch.equal(Ds[0].model[0].weight.detach().cpu(), sd_d[0]["model.0.weight"])


def test_graphed_step_matches_eager():
    """The CUDA-graph replay of the iteration trains exactly like the eagerly launched one (gradient penalty off so
    that no random draw is involved; tolerance covers the order of the fp32 atomics in the weight-gradient kernels)."""
    dev = _dev()
    import models
    from srb200.train_step import GANTrainStep, GraphedStep
    hr, lr = O.synthetic_jets(4, 80, 4, 91)
    hr, lr = hr.to(dev), lr.to(dev)

    def make():
        G = models.GeneratorRRDB(1, filters=64, num_res_blocks=1, num_upsample=2, res_scale=0.1)
        G.load_state_dict(O.make_generator_sd(92, num_res_blocks=1, num_upsample=2)); G = G.to(dev).train()
        Ds = []
        for k in range(2):
            d = models.Markovian_Discriminator((1, 80, 80), [16, 32, 32, 64])
            d.load_state_dict(O.make_discriminator_sd(93 + k, gain=2.0)); Ds.append(d.to(dev))
        return GANTrainStep(G, Ds, factor=4, lambda_reg=0.0)

    eager = make()
    for _ in range(4):
        le = eager.step(lr, hr)
    gr = GraphedStep(make(), lr, hr, warmup=2)        # two eager steps inside, then the capture (not executed)
    for _ in range(2):
        lg = gr.step(lr, hr)
    torch.cuda.synchronize()
    for k in ("g_loss", "d_loss0", "d_loss1", "hr0", "adv1"):
        assert abs(le[k].item() - lg[k].item()) <= 1e-3 * abs(le[k].item()) + 1e-6, (k, le[k].item(), lg[k].item())
    # Adam moves every weight by ~lr = 2e-4 per step in the direction of sign-like ratios m/sqrt(v): where a gradient
    # is at rounding-noise level the fp32 atomics' order can flip that direction, so single weights may differ by a few
    # lr; the bulk must agree.
    for a, b in ((eager.opt_G.flat_p, gr.tr.opt_G.flat_p), (eager.opt_D[0].flat_p, gr.tr.opt_D[0].flat_p)):
        d = (a - b).abs()
        assert d.max().item() <= 4 * 4 * 2e-4, d.max().item()
        assert d.mean().item() <= 2e-5, d.mean().item()
