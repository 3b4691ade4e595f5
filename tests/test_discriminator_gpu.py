"""GPU parity: discriminator conv stack (fwd, bwd, double backward through the gradient penalty) and SumPool2d
vs the golden vectors the reference produced.  fp32 kernels: tolerance 1e-4 relative."""
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda:0")


def test_patch_discriminator_losses_and_grads(golden):
    dev = _dev()
    import models
    case = golden["d_patch"]
    seed, batch = case["cfg"]["seed"], case["cfg"]["batch"]
    d = models.Markovian_Discriminator((1, 80, 80), [16, 32, 32, 64])
    d.load_state_dict(O.make_discriminator_sd(seed, gain=2.0), strict=True)
    d = d.to(dev)
    hr, _ = O.synthetic_jets(batch, 80, 4, seed + 3)
    fake, _ = O.synthetic_jets(batch, 80, 4, seed + 4)
    fake = fake * 0.7 + 0.01
    hr, fake = hr.to(dev), fake.to(dev)
    pred_real = d(hr)
    torch.testing.assert_close(pred_real.cpu(), case["pred_real"], rtol=1e-4, atol=1e-5)
    pred_fake = d(fake)
    loss = O.ragan_loss_d(pred_real, pred_fake)
    assert abs(loss.item() - case["loss_ragan"]) < 1e-5
    eps = case["epsilon"].to(dev)
    interp = (eps * hr + (1 - eps) * fake).requires_grad_(True)
    pred_i = d(interp)
    grads = torch.autograd.grad(pred_i, interp, torch.ones_like(pred_i), create_graph=True, retain_graph=True)[0]
    gp = ((grads.view(batch, -1).norm(2, dim=1) - 1) ** 2).mean() * 0.01 / 2
    assert abs(gp.item() - case["gp"]) <= 1e-4 * max(1.0, case["gp"])
    (loss + gp).backward()
    for k, p in d.named_parameters():
        ref = case["grads"][k]
        err = (p.grad.cpu() - ref).abs().max().item()
        assert err <= 1e-3 * ref.abs().max().item() + 1e-6, (k, err, ref.abs().max().item())


def test_sumpool(golden):
    dev = _dev()
    import models
    c = golden["sumpool4"]
    x = c["x"].to(dev).requires_grad_(True)
    y = models.SumPool2d(4)(x)
    torch.testing.assert_close(y.cpu(), c["y"])
    y.sum().backward()
    assert torch.equal(x.grad.cpu(), torch.ones_like(c["x"]))


@pytest.mark.parametrize("B,Cin,Cout,H,W,s", [(3, 1, 16, 80, 80, 1), (2, 16, 16, 80, 80, 2), (2, 16, 32, 13, 9, 1),
                                              (2, 32, 32, 13, 9, 2), (2, 64, 64, 10, 10, 2), (3, 64, 1, 5, 5, 1),
                                              (2, 32, 64, 10, 10, 1), (1, 16, 16, 7, 6, 2)])
@pytest.mark.parametrize("simple", [False, True])
def test_dconv_kernels_vs_torch(B, Cin, Cout, H, W, s, simple):
    """forward (+ fused LeakyReLU), input gradient and weight gradient of the tuned and the simple kernels."""
    dev = _dev()
    import torch.nn.functional as F
    from srb200 import discriminator as D
    D.set_impl(simple)
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False          # the torch reference must be true fp32
    try:
        g = torch.Generator().manual_seed(B * 1000 + Cin + Cout + H)
        x = torch.randn(B, Cin, H, W, generator=g).to(dev).requires_grad_(True)
        w = (torch.randn(Cout, Cin, 3, 3, generator=g) * 0.2).to(dev).requires_grad_(True)
        b = torch.randn(Cout, generator=g).to(dev).requires_grad_(True)
        xr, wr, br = [t.detach().clone().requires_grad_(True) for t in (x, w, b)]
        y = D.ConvLReLU.apply(x, w, b, s, 0.2)
        yr = F.leaky_relu(F.conv2d(xr, wr, br, s, 1), 0.2)
        torch.testing.assert_close(y, yr, rtol=1e-4, atol=1e-4)
        gy = torch.randn(yr.shape, generator=g).to(dev)
        y.backward(gy); yr.backward(gy)
        torch.testing.assert_close(x.grad, xr.grad, rtol=1e-4, atol=1e-4)
        torch.testing.assert_close(w.grad, wr.grad, rtol=1e-4, atol=2e-4 * max(1.0, wr.grad.abs().max().item()))
        torch.testing.assert_close(b.grad, br.grad, rtol=1e-4, atol=2e-4 * max(1.0, br.grad.abs().max().item()))
    finally:
        D.set_impl(False)
        torch.backends.cudnn.allow_tf32 = tf32


def test_input_gradient_matches_autograd():
    dev = _dev()
    import models
    from srb200 import discriminator as D
    d = models.Markovian_Discriminator((1, 80, 80), [16, 32, 32, 64])
    d.load_state_dict(O.make_discriminator_sd(9, gain=2.0))
    d = d.to(dev)
    x, _ = O.synthetic_jets(3, 80, 4, 10)
    x = (x + 0.01).to(dev)
    torch.backends.cudnn.allow_tf32 = False
    xa = x.clone().requires_grad_(True)
    pred = d(xa)
    (ga,) = torch.autograd.grad(pred, xa, torch.ones_like(pred), create_graph=True)
    pred2, gb = D.input_gradient(d.model, x)
    torch.testing.assert_close(pred2, pred.detach(), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(gb.detach(), ga.detach(), rtol=1e-4, atol=1e-7)
    # second order: d/dw of a function of the input gradient
    la = (ga ** 2).sum(); lb = (gb ** 2).sum()
    grads_a = torch.autograd.grad(la, list(d.parameters()), allow_unused=True)
    grads_b = torch.autograd.grad(lb, list(d.parameters()), allow_unused=True)
    for (n, _), a, b in zip(d.named_parameters(), grads_a, grads_b):
        if a is None or b is None:
            assert a is None and b is None or (a if a is not None else b).abs().max().item() == 0, n
            continue
        torch.testing.assert_close(b, a, rtol=1e-3, atol=1e-6 + 1e-4 * a.abs().max().item())
