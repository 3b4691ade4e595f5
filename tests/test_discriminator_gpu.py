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
