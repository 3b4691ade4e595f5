"""Loss reductions of the GAN step on the C ABI kernels (esrgan.py:471-508, 578-606)."""
from __future__ import annotations

import torch

from . import _lib as L

EPS = 1e-7   # esrgan.py:319


def _st() -> int:
    return torch.cuda.current_stream().cuda_stream


def _c(t):
    return t.detach().contiguous().float()


class PixelLossWork:
    """Scratch for srb_pixel_losses for one (B, H, W, factor)."""

    def __init__(self, B, H, W, k, dev):
        self.B, self.H, self.W, self.k = B, H, W, k
        self.msign = torch.empty(H * W, device=dev)
        self.dsign = torch.empty(B * (H // k) * (W // k), device=dev)
        self.acc = torch.zeros(2, device=dev)


def pixel_losses(gen, gt, lr, k, w_hr, w_lr, grad_out, accumulate, work: PixelLossWork, lr_power: float = 1.0):
    """loss_hr = L1(mean_b gen, mean_b gt) (esrgan.py:487); loss_lr = L1(SumPool_k(gen), lr) (471,490), or with
    ``lr_power`` p != 1 L1(SumPool_k(gen)^p, lr^p) (472-473: the scaling_power branch).
    Adds w_hr * d loss_hr/d gen + w_lr * d loss_lr/d gen into ``grad_out`` (or overwrites it).
    Returns (loss_hr, loss_lr) as 0-dim device tensors."""
    gen, gt, lr = _c(gen), _c(gt), _c(lr)
    B, Cc, H, W = gen.shape
    assert Cc == 1, "pixel loss kernel is single-channel (jet images)"
    L.check(L.lib().srb_pixel_losses_pow(gen.data_ptr(), gt.data_ptr(), lr.data_ptr(), B, H, W, k, float(w_hr),
                                         float(w_lr), float(lr_power), int(accumulate), work.msign.data_ptr(),
                                         work.dsign.data_ptr(), work.acc.data_ptr(),
                                         grad_out.data_ptr() if grad_out is not None else None, _st()), "pixel_losses")
    acc = work.acc.clone()
    return acc[0] / (H * W), acc[1] / (B * (H // k) * (W // k))


def ragan_raw(u, v):
    """One kernel: the relativistic-average loss of RaGAN below plus d loss/d u and d loss/d v (no autograd)."""
    uc, vc = _c(u), _c(v)
    B = uc.shape[0]
    n = uc.numel() // B
    loss = torch.empty((), device=u.device)
    gu, gv = torch.empty_like(uc), torch.empty_like(vc)
    L.check(L.lib().srb_ragan(uc.data_ptr(), vc.data_ptr(), B, n, EPS, loss.data_ptr(), gu.data_ptr(),
                              gv.data_ptr(), _st()), "ragan")
    return loss, gu, gv


def gradient_penalty_raw(g, coef):
    """((||g_b||_2 - 1)^2).mean() * coef and its gradient w.r.t. g (esrgan.py:604-605), two small kernels."""
    gc = _c(g)
    B = gc.shape[0]
    n = gc.numel() // B
    norms = torch.empty(B, device=g.device)
    loss = torch.zeros((), device=g.device)
    L.check(L.lib().srb_gp_fwd(gc.data_ptr(), B, n, float(coef), norms.data_ptr(), loss.data_ptr(), _st()), "gp_fwd")
    gg = torch.empty_like(gc)
    one = torch.ones((), device=g.device)
    L.check(L.lib().srb_gp_bwd(gc.data_ptr(), norms.data_ptr(), B, n, float(coef), one.data_ptr(), gg.data_ptr(),
                               _st()), "gp_bwd")
    return loss, gg


class RaGAN(torch.autograd.Function):
    """0.5*[BCE(eps + u - mean_b v, 1) + BCE(eps + v - mean_b u, 0)]  (esrgan.py:507-508 with u=fake, v=real;
    esrgan.py:578-588 with u=real, v=fake).  Loss and both gradients come out of one kernel."""

    @staticmethod
    def forward(ctx, u, v):
        uc, vc = _c(u), _c(v)
        B = uc.shape[0]
        n = uc.numel() // B
        loss = torch.empty((), device=u.device)
        gu, gv = torch.empty_like(uc), torch.empty_like(vc)
        L.check(L.lib().srb_ragan(uc.data_ptr(), vc.data_ptr(), B, n, EPS, loss.data_ptr(), gu.data_ptr(),
                                  gv.data_ptr(), _st()), "ragan")
        ctx.save_for_backward(gu, gv)
        return loss

    @staticmethod
    def backward(ctx, g):
        gu, gv = ctx.saved_tensors
        return (g * gu if ctx.needs_input_grad[0] else None), (g * gv if ctx.needs_input_grad[1] else None)


class GradientPenalty(torch.autograd.Function):
    """((||g_b||_2 - 1)^2).mean() * coef   (esrgan.py:604-605 with coef = lambda_reg / 2)."""

    @staticmethod
    def forward(ctx, g, coef):
        gc = _c(g)
        B = gc.shape[0]
        n = gc.numel() // B
        norms = torch.empty(B, device=g.device)
        loss = torch.zeros((), device=g.device)
        L.check(L.lib().srb_gp_fwd(gc.data_ptr(), B, n, float(coef), norms.data_ptr(), loss.data_ptr(), _st()),
                "gp_fwd")
        ctx.save_for_backward(gc, norms)
        ctx.coef = coef
        return loss

    @staticmethod
    def backward(ctx, gl):
        gc, norms = ctx.saved_tensors
        B = gc.shape[0]
        n = gc.numel() // B
        gg = torch.empty_like(gc)
        glc = gl.detach().float().contiguous()
        L.check(L.lib().srb_gp_bwd(gc.data_ptr(), norms.data_ptr(), B, n, float(ctx.coef), glc.data_ptr(),
                                   gg.data_ptr(), _st()), "gp_bwd")
        return gg, None


# ---------------------------------------------------------------------------------------------- optional terms
class OptionalLossWork:
    """Scratch + constants of the optional generator loss terms (csrc/opt_losses.cu) for one (B, H, W)."""

    def __init__(self, B, H, W, factor, dev, binedges=None):
        self.B, self.H, self.W, self.f = B, H, W, factor
        self.per_image = torch.empty(B, device=dev)
        self.hit = torch.zeros(2 * factor * factor, device=dev)
        self.bins = None
        if binedges is not None:       # per branch k: (centers, deltas) of DiffableHistogram built from the bin edges
            self.bins = []
            for e in binedges:
                e = torch.as_tensor(e, dtype=torch.float32, device=dev)
                d = (e[1:] - e[:-1]).contiguous()
                self.bins.append(dict(centers=(e[:-1] + 0.5 * d).contiguous(), deltas=d, n=int(d.numel()),
                                      hq=torch.zeros(d.numel(), device=dev), hp=torch.zeros(d.numel(), device=dev),
                                      dq=torch.zeros(d.numel(), device=dev)))


def nnz_loss(gen, gt, w, grad, work: OptionalLossWork, sigma=50000.0):
    """esrgan.py:530-533; adds w * d loss / d gen into ``grad``."""
    g, t = _c(gen), _c(gt)
    B = g.shape[0]
    loss = torch.empty((), device=g.device)
    L.check(L.lib().srb_nnz_loss(g.data_ptr(), t.data_ptr(), B, g[0].numel(), float(sigma), float(w), 1,
                                 work.per_image.data_ptr(), loss.data_ptr(), grad.data_ptr(), _st()), "nnz_loss")
    return loss


def mask_loss(gen, gt, w, grad, sigma=5e4):
    """esrgan.py:534-537."""
    g, t = _c(gen), _c(gt)
    loss = torch.empty((), device=g.device)
    L.check(L.lib().srb_mask_loss(g.data_ptr(), t.data_ptr(), g.numel(), float(sigma), float(w), 1, loss.data_ptr(),
                                  grad.data_ptr(), _st()), "mask_loss")
    return loss


def hit_loss(gen, gt, factor, thr, sig, w, grad, work: OptionalLossWork):
    """esrgan.py:546-550."""
    g, t = _c(gen), _c(gt)
    B, _, H, W = g.shape
    loss = torch.empty((), device=g.device)
    L.check(L.lib().srb_hit_loss(g.data_ptr(), t.data_ptr(), B, H, W, int(factor), float(thr), float(sig), float(w), 1,
                                 work.hit.data_ptr(), loss.data_ptr(), grad.data_ptr(), _st()), "hit_loss")
    return loss


def hist_loss(gen, gt, k, sigma, w, grad, work: OptionalLossWork):
    """esrgan.py:538-545: KLD_hist(DiffableHistogram(gen[gen > 0]), DiffableHistogram(gt[gt > 0]))."""
    g, t = _c(gen), _c(gt)
    b = work.bins[k]
    lib = L.lib()
    L.check(lib.srb_soft_hist(g.data_ptr(), g.numel(), b["centers"].data_ptr(), b["deltas"].data_ptr(), b["n"],
                              float(sigma), b["hq"].data_ptr(), _st()), "soft_hist")
    L.check(lib.srb_soft_hist(t.data_ptr(), t.numel(), b["centers"].data_ptr(), b["deltas"].data_ptr(), b["n"],
                              float(sigma), b["hp"].data_ptr(), _st()), "soft_hist")
    loss = torch.empty((), device=g.device)
    L.check(lib.srb_kld_hist(b["hq"].data_ptr(), b["hp"].data_ptr(), b["deltas"].data_ptr(), b["n"], loss.data_ptr(),
                             b["dq"].data_ptr(), _st()), "kld_hist")
    L.check(lib.srb_soft_hist_bwd(g.data_ptr(), g.numel(), b["centers"].data_ptr(), b["deltas"].data_ptr(), b["n"],
                                  float(sigma), b["dq"].data_ptr(), float(w), 1, grad.data_ptr(), _st()), "soft_hist_bwd")
    return loss


class SinkhornWork:
    """Workspace of srb_sinkhorn_loss for one (B, H*W): sorted values, the duals of every iteration, adjoints."""

    def __init__(self, B, HW, dev, max_iter=100, cap1=None, cap2=None):
        if cap1 is None:
            cap1 = min(HW, 12288)
        if cap2 is None:
            cap2 = min(HW, (220 * 1024 - 12 * cap1) // 16)
        self.B, self.HW, self.max_iter, self.cap1, self.cap2 = B, HW, int(max_iter), int(cap1), int(cap2)
        n = L.lib().srb_sinkhorn_workspace_bytes(B, HW, self.cap1, self.cap2, self.max_iter)
        if n == 0:
            raise RuntimeError("CUDA srb200: bad Sinkhorn workspace arguments")
        self.ws = torch.empty(n, dtype=torch.uint8, device=dev)
        self.info = torch.zeros(4, dtype=torch.int32, device=dev)


def wasserstein_loss(gen, gt, eps, w, grad, work: SinkhornWork, thresh=0.1):
    """esrgan.py:539-542: SinkhornDistance(eps, 100, 'sum') between the per-image sorted pixel values, cut to the
    batch's largest positive count; adds w * d loss / d gen into ``grad`` (None: value only).  ``work.info`` then holds
    (points generated, points true, iterations run, capacity overflow)."""
    g, t = _c(gen), _c(gt)
    B = g.shape[0]
    HW = g[0].numel()
    assert (B, HW) == (work.B, work.HW)
    loss = torch.empty((), device=g.device)
    L.check(L.lib().srb_sinkhorn_loss(g.data_ptr(), t.data_ptr(), B, HW, float(eps), work.max_iter, float(thresh),
                                      work.cap1, work.cap2, float(w), work.ws.data_ptr(), work.ws.numel(),
                                      loss.data_ptr(), grad.data_ptr() if grad is not None else None,
                                      work.info.data_ptr(), _st()), "sinkhorn_loss")
    return loss
