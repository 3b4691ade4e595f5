"""Small host-side wrappers: SumPool2d (models.py:297-305) and the stand-alone DenseResidualBlock forward."""
from __future__ import annotations

import torch

from . import _lib as L
from . import ops


def _st() -> int:
    return torch.cuda.current_stream().cuda_stream


class SumPoolFn(torch.autograd.Function):
    """k x k block sum, stride k (what k*k*AvgPool2d(k) computes); linear, so backward is a broadcast."""

    @staticmethod
    def forward(ctx, x, k):
        if not x.is_cuda:
            raise RuntimeError("CUDA tensor required: super-resolution_b200 has no CPU path for SumPool2d")
        xc = x.detach().contiguous().float()
        B, Cc, H, W = xc.shape
        y = torch.empty(B, Cc, H // k, W // k, device=x.device)
        L.check(L.lib().srb_sumpool_fwd(xc.data_ptr(), y.data_ptr(), B * Cc, H, W, k, _st()), "sumpool_fwd")
        ctx.k, ctx.shape = k, xc.shape
        return y

    @staticmethod
    def backward(ctx, gy):
        return SumPoolBwdFn.apply(gy, ctx.k, ctx.shape), None


class SumPoolBwdFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gy, k, shape):
        g = gy.detach().contiguous().float()
        B, Cc, H, W = shape
        gx = torch.empty(shape, device=gy.device)
        L.check(L.lib().srb_sumpool_bwd(g.data_ptr(), gx.data_ptr(), B * Cc, H, W, k, _st()), "sumpool_bwd")
        ctx.k = k
        return gx

    @staticmethod
    def backward(ctx, ggx):
        return SumPoolFn.apply(ggx, ctx.k), None, None


def sum_pool(x, k):
    return SumPoolFn.apply(x, k)


def dense_block_forward(block, x):
    """DenseResidualBlock.forward used on its own (models.py:34-41), inference only: NCHW float in/out, run on the
    same conv kernel through a temporary growth buffer.  The generator never takes this path."""
    if torch.is_grad_enabled() and any(p.requires_grad for p in block.parameters()):
        raise RuntimeError("CUDA srb200: stand-alone DenseResidualBlock is inference-only; train through GeneratorRRDB")
    if block.drop and block.training:
        raise RuntimeError("CUDA srb200: Dropout2d variant is constructor-surface only in this build")
    B, F_, H, W = x.shape
    g = ops.to_pnhwc(x, 5 * F_, 0)
    out = ops.pnhwc_zeros(B, H, W, F_, x.device)
    for k in range(1, 6):
        conv = getattr(block, "b%d" % k)[0]
        wp = ops.pack_fwd(conv.weight)
        b = conv.bias.detach().float().contiguous()
        if k < 5:
            ops.conv3x3(g, 0, k * F_, wp, b, g, k * F_, F_, act=L.ACT_LRELU, slope=0.01)
        else:
            ops.conv3x3(g, 0, 5 * F_, wp, b, out, 0, F_, alpha=float(block.res_scale), res1=(g, 0, 1.0))
    return ops.from_pnhwc(out, 0, F_)
