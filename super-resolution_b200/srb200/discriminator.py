"""Discriminator conv stack on the C ABI kernels (reference models.py:140-174).

Three linear kernels - forward conv, input gradient, weight gradient - are wrapped as autograd Functions whose
backward passes are again expressed through those same Functions, so autograd can differentiate them twice.  That is
what the gradient penalty needs (esrgan.py:598-606: autograd.grad(..., create_graph=True) then loss.backward()).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _lib as L


def _st() -> int:
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(t):
    if not t.is_cuda:
        raise RuntimeError("CUDA tensor required: super-resolution_b200 has no CPU path for the discriminator")


def _f32(t):
    return t.detach().contiguous().float()


def _raw_fwd(x, w, b, stride):
    B, Cin, H, W = x.shape
    Cout = w.shape[0]
    Ho, Wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    y = torch.empty(B, Cout, Ho, Wo, device=x.device)
    L.check(L.lib().srb_dconv_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr() if b is not None else None, y.data_ptr(),
                                  B, Cin, H, W, Cout, stride, _st()), "dconv_fwd")
    return y


def _raw_dgrad(gy, w, x_shape, stride):
    B, Cin, H, W = x_shape
    gx = torch.empty(x_shape, device=gy.device)
    L.check(L.lib().srb_dconv_dgrad(gy.data_ptr(), w.data_ptr(), gx.data_ptr(), B, Cin, H, W, w.shape[0], stride,
                                    _st()), "dconv_dgrad")
    return gx


def _raw_wgrad(gy, x, w_shape, stride, want_bias):
    B, Cin, H, W = x.shape
    dw = torch.empty(w_shape, device=x.device)
    db = torch.empty(w_shape[0], device=x.device) if want_bias else None
    L.check(L.lib().srb_dconv_wgrad(gy.data_ptr(), x.data_ptr(), dw.data_ptr(), db.data_ptr() if want_bias else None,
                                    B, Cin, H, W, w_shape[0], stride, _st()), "dconv_wgrad")
    return dw, db


class ConvFwd(torch.autograd.Function):
    """y = conv3x3(x, w, stride, pad 1) + b"""

    @staticmethod
    def forward(ctx, x, w, b, stride):
        _need_cuda(x)
        ctx.save_for_backward(x, w)
        ctx.stride, ctx.has_b = stride, b is not None
        return _raw_fwd(_f32(x), _f32(w), _f32(b) if b is not None else None, stride)

    @staticmethod
    def backward(ctx, gy):
        x, w = ctx.saved_tensors
        gx = ConvDgrad.apply(gy, w, x.shape, ctx.stride) if ctx.needs_input_grad[0] else None
        gw = gb = None
        if ctx.needs_input_grad[1] or (ctx.has_b and ctx.needs_input_grad[2]):
            gw, gb = ConvWgrad.apply(gy, x, w.shape, ctx.stride, ctx.has_b)
        return gx, gw, (gb if ctx.has_b else None), None


class ConvDgrad(torch.autograd.Function):
    """gx = conv_transpose(gy, w): linear in gy and in w."""

    @staticmethod
    def forward(ctx, gy, w, x_shape, stride):
        ctx.save_for_backward(gy, w)
        ctx.stride, ctx.x_shape = stride, x_shape
        return _raw_dgrad(_f32(gy), _f32(w), tuple(x_shape), stride)

    @staticmethod
    def backward(ctx, ggx):
        gy, w = ctx.saved_tensors
        g_gy = ConvFwd.apply(ggx, w, None, ctx.stride) if ctx.needs_input_grad[0] else None
        g_w = None
        if ctx.needs_input_grad[1]:
            g_w, _ = ConvWgrad.apply(gy, ggx, w.shape, ctx.stride, False)
        return g_gy, g_w, None, None


class ConvWgrad(torch.autograd.Function):
    """dw = sum gy (x) x, db = sum gy: linear in gy and in x."""

    @staticmethod
    def forward(ctx, gy, x, w_shape, stride, want_bias):
        ctx.save_for_backward(gy, x)
        ctx.stride, ctx.want_bias = stride, want_bias
        dw, db = _raw_wgrad(_f32(gy), _f32(x), tuple(w_shape), stride, want_bias)
        if not want_bias:
            db = torch.zeros(0, device=x.device)
            ctx.mark_non_differentiable(db)
        return dw, db

    @staticmethod
    def backward(ctx, g_dw, g_db):
        gy, x = ctx.saved_tensors
        g_gy = g_x = None
        if ctx.needs_input_grad[0]:
            g_gy = ConvFwd.apply(x, g_dw, g_db if ctx.want_bias else None, ctx.stride)
        if ctx.needs_input_grad[1]:
            g_x = ConvDgrad.apply(gy, g_dw, x.shape, ctx.stride)
        return g_gy, g_x, None, None, None


class LReLU(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, slope):
        _need_cuda(x)
        xc = _f32(x)
        y = torch.empty_like(xc)
        L.check(L.lib().srb_lrelu_fwd(xc.data_ptr(), y.data_ptr(), xc.numel(), float(slope), _st()), "lrelu_fwd")
        ctx.save_for_backward(y)
        ctx.slope = slope
        return y

    @staticmethod
    def backward(ctx, g):
        (y,) = ctx.saved_tensors
        return LReLUBwd.apply(g, y, ctx.slope), None


class LReLUBwd(torch.autograd.Function):
    """g * (ref > 0 ? 1 : slope): linear in g, piecewise constant in ref."""

    @staticmethod
    def forward(ctx, g, ref, slope):
        gc = _f32(g)
        out = torch.empty_like(gc)
        L.check(L.lib().srb_lrelu_bwd(gc.data_ptr(), ref.data_ptr(), out.data_ptr(), gc.numel(), float(slope), _st()),
                "lrelu_bwd")
        ctx.save_for_backward(ref)
        ctx.slope = slope
        return out

    @staticmethod
    def backward(ctx, gg):
        (ref,) = ctx.saved_tensors
        return LReLUBwd.apply(gg, ref, ctx.slope), None, None


def conv2d(x, conv: nn.Conv2d, detach_params: bool = False):
    if conv.kernel_size != (3, 3) or conv.padding != (1, 1) or conv.stride[0] != conv.stride[1] \
            or conv.stride[0] not in (1, 2):
        raise RuntimeError("CUDA srb200: discriminator kernels cover 3x3 / padding 1 / stride 1 or 2")
    w, b = conv.weight, conv.bias
    if detach_params:
        w, b = w.detach(), (b.detach() if b is not None else None)
    return ConvFwd.apply(x, w, b, conv.stride[0])


def sequential_forward(seq: nn.Sequential, x, detach_params: bool = False):
    """Run an nn.Sequential of Conv2d / LeakyReLU (the discriminator_block pattern, models.py:140-146).
    ``detach_params`` treats the weights as constants (the generator step only needs d/d input)."""
    for m in seq:
        if isinstance(m, nn.Conv2d):
            x = conv2d(x, m, detach_params)
        elif isinstance(m, nn.LeakyReLU):
            x = LReLU.apply(x, m.negative_slope)
        else:
            x = m(x)
    return x
