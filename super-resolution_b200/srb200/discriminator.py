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


def _raw_fwd(x, w, b, stride, slope=None):
    """conv (+ bias) (+ LeakyReLU(slope) fused in the kernel epilogue when slope is given)"""
    B, Cin, H, W = x.shape
    Cout = w.shape[0]
    Ho, Wo = (H - 1) // stride + 1, (W - 1) // stride + 1
    y = torch.empty(B, Cout, Ho, Wo, device=x.device)
    L.check(L.lib().srb_dconv_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr() if b is not None else None, y.data_ptr(),
                                  B, Cin, H, W, Cout, stride, int(slope is not None),
                                  float(slope) if slope is not None else 0.0, _st()), "dconv_fwd")
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


class ConvLReLU(torch.autograd.Function):
    """y = LeakyReLU(conv3x3(x, w) + b), one kernel (discriminator_block, models.py:142-145).  Backward is composed of
    the differentiable pieces LReLUBwd / ConvDgrad / ConvWgrad, so it can itself be differentiated."""

    @staticmethod
    def forward(ctx, x, w, b, stride, slope):
        _need_cuda(x)
        y = _raw_fwd(_f32(x), _f32(w), _f32(b) if b is not None else None, stride, slope)
        ctx.save_for_backward(x, w, y)
        ctx.stride, ctx.has_b, ctx.slope = stride, b is not None, slope
        return y

    @staticmethod
    def backward(ctx, gy):
        x, w, y = ctx.saved_tensors
        gz = LReLUBwd.apply(gy, y, ctx.slope)
        gx = ConvDgrad.apply(gz, w, x.shape, ctx.stride) if ctx.needs_input_grad[0] else None
        gw = gb = None
        if ctx.needs_input_grad[1] or (ctx.has_b and ctx.needs_input_grad[2]):
            gw, gb = ConvWgrad.apply(gz, x, w.shape, ctx.stride, ctx.has_b)
        return gx, gw, (gb if ctx.has_b else None), None, None


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


def _check_conv(conv: nn.Conv2d):
    if conv.kernel_size != (3, 3) or conv.padding != (1, 1) or conv.stride[0] != conv.stride[1] \
            or conv.stride[0] not in (1, 2):
        raise RuntimeError("CUDA srb200: discriminator kernels cover 3x3 / padding 1 / stride 1 or 2")


def conv2d(x, conv: nn.Conv2d, detach_params: bool = False, slope=None):
    _check_conv(conv)
    w, b = conv.weight, conv.bias
    if detach_params:
        w, b = w.detach(), (b.detach() if b is not None else None)
    if slope is not None:
        return ConvLReLU.apply(x, w, b, conv.stride[0], slope)
    return ConvFwd.apply(x, w, b, conv.stride[0])


def _layers(seq: nn.Sequential):
    """[(conv, slope-or-None)] for a Conv2d / LeakyReLU chain; None if the Sequential holds anything else."""
    mods, out, i = list(seq), [], 0
    while i < len(mods):
        m = mods[i]
        if not isinstance(m, nn.Conv2d):
            return None
        slope = None
        if i + 1 < len(mods) and isinstance(mods[i + 1], nn.LeakyReLU):
            slope = mods[i + 1].negative_slope
            i += 1
        out.append((m, slope))
        i += 1
    return out


def sequential_forward(seq: nn.Sequential, x, detach_params: bool = False):
    """Run an nn.Sequential of Conv2d / LeakyReLU (the discriminator_block pattern, models.py:140-146); each
    conv + LeakyReLU pair is one fused kernel.  ``detach_params`` treats the weights as constants (the generator
    step only needs d/d input)."""
    layers = _layers(seq)
    if layers is None:
        for m in seq:
            if isinstance(m, nn.Conv2d):
                x = conv2d(x, m, detach_params)
            elif isinstance(m, nn.LeakyReLU):
                x = LReLU.apply(x, m.negative_slope)
            else:
                x = m(x)
        return x
    for conv, slope in layers:
        x = conv2d(x, conv, detach_params, slope)
    return x


def input_gradient(seq: nn.Sequential, x):
    """d sum(D(x)) / d x as a function that is differentiable w.r.t. the weights: what
    autograd.grad(D(x), x, ones, create_graph=True) returns at esrgan.py:602, built explicitly from the input-gradient
    kernels so that no weight gradient is computed (and thrown away) on the way.  Returns (pred, grad_x)."""
    layers = _layers(seq)
    if layers is None:
        raise RuntimeError("CUDA srb200: input_gradient expects a Conv2d / LeakyReLU chain")
    acts, h = [], _f32(x)
    with torch.no_grad():
        for conv, slope in layers:
            _check_conv(conv)
            acts.append(h)
            h = _raw_fwd(h, _f32(conv.weight), _f32(conv.bias) if conv.bias is not None else None, conv.stride[0],
                         slope)
            acts.append(h)
    pred = h
    g = torch.ones_like(pred)
    for n in range(len(layers) - 1, -1, -1):
        conv, slope = layers[n]
        xin, y = acts[2 * n], acts[2 * n + 1]
        if slope is not None:
            g = LReLUBwd.apply(g, y, slope)
        g = ConvDgrad.apply(g, conv.weight, xin.shape, conv.stride[0])
    return pred, g


def set_impl(simple: bool) -> None:
    """Switch the discriminator convolutions to the simple cross-check kernels (tests) or back."""
    L.check(L.lib().srb_set_dconv_impl(1 if simple else 0), "set_dconv_impl")
