"""Tensor-level wrappers over the C ABI.  torch is used for device memory and streams only."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from . import _lib as L

Tensor = torch.Tensor


PROFILE = None    # when set to a list, conv3x3 / wgrad3x3 append (kind, algorithmic_flops, start_event, end_event)


class _Prof:
    def __init__(self, kind, flops):
        self.kind, self.flops = kind, flops

    def __enter__(self):
        if PROFILE is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *exc):
        if PROFILE is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            PROFILE.append((self.kind, self.flops, self.e0, e1))
        return False


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _require_cuda(t: Tensor, name: str) -> None:
    if not t.is_cuda:
        raise RuntimeError(f"CUDA tensor required for {name}: super-resolution_b200 has no CPU path")


# ------------------------------------------------------------------ PNHWC activation buffers
def pnhwc_zeros(batch: int, H: int, W: int, C_: int, device) -> Tensor:
    """bf16 [batch, H+1, W+1, C] with the zero pad row/column (include/srb200.h)."""
    return torch.zeros(batch, H + 1, W + 1, C_, dtype=torch.bfloat16, device=device)


def to_pnhwc(x: Tensor, C_total: Optional[int] = None, coff: int = 0) -> Tensor:
    """NCHW float -> PNHWC bf16 (test / glue helper, not on the hot path)."""
    B, Cc, H, W = x.shape
    buf = pnhwc_zeros(B, H, W, C_total or Cc, x.device)
    buf[:, :H, :W, coff:coff + Cc] = x.permute(0, 2, 3, 1).to(torch.bfloat16)
    return buf


def from_pnhwc(buf: Tensor, coff: int = 0, c: Optional[int] = None) -> Tensor:
    """PNHWC bf16 -> NCHW fp32."""
    B, Hp, Wp, Ct = buf.shape
    c = Ct - coff if c is None else c
    return buf[:, :Hp - 1, :Wp - 1, coff:coff + c].permute(0, 3, 1, 2).float().contiguous()


# ------------------------------------------------------------------ weight packs
def pack_fwd(w: Tensor, shuffle: bool = False, scale: float = 1.0, out: Optional[Tensor] = None) -> Tensor:
    """fp32 OIHW -> bf16 [9][cin/64][cout][64]."""
    _require_cuda(w, "pack_fwd")
    cout, cin = w.shape[0], w.shape[1]
    w = w.detach().contiguous().float()
    if out is None:
        out = torch.empty(9 * cin * cout, dtype=torch.bfloat16, device=w.device)
    L.check(L.lib().srb_pack_conv3x3_fwd(w.data_ptr(), cout, cin, int(shuffle), float(scale), out.data_ptr(),
                                         _stream()), "pack_fwd")
    return out


def pack_dgrad(w: Tensor, ci_off: int, n_out: int, out: Tensor, kb_dst: int, kb_total: int, *,
               unshuffle: bool = False, scale: float = 1.0) -> Tensor:
    """dgrad pack of forward weight ``w`` (fp32 OIHW) for input channels [ci_off, ci_off+n_out), written at
    k-block ``kb_dst`` of ``out`` = bf16 [9][kb_total][n_out][64]."""
    _require_cuda(w, "pack_dgrad")
    cout, cin = w.shape[0], w.shape[1]
    w = w.detach().contiguous().float()
    L.check(L.lib().srb_pack_conv3x3_dgrad(w.data_ptr(), cout, cin, ci_off, n_out, int(unshuffle), float(scale),
                                           out.data_ptr(), kb_dst, kb_total, _stream()), "pack_dgrad")
    return out


# ------------------------------------------------------------------ convolution
def conv3x3_args(x: Tensor, x_coff: int, cin: int, wp: Tensor, bias: Optional[Tensor], y: Tensor, y_coff: int, cout: int,
                 *, alpha: float = 1.0, act: int = L.ACT_NONE, slope: float = 0.01,
                 mask: Optional[Tuple[Tensor, int]] = None, res1: Optional[Tuple[Tensor, int, float]] = None,
                 res2: Optional[Tuple[Tensor, int, float]] = None, shuffle: bool = False, impl: int = L.IMPL_TCGEN05,
                 flags: int = 0, colsum: Optional[Tensor] = None, sign_out: Optional[Tuple[Tensor, int]] = None,
                 mask_bits: Optional[Tuple[Tensor, int]] = None, y32: Optional[Tuple[Tensor, int]] = None,
                 res3: Optional[Tuple[Tensor, int, float]] = None, chan_scale: Optional[Tensor] = None):
    """srb_conv3x3_args for y[..., y_coff:y_coff+cout] = epilogue(conv3x3(x[..., x_coff:x_coff+cin])) and its
    algorithmic FLOPs; x, y PNHWC bf16 buffers."""
    _require_cuda(x, "conv3x3")
    B, Hp, Wp, Cx = x.shape
    a = L.Conv3x3Args()
    a.batch, a.H, a.W = B, Hp - 1, Wp - 1
    a.x, a.x_cstride, a.x_coff, a.cin = x.data_ptr(), Cx, x_coff, cin
    a.w, a.bias, a.cout = wp.data_ptr(), _ptr(bias), cout
    a.y, a.y_cstride, a.y_coff = y.data_ptr(), y.shape[-1], y_coff
    a.store_mode = L.STORE_SHUFFLE if shuffle else L.STORE_PLAIN
    if shuffle:
        assert y.shape[1] == 2 * a.H + 1 and y.shape[2] == 2 * a.W + 1 and y.shape[0] == B
    else:
        assert y.shape[:3] == x.shape[:3]
    a.alpha, a.act, a.slope = alpha, act, slope
    if mask is not None:
        a.mask, a.mask_cstride, a.mask_coff = mask[0].data_ptr(), mask[0].shape[-1], mask[1]
    f32 = 0
    if res1 is not None:     # (tensor, channel offset, scale): bf16 PNHWC, or an fp32 buffer of the same geometry
        a.res1, a.res1_cstride, a.res1_coff, a.res1_scale = res1[0].data_ptr(), res1[0].shape[-1], res1[1], res1[2]
        assert res1[0].dtype in (torch.bfloat16, torch.float32) and res1[0].shape[:3] == x.shape[:3]
        f32 |= 1 if res1[0].dtype == torch.float32 else 0
    if res2 is not None:
        a.res2, a.res2_cstride, a.res2_coff, a.res2_scale = res2[0].data_ptr(), res2[0].shape[-1], res2[1], res2[2]
        assert res2[0].dtype in (torch.bfloat16, torch.float32) and res2[0].shape[:3] == x.shape[:3]
        f32 |= 2 if res2[0].dtype == torch.float32 else 0
    a.res_f32 = f32
    if res3 is not None:
        assert res3[0].dtype == torch.bfloat16 and res3[0].shape[:3] == x.shape[:3]
        a.res3, a.res3_cstride, a.res3_coff, a.res3_scale = res3[0].data_ptr(), res3[0].shape[-1], res3[1], res3[2]
    if y32 is not None:      # (fp32 tensor of y's pixel geometry, channel offset): unrounded copy of the output
        assert y32[0].dtype == torch.float32 and y32[0].shape[:3] == y.shape[:3] and not shuffle
        a.y32, a.y32_cstride, a.y32_coff = y32[0].data_ptr(), y32[0].shape[-1], y32[1]
    if chan_scale is not None:   # fp32 [batch, cout]
        assert chan_scale.dtype == torch.float32 and tuple(chan_scale.shape) == (B, cout) and chan_scale.is_contiguous()
        a.chan_scale = chan_scale.data_ptr()
    a.impl, a.flags = impl, flags
    def sign_layout(t):         # row-major [B, H+1, W+1, words] -> +words; plane-major [words, B*(H+1)*(W+1)] -> -pixels
        assert t.dtype == torch.int32 and t.is_contiguous()
        if t.dim() == 2:
            assert t.shape[1] == y.shape[0] * y.shape[1] * y.shape[2]
            return -t.shape[1]
        assert t.shape[:3] == y.shape[:3]
        return t.shape[-1]
    if sign_out is not None:    # (int32 sign words, first word): (output > 0) bits, 32 channels per word
        a.sign_out, a.sign_cstride, a.sign_coff = sign_out[0].data_ptr(), sign_layout(sign_out[0]), sign_out[1]
    if mask_bits is not None:   # LeakyReLU' from such sign words instead of the bf16 activation
        a.mask_bits, a.maskb_cstride, a.maskb_coff = mask_bits[0].data_ptr(), sign_layout(mask_bits[0]), mask_bits[1]
    if colsum is not None:      # fp32 [cout] accumulator of per-channel output sums (bias gradient), cout == 64
        assert colsum.dtype == torch.float32 and colsum.numel() >= cout and colsum.is_contiguous()
        a.colsum = colsum.data_ptr()
    return a, 2.0 * B * a.H * a.W * 9 * cin * cout


def conv3x3(x: Tensor, x_coff: int, cin: int, wp: Tensor, bias: Optional[Tensor], y: Tensor, y_coff: int, cout: int,
            **kw) -> None:
    """One convolution launch (keyword arguments as conv3x3_args)."""
    a, flops = conv3x3_args(x, x_coff, cin, wp, bias, y, y_coff, cout, **kw)
    with _Prof("conv3x3", flops):
        L.check(L.lib().srb_conv3x3(C.byref(a), _stream()), "conv3x3")


CHAIN_MAX = 16


def chain_flag_ints(batch: int, H: int, W: int, n: int) -> int:
    return int(L.lib().srb_conv3x3_chain_flag_bytes(batch, H, W, n)) // 4


def conv3x3_chain(layers, flags: Tensor) -> None:
    """``layers`` = list of (args, flops) from conv3x3_args, executed in order inside ONE persistent launch
    (csrc/conv_chain.cu).  ``flags``: int32 device workspace of chain_flag_ints(...) entries, zero on entry."""
    n = len(layers)
    arr = (L.Conv3x3Args * n)(*[a for a, _ in layers])
    with _Prof("conv3x3", sum(f for _, f in layers)):
        L.check(L.lib().srb_conv3x3_chain(C.cast(arr, C.c_void_p), n, flags.data_ptr(), flags.numel() * 4, _stream()),
                "conv3x3_chain")


def wgrad_workspace_bytes(batch: int, H: int, W: int, cin: int, cout: int) -> int:
    return int(L.lib().srb_wgrad3x3_workspace_bytes(batch, H, W, cin, cout))


def wgrad3x3(x: Tensor, x_coff: int, cin: int, gy: Tensor, gy_coff: int, cout: int, dw: Tensor,
             dbias: Optional[Tensor], *, alpha: float = 1.0, accumulate: bool = False, unshuffle: bool = False,
             workspace: Optional[Tensor] = None, impl: int = L.IMPL_TCGEN05, flags: int = 0) -> None:
    """dw (fp32 OIHW) = alpha * sum_pixels gy (x) x; both PNHWC bf16 buffers of the same geometry."""
    _require_cuda(x, "wgrad3x3")
    B, Hp, Wp, Cx = x.shape
    assert gy.shape[:3] == x.shape[:3]
    assert dw.dtype == torch.float32 and dw.is_contiguous() and tuple(dw.shape) == (cout, cin, 3, 3)
    a = L.Wgrad3x3Args()
    a.batch, a.H, a.W = B, Hp - 1, Wp - 1
    a.x, a.x_cstride, a.x_coff, a.cin = x.data_ptr(), Cx, x_coff, cin
    a.gy, a.gy_cstride, a.gy_coff, a.cout = gy.data_ptr(), gy.shape[-1], gy_coff, cout
    a.dw, a.dbias, a.alpha = dw.data_ptr(), _ptr(dbias), alpha
    a.accumulate, a.unshuffle = int(accumulate), int(unshuffle)
    if impl == L.IMPL_TCGEN05:
        need = wgrad_workspace_bytes(B, a.H, a.W, cin, cout)
        if workspace is None or workspace.numel() * workspace.element_size() < need:
            workspace = torch.empty(need, dtype=torch.uint8, device=x.device)
        a.workspace, a.workspace_bytes = workspace.data_ptr(), workspace.numel() * workspace.element_size()
    a.impl, a.flags = impl, flags
    with _Prof("wgrad3x3", 2.0 * B * a.H * a.W * 9 * cin * cout):
        L.check(L.lib().srb_wgrad3x3(C.byref(a), _stream()), "wgrad3x3")


def wgrad_group_workspace_bytes(batch: int, H: int, W: int, cins, couts) -> int:
    n = len(cins)
    ci = (C.c_int * n)(*cins)
    co = (C.c_int * n)(*couts)
    return int(L.lib().srb_wgrad3x3_group_workspace_bytes(batch, H, W, n, C.cast(ci, C.c_void_p), C.cast(co, C.c_void_p)))


def wgrad3x3_group(x: Tensor, gy: Tensor, problems, workspace: Tensor, *, accumulate: bool = False,
                   gy_colsum: Optional[Tensor] = None) -> None:
    """Several weight gradients over the same buffers in one launch.  ``problems`` = list of
    (x_coff, cin, gy_coff, cout, dw, dbias_or_None, alpha).  ``gy_colsum``: fp32 [channels of gy] column sums of gy
    accumulated by the convolutions that wrote it (conv3x3_args(colsum=...)); consumed and reset to zero."""
    _require_cuda(x, "wgrad3x3_group")
    B, Hp, Wp, Cx = x.shape
    assert gy.shape[:3] == x.shape[:3]
    n = len(problems)
    arr = (L.Wgrad3x3Args * n)()
    flops = 0.0
    for a, (x_coff, cin, gy_coff, cout, dw, db, alpha) in zip(arr, problems):
        assert dw.dtype == torch.float32 and dw.is_contiguous() and tuple(dw.shape) == (cout, cin, 3, 3)
        a.batch, a.H, a.W = B, Hp - 1, Wp - 1
        a.x, a.x_cstride, a.x_coff, a.cin = x.data_ptr(), Cx, x_coff, cin
        a.gy, a.gy_cstride, a.gy_coff, a.cout = gy.data_ptr(), gy.shape[-1], gy_coff, cout
        a.dw, a.dbias, a.alpha = dw.data_ptr(), _ptr(db), alpha
        a.accumulate, a.unshuffle = int(accumulate), 0
        a.workspace, a.workspace_bytes = workspace.data_ptr(), workspace.numel() * workspace.element_size()
        if gy_colsum is not None:
            assert gy_colsum.dtype == torch.float32 and gy_colsum.numel() == gy.shape[-1]
            a.gy_colsum = gy_colsum.data_ptr()
        flops += 2.0 * B * a.H * a.W * 9 * cin * cout
    with _Prof("wgrad3x3", flops):
        L.check(L.lib().srb_wgrad3x3_group(C.cast(arr, C.c_void_p), n, _stream()), "wgrad3x3_group")


def read_fault() -> int:
    v = C.c_uint(0)
    L.check(L.lib().srb_read_fault(C.byref(v)), "read_fault")
    return int(v.value)
