"""Hand-scheduled execution engine for the patch discriminator (reference models.py:140-174) on the bf16
tensor-core kernels of csrc/disc_mma.cu.  It evaluates exactly the quantities the training step differentiates
(esrgan.py:493-508 generator side; 569-616 discriminator side incl. the gradient penalty's second derivative)
as a fixed sequence of kernel launches - no autograd graph, no per-layer torch ops.

Notation for one conv stack  z_l = conv_l(a_{l-1}) + b_l,  a_l = LeakyReLU(z_l)  (l = 1..L, a_L = z_L = pred):

  forward(imgs)            a_1..a_L                       stores bf16 NHWC activations
  backward(acts, g_pred)   gz_L = g_pred;  gz_{l-1} = dgrad_l(gz_l) * LeakyReLU'(a_{l-1});  dW_l += wgrad(gz_l, a_{l-1})
  input_grad(acts, g_pred) the same chain down to the image, no weight gradients (generator step / penalty)
  penalty second order     GP = coef * mean_b (||h_0||_2 - 1)^2 with h the chain started from ones.  d GP / d W_l has
                           only the "direct" term because LeakyReLU' is piecewise constant:
                             q_0 = d GP / d h_0;  dW_l += wgrad(hz_l, q_{l-1});  q_l = conv_l(q_{l-1}) * LeakyReLU'(a_l)

The module API (models.Markovian_Discriminator.forward under autograd) keeps using the general fp32 Functions of
srb200/discriminator.py; this engine is what srb200.train_step drives.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import torch
import torch.nn as nn

from . import _lib as L
from . import generator as _gen

Tensor = torch.Tensor
_OK = (16, 32, 64)


def _st() -> int:
    return torch.cuda.current_stream().cuda_stream


def supported(seq: nn.Sequential, input_shape) -> bool:
    """True when ``seq`` is a Conv2d(3x3, pad 1, stride 1|2)/LeakyReLU chain with channel widths the MMA kernels
    cover: 1 -> {16,32,64} ... -> 1, even maps in front of every stride-2 layer."""
    from .discriminator import _layers
    layers = _layers(seq)
    if not layers or len(layers) < 2:
        return False
    c, h, w = input_shape
    if c != 1:
        return False
    for i, (conv, slope) in enumerate(layers):
        if conv.kernel_size != (3, 3) or conv.padding != (1, 1) or conv.stride[0] != conv.stride[1] \
                or conv.stride[0] not in (1, 2) or conv.bias is None or conv.groups != 1 or conv.dilation != (1, 1):
            return False
        last = i == len(layers) - 1
        if i == 0 and (conv.in_channels != 1 or conv.stride[0] != 1 or conv.out_channels not in _OK):
            return False
        if i > 0 and conv.in_channels not in _OK:
            return False
        if not last and (conv.out_channels not in _OK or slope is None):
            return False
        if last and (conv.out_channels != 1 or slope is not None):
            return False
        if conv.stride[0] == 2 and (h % 2 or w % 2):
            return False
        h, w = (h - 1) // conv.stride[0] + 1, (w - 1) // conv.stride[0] + 1
    return True


class _Acts:
    """bf16 NHWC activations a_1..a_{L-1} and the fp32 prediction for one batch size."""

    def __init__(self, eng: "DiscEngine", B: int, dev):
        self.B = B
        self.a = [torch.empty(B, h, w, c, dtype=torch.bfloat16, device=dev) for (c, h, w) in eng.geo[1:-1]]
        c, h, w = eng.geo[-1]
        self.pred = torch.empty(B, 1, h, w, device=dev)
        self.imgs: List[Tuple[Tensor, int]] = []          # (fp32 image tensor, batch offset)


class DiscEngine:
    def __init__(self, seq: nn.Sequential, input_shape):
        from .discriminator import _layers
        if not supported(seq, input_shape):
            raise RuntimeError("CUDA srb200: discriminator configuration not covered by the tensor-core engine")
        self.layers = _layers(seq)
        self.L = len(self.layers)
        self.convs = [c for c, _ in self.layers]
        self.slopes = [s for _, s in self.layers]
        c, h, w = input_shape
        self.geo = [(c, h, w)]                              # geo[l] = (channels, H, W) of a_l; geo[0] = image
        for conv in self.convs:
            s = conv.stride[0]
            h, w = (h - 1) // s + 1, (w - 1) // s + 1
            self.geo.append((conv.out_channels, h, w))
        self._pack_epoch = None
        self._packs = None
        self._acts = {}
        self._chain = {}

    # ------------------------------------------------------------------ weight packs (bf16, once per optimizer step)
    def _ensure_packs(self) -> None:
        key = (_gen.PARAM_EPOCH[0], sum(c.weight._version for c in self.convs), self.convs[0].weight.data_ptr())
        if key == self._pack_epoch:
            return
        dev = self.convs[0].weight.device
        ws = [c.weight for c in self.convs]
        # all parameters of one FusedAdam live in one flat buffer: address the layers by offset from the lowest one
        base = min(w.data_ptr() for w in ws)
        items = []          # (src_off, cout, cin, kind, rows, kpad)
        for l in range(1, self.L):                           # layer index l (0-based), layer 0 is the fp32 image end
            w = ws[l]
            co, ci = w.shape[0], w.shape[1]
            items.append((l, 0, (w.data_ptr() - base) // 4, co, ci))
            items.append((l, 1 if self.convs[l].stride[0] == 1 else 2, (w.data_ptr() - base) // 4, co, ci))
        pad8 = lambda v: (v + 7) // 8 * 8
        pad16 = lambda v: (v + 15) // 16 * 16
        offs, total = [], 0
        for (_, kind, _, co, ci) in items:
            rows, kp = (pad8(co), pad16(ci)) if kind == 0 else (pad8(ci), pad16(co))
            offs.append(total)
            total += 9 * rows * kp
        if self._packs is None or self._packs.numel() != total or self._packs.device != dev:
            self._packs = torch.empty(total, dtype=torch.bfloat16, device=dev)
        # the flat view must cover every layer: verify the parameters really are in one allocation (FusedAdam)
        span_end = max(w.data_ptr() + w.numel() * 4 for w in ws)
        stor = ws[0].untyped_storage()
        src_offs = [it[2] for it in items]
        if stor.data_ptr() <= base and span_end <= stor.data_ptr() + stor.nbytes():
            base_ptr = base
        else:                                               # not flattened: pack from a private concatenation
            flat = torch.cat([w.detach().reshape(-1).float() for w in ws])
            o, starts = 0, {}
            for l, w in enumerate(ws):
                starts[l] = o
                o += w.numel()
            src_offs = [starts[it[0]] for it in items]
            base_ptr = flat.data_ptr()
            self._flat_keepalive = flat
        for i0 in range(0, len(items), 24):
            sub = items[i0:i0 + 24]
            m = len(sub)
            src = (C.c_longlong * m)(*src_offs[i0:i0 + m])
            dst = (C.c_longlong * m)(*offs[i0:i0 + m])
            co_ = (C.c_int * m)(*[it[3] for it in sub])
            ci_ = (C.c_int * m)(*[it[4] for it in sub])
            kd = (C.c_int * m)(*[it[1] for it in sub])
            L.check(L.lib().srb_dpack(base_ptr, self._packs.data_ptr(), m, C.cast(src, C.c_void_p),
                                      C.cast(dst, C.c_void_p), C.cast(co_, C.c_void_p), C.cast(ci_, C.c_void_p),
                                      C.cast(kd, C.c_void_p), _st()), "dpack")
        self.wf, self.wd = {}, {}
        for (l, kind, _, co, ci), o in zip(items, offs):
            rows, kp = (pad8(co), pad16(ci)) if kind == 0 else (pad8(ci), pad16(co))
            view = self._packs[o:o + 9 * rows * kp]
            if kind == 0:
                self.wf[l] = view
            else:
                self.wd[l] = view
        self._pack_epoch = key

    # ------------------------------------------------------------------ buffers
    def acts(self, B: int, dev, tag: str = "") -> _Acts:
        key = (B, str(dev), tag)
        a = self._acts.get(key)
        if a is None:
            a = _Acts(self, B, dev)
            self._acts[key] = a
        return a

    def _chain_bufs(self, B: int, dev, tag: str):
        """gradient chain buffers gz_1..gz_{L-1} (bf16, geometry of a_l)."""
        key = (B, str(dev), tag)
        c = self._chain.get(key)
        if c is None:
            c = [torch.empty(B, h, w, ch, dtype=torch.bfloat16, device=dev) for (ch, h, w) in self.geo[1:-1]]
            self._chain[key] = c
        return c

    # ------------------------------------------------------------------ kernels
    def _conv(self, l: int, src_bf16, src_f32, out_bf16, out_f32, B, *, bias: bool, act: int, ref=None,
              slope: float = 0.0):
        """forward conv of layer l (0-based, l >= 1)."""
        conv = self.convs[l]
        ci, hs, ws = self.geo[l]
        co, hd, wd = self.geo[l + 1]
        L.check(L.lib().srb_dmma_conv(src_bf16.data_ptr() if src_bf16 is not None else None,
                                      src_f32.data_ptr() if src_f32 is not None else None, ci,
                                      self.wf[l].data_ptr(), conv.bias.data_ptr() if bias else None, co,
                                      out_bf16.data_ptr() if out_bf16 is not None else None,
                                      out_f32.data_ptr() if out_f32 is not None else None,
                                      ref.data_ptr() if ref is not None else None, B, hs, ws, hd, wd, conv.stride[0], 0,
                                      act, float(slope), _st()), "dmma_conv")

    def _dgrad(self, l: int, gz_bf16, gz_f32, out_bf16, B, *, ref, slope):
        """input gradient of layer l (0-based, l >= 1): out = dgrad_l(gz) * LeakyReLU'(ref)."""
        conv = self.convs[l]
        ci, hi, wi = self.geo[l]
        co, ho, wo = self.geo[l + 1]
        s = conv.stride[0]
        L.check(L.lib().srb_dmma_conv(gz_bf16.data_ptr() if gz_bf16 is not None else None,
                                      gz_f32.data_ptr() if gz_f32 is not None else None, co, self.wd[l].data_ptr(),
                                      None, ci, out_bf16.data_ptr(), None, ref.data_ptr(), B, ho, wo, hi, wi, 1,
                                      1 if s == 2 else 0, 2, float(slope), _st()), "dmma_dgrad")

    def _wgrad(self, l: int, gz_bf16, gz_f32, a_bf16, B, dw: Tensor, db: Optional[Tensor]):
        conv = self.convs[l]
        ci, hi, wi = self.geo[l]
        co, ho, wo = self.geo[l + 1]
        L.check(L.lib().srb_dmma_wgrad(gz_bf16.data_ptr() if gz_bf16 is not None else None,
                                       gz_f32.data_ptr() if gz_f32 is not None else None, co, a_bf16.data_ptr(), ci,
                                       dw.data_ptr(), db.data_ptr() if db is not None else None, B, hi, wi, ho, wo,
                                       conv.stride[0], _st()), "dmma_wgrad")

    # ------------------------------------------------------------------ forward
    def forward(self, imgs: Sequence[Tensor], tag: str = "") -> _Acts:
        """Run the stack on the concatenation of ``imgs`` (fp32 [b_i,1,H,W] each); activations are kept."""
        self._ensure_packs()
        imgs = [t.detach().float().contiguous() for t in imgs]
        for t in imgs:
            if not t.is_cuda:
                raise RuntimeError("CUDA tensor required: super-resolution_b200 has no CPU path for the discriminator")
        B = sum(t.shape[0] for t in imgs)
        dev = imgs[0].device
        A = self.acts(B, dev, tag)
        A.imgs = []
        c1, h, w = self.geo[1]
        conv0 = self.convs[0]
        off = 0
        for t in imgs:
            b = t.shape[0]
            L.check(L.lib().srb_dl1_fwd(t.data_ptr(), conv0.weight.data_ptr(), conv0.bias.data_ptr(),
                                        A.a[0][off:].data_ptr(), None, c1, b, h, w, 1, float(self.slopes[0]), _st()),
                    "dl1_fwd")
            A.imgs.append((t, off))
            off += b
        for l in range(1, self.L):
            last = l == self.L - 1
            self._conv(l, A.a[l - 1], None, None if last else A.a[l], A.pred if last else None, B, bias=True,
                       act=0 if last else 1, slope=0.0 if last else self.slopes[l])
        return A

    # ------------------------------------------------------------------ first-order chains
    def _chain_down(self, A: _Acts, g_pred: Tensor, b0: int, B: int, tag: str, want_wgrad: bool, grads):
        """gz_l for l = L-1 .. 1 over the images [b0, b0+B) of ``A``; optionally accumulates dW_l, db_l (l >= 1, 0-based
        layer indices 1..L-1) into ``grads`` = list of (dw, db).  Returns the chain buffers (gz of a_1 .. a_{L-1})."""
        dev = g_pred.device
        ch = self._chain_bufs(B, dev, tag)
        g_pred = g_pred.detach().float().contiguous()
        for l in range(self.L - 1, 0, -1):
            last = l == self.L - 1
            a_in = A.a[l - 1][b0:b0 + B]
            gz_bf16 = None if last else ch[l]
            gz_f32 = g_pred if last else None
            if want_wgrad:
                dw, db = grads[l]
                if last:        # single fp32 channel: its bias gradient is a plain sum, kept in fp32 so that the
                    self._wgrad(l, gz_bf16, gz_f32, a_in, B, dw, None)      # relativistic loss's exact cancellation
                    db.add_(g_pred.sum().reshape(db.shape))                 # (sum of d loss / d pred == 0) survives
                else:
                    self._wgrad(l, gz_bf16, gz_f32, a_in, B, dw, db)
            self._dgrad(l, gz_bf16, gz_f32, ch[l - 1], B, ref=a_in, slope=self.slopes[l - 1])
        return ch

    def input_grad(self, A: _Acts, g_pred: Tensor, b0: int, B: int, tag: str = "ig") -> Tuple[Tensor, list]:
        """d <g_pred, D(x)> / d x for the images [b0, b0+B) of the forward ``A``: fp32 [B,1,H,W] (+ the chain)."""
        ch = self._chain_down(A, g_pred, b0, B, tag, False, None)
        c1, h, w = self.geo[1]
        gx = torch.empty(B, 1, h, w, device=g_pred.device)
        L.check(L.lib().srb_dl1_dgrad(ch[0].data_ptr(), self.convs[0].weight.data_ptr(), gx.data_ptr(), c1, B, h, w,
                                      _st()), "dl1_dgrad")
        return gx, ch

    def backward(self, A: _Acts, g_pred: Tensor, grads, tag: str = "bw") -> None:
        """Accumulate d <g_pred, D(imgs)> / d parameters into ``grads`` (list over layers of (dw, db) fp32 views)."""
        B = A.B
        ch = self._chain_down(A, g_pred, 0, B, tag, True, grads)
        c1, h, w = self.geo[1]
        dw, db = grads[0]
        for t, off in A.imgs:
            b = t.shape[0]
            L.check(L.lib().srb_dl1_wgrad(ch[0][off:].data_ptr(), t.data_ptr(), dw.data_ptr(), db.data_ptr(), c1, b, h, w,
                                          _st()), "dl1_wgrad")

    # ------------------------------------------------------------------ gradient penalty (esrgan.py:596-606)
    def penalty_backward(self, A: _Acts, hz: list, q0: Tensor, grads) -> None:
        """Second-order term: given the penalty's chain ``hz`` (from input_grad with g_pred = ones) and
        q0 = d GP / d (input gradient) (fp32 [B,1,H,W]), accumulate d GP / d W_l into ``grads``."""
        B = A.B
        dev = q0.device
        q0 = q0.detach().float().contiguous()
        c1, h, w = self.geo[1]
        dw0, _ = grads[0]
        # layer 0: dW_0 += wgrad(hz_1, q_0)   (bias has no part in the input gradient)
        L.check(L.lib().srb_dl1_wgrad(hz[0].data_ptr(), q0.data_ptr(), dw0.data_ptr(), None, c1, B, h, w, _st()),
                "dl1_wgrad(q)")
        q = self._chain_bufs(B, dev, "q")
        conv0 = self.convs[0]
        # q_1 = conv_0(q_0) * LeakyReLU'(a_1)
        L.check(L.lib().srb_dl1_fwd(q0.data_ptr(), conv0.weight.data_ptr(), None, q[0].data_ptr(), A.a[0].data_ptr(), c1,
                                    B, h, w, 2, float(self.slopes[0]), _st()), "dl1_fwd(q)")
        ones = None
        for l in range(1, self.L):
            last = l == self.L - 1
            dw, _ = grads[l]
            if last:
                co, ho, wo = self.geo[l + 1]
                ones = torch.ones(B, ho, wo, device=dev)
                self._wgrad(l, None, ones, q[l - 1], B, dw, None)
            else:
                self._wgrad(l, hz[l], None, q[l - 1], B, dw, None)
                self._conv(l, q[l - 1], None, q[l], None, B, bias=False, act=2, ref=A.a[l], slope=self.slopes[l])
