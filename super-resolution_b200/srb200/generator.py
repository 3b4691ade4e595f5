"""Execution engine for GeneratorRRDB on the C ABI kernels (forward + hand-scheduled backward).

Mirrors models.py:9-135 of the reference.  The module objects in ``models.py`` of this package only hold
parameters (same names/shapes as the reference state_dict); this engine turns a forward into a fixed sequence of
kernel launches over preallocated PNHWC bf16 buffers:

* each DenseResidualBlock owns a growth buffer [B, h+1, w+1, 5F]; conv k reads channels [0, kF) and writes slot k
  in place (no torch.cat, models.py:38); conv 5's epilogue applies ``out*res_scale + x`` (models.py:41) and, for
  the third block of an RRDB, the outer ``*res_scale + x`` too (models.py:53), storing into slot 0 of the next
  block's buffer;
* backward mirrors it on a gradient growth buffer [g_out, gz5, gz4, ..., gz1]: the gradient of slot j is ONE
  convolution over the already-final gradient slots with a concatenated transposed/flipped weight pack, so the
  dense block's input gradients are never read-modify-written (hard part 2).

Constructor variants (models.py:58-106), all on the same kernels:
* ``num_final_layer_res`` (models.py:93-94,128-130): a second RRDB trunk at the upsampled geometry; the last
  upsampling stage stores straight into its first growth buffer, ``out3 + out`` is the third residual of its last
  convolution (srb_conv3x3_args.res3) and of the matching input-gradient convolution;
* ``use_transposed_conv`` / ``fully_tconv_upsample`` (models.py:70-83): ConvTranspose2d(F, F, k=2, s=2) + LeakyReLU is a
  1x1 convolution to 4F channels followed by PixelShuffle(2) with W3[4co+2i+j][ci] = Wt[ci][co][i][j].  It runs as the
  same convolution + shuffle-store kernel as the default stage through a "virtual" 3x3 weight whose other eight taps are
  zero (rebuilt from the module's [F,F,2,2] parameter whenever it changes); its weight gradient is the centre tap of
  the virtual weight's gradient, folded back into the parameter's layout;
* ``drop_rate`` (models.py:29-31,39-40): Dropout2d on out5 in training = a per (image, channel) factor on the fifth
  convolution's accumulator (srb_conv3x3_args.chan_scale); backward masks the block's output gradient for the weight /
  input gradients of conv 5 and keeps an unmasked copy for the identity paths (srb_chan_mask_split).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from . import _lib as L
from . import ops

Tensor = torch.Tensor
G_SLOPE = 0.01      # nn.LeakyReLU() default (models.py:21,88,98)
DRB_SCALE = 0.2     # DenseResidualBlock default res_scale; never overridden (models.py:49)

PARAM_EPOCH = [0]   # bumped by srb200.optim.FusedAdam.step(); engines re-pack weights when it moves


def _st() -> int:
    return torch.cuda.current_stream().cuda_stream


class _Bufs:
    """Activation / gradient buffers for one (batch, h, w, training) shape."""
    pass


class _FlagArena:
    """Per-tile completion counters of the chained convolution launches (csrc/conv_chain.cu): every chain launch of a
    pass takes its own zeroed slice; ONE clear per pass, at its start, so no memset sits between two launches."""

    def __init__(self, n_ints: int, dev):
        self.t = torch.zeros(max(n_ints, 1), dtype=torch.int32, device=dev)
        self.off = 0

    def reset(self) -> None:
        self.t.zero_()
        self.off = 0

    def take(self, n: int) -> Tensor:
        if self.off + n > self.t.numel():
            raise RuntimeError("CUDA srb200: chain flag arena exhausted")
        v = self.t[self.off:self.off + n]
        self.off += n
        return v


def _chain_default() -> int:
    """Convolutions per chained launch in the RRDB trunk (multiple of 5 = whole dense blocks; 0 = one launch per
    convolution).  Default 5: one launch per dense block, forward and input-gradient pass (measured at batch 512:
    GAN step 34.4 ms with 0, 33.9 ms with 5, 33.95 ms with 15).  SRB_CONV_CHAIN overrides."""
    v = int(os.environ.get("SRB_CONV_CHAIN", "5"))
    return max(0, min(v, ops.CHAIN_MAX))


class GeneratorEngine:
    def __init__(self, params: Dict[str, torch.nn.Parameter], *, channels: int, filters: int, num_res_blocks: int,
                 num_upsample: int, res_scale: float, impl: int = L.IMPL_TCGEN05,
                 up_stages: Optional[Sequence[Tuple[str, str]]] = None, num_final: int = 0, drop_rate: float = 0.0):
        """``up_stages``: per upsampling stage (kind, parameter prefix) with kind 'ps' (conv + LeakyReLU + PixelShuffle)
        or 'tconv' (ConvTranspose2d + LeakyReLU); default: ``num_upsample`` 'ps' stages named upsampling.{3u}."""
        if filters % 64:
            raise RuntimeError("CUDA srb200: GeneratorRRDB engine needs filters to be a multiple of 64")
        if channels > 4:
            raise RuntimeError("CUDA srb200: at most 4 image channels")
        self.p = params
        self.ch, self.F, self.nb, self.nup, self.rs = channels, filters, num_res_blocks, num_upsample, float(res_scale)
        self.ups = list(up_stages) if up_stages is not None else [("ps", f"upsampling.{3 * u}") for u in range(num_upsample)]
        if len(self.ups) != num_upsample:
            raise RuntimeError("CUDA srb200: upsampling stage list does not match num_upsample")
        self.nfinal, self.drop_rate = int(num_final), float(drop_rate)
        if self.nfinal and not self.nup:
            raise RuntimeError("CUDA srb200: num_final_layer_res needs at least one upsampling stage")
        self.impl = impl
        self.chain = _chain_default() if impl == L.IMPL_TCGEN05 else 0
        self.conv_flags = 0          # srb_conv3x3_args.flags of every tcgen05 convolution (kernel selection bits, tests)
        self._bufs: Dict[Tuple, _Bufs] = {}
        self._pack_key = None
        self._packs_bwd_ready = False
        self._saved = None
        self._fwd_gen = 0            # forward(save_for_backward=True) counter: backward must match the LAST such forward
        self._bias_perm = {}
        self._virt: Dict[str, Dict[str, Tensor]] = {}     # tconv stages: virtual 3x3 weight / bias and their gradients
        self._main_marker = None
        # The residual stream of the RRDB trunk (models.py:41,53) is rounded to bf16 once per dense block when it only
        # lives in the growth buffers: 3 nb roundings that accumulate like a random walk - rel-L2 1.35e-2 / energy 3.0e-3
        # at nb = 10 but 2.25e-2 / 5.3e-3 at nb = 23, past SURVEY 8c's 2e-2 / 5e-3.  Deep trunks therefore carry the
        # stream ACROSS RRDBs in fp32 (xs ring: the outer skip `x + res_scale * (...)` reads and writes fp32; the two
        # inner dense-block adds stay bf16, their rounding enters scaled by res_scale): 1.40e-2 / 2.9e-3 at nb = 23 for one
        # fp32 read + write per RRDB.  Carrying every dense-block add in fp32 gives 1.33e-2 and costs 24 % of the forward
        # chain (profiles/r02_s3_exp_engine_fp32_stream.txt), so that is not done.  SRB_FP32_STREAM=0/1 overrides the
        # depth rule (on from 16 RRDBs).
        env = os.environ.get("SRB_FP32_STREAM")
        self.fp32_stream = (num_res_blocks >= 16) if env not in ("0", "1") else env == "1"
        self.grad_views = None       # optional name -> fp32 view: where backward writes parameter gradients
        self.ready_hook = None       # optional callable(param_name): gradients of this and all later parameters are final
        self.names3x3 = self._conv_names()

    # ------------------------------------------------------------------ parameter bookkeeping
    @staticmethod
    def _trunk_names(prefix: str, nb: int) -> List[str]:
        return [f"{prefix}.{i}.dense_blocks.{j}.b{k}.0" for i in range(nb) for j in range(3) for k in range(1, 6)]

    def _conv_names(self) -> List[str]:
        n = self._trunk_names("res_blocks", self.nb)
        n.append("conv2")
        n += [name for _, name in self.ups]
        n += self._trunk_names("res_blocks_final", self.nfinal)
        n.append("conv3.0")
        return n

    def _w(self, name: str) -> Tensor:
        v = self._virt.get(name)
        return v["w3"] if v is not None else self.p[name + ".weight"]

    def _b(self, name: str) -> Tensor:
        v = self._virt.get(name)
        return v["b3"] if v is not None else self.p[name + ".bias"]

    def _params_key(self):
        dev = self.p["conv1.weight"].device
        return (PARAM_EPOCH[0], str(dev), sum(p._version for p in self.p.values()),
                self.p["conv2.weight"].data_ptr())

    def _refresh_virtual(self) -> None:
        """tconv stages: W3[4co+2i+j][ci][1][1] = Wt[ci][co][i][j], other taps zero; b3[4co+2i+j] = bt[co]."""
        F_ = self.F
        for kind, name in self.ups:
            if kind != "tconv":
                continue
            wt, bt = self.p[name + ".weight"], self.p[name + ".bias"]
            v = self._virt.get(name)
            if v is None or v["w3"].device != wt.device:
                v = dict(w3=torch.zeros(4 * F_, F_, 3, 3, device=wt.device), b3=torch.zeros(4 * F_, device=wt.device),
                         gw3=torch.zeros(4 * F_, F_, 3, 3, device=wt.device), gb3=torch.zeros(4 * F_, device=wt.device))
                self._virt[name] = v
            with torch.no_grad():
                v["w3"][:, :, 1, 1].copy_(wt.detach().permute(1, 2, 3, 0).reshape(4 * F_, F_))
                v["b3"].copy_(bt.detach().repeat_interleave(4))

    def _ensure_packs(self, need_bwd: bool) -> None:
        key = self._params_key()
        if key != self._pack_key:
            self._refresh_virtual()
            self._build_fwd_packs()
            self._pack_key = key
            self._packs_bwd_ready = False
        if need_bwd and not self._packs_bwd_ready:
            self._build_bwd_packs()
            self._packs_bwd_ready = True

    # Packs are rebuilt after every optimizer step, so they go through ONE table-driven launch per direction
    # (srb_pack_batch): the table (source / destination pointers and geometry of every pack item) is built once and
    # reused for as long as the parameter storage and the pack buffers stay where they are.
    def _upload_table(self, items, dev):
        n = len(items)
        arr = (L.PackItem * n)(*items)
        raw = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).to(dev)
        prefix, total = [], 0
        for it in items:
            prefix.append(total)
            pairs = it.cout * (it.cin if it.kind == 0 else it.n_out)      # one thread per (out, in) channel pair
            total += (pairs + 255) // 256
        pre = torch.tensor(prefix, dtype=torch.int32, device=dev)
        return dict(raw=raw, prefix=pre, n=n, blocks=total)

    def _run_table(self, tab) -> None:
        L.check(L.lib().srb_pack_batch(tab["raw"].data_ptr(), tab["prefix"].data_ptr(), tab["n"], tab["blocks"], _st()),
                "pack_batch")

    def _ptr_key(self):
        return (str(self.p["conv1.weight"].device),) + tuple(self._w(n).data_ptr() for n in self.names3x3)

    def _build_fwd_packs(self) -> None:
        dev = self.p["conv1.weight"].device
        key = self._ptr_key()
        if getattr(self, "_fwd_tab_key", None) != key:
            sizes, total = {}, 0
            for n in self.names3x3:
                w = self._w(n)
                sizes[n] = (total, w.numel())
                total += w.numel()
            self._fwd_flat = torch.empty(total, dtype=torch.bfloat16, device=dev)
            self.wp = {}
            items = []
            for n in self.names3x3:
                off, cnt = sizes[n]
                w = self._w(n)
                if w.dtype != torch.float32 or not w.is_contiguous():
                    raise RuntimeError("CUDA srb200: generator weights must be contiguous fp32")
                self.wp[n] = self._fwd_flat[off:off + cnt]
                items.append(L.PackItem(w.data_ptr(), self.wp[n].data_ptr(), w.shape[0], w.shape[1], 0, 0,
                                        int(n.startswith("upsampling.")), 0, 0, 0, 1.0, 0))
            self._fwd_tab = self._upload_table(items, dev)
            self._fwd_tab_key = key
        self._run_table(self._fwd_tab)
        self.bias_packed = {}
        for n in self.names3x3:
            b = self._b(n).detach()
            if n.startswith("upsampling."):
                cout = b.numel()
                idx = self._bias_perm.get((n, str(dev)))
                if idx is None:                    # built once: the re-pack after every optimizer step is one gather
                    i = torch.arange(cout, device=dev)
                    idx = self._bias_perm[(n, str(dev))] = 4 * (i % (cout // 4)) + i // (cout // 4)
                b = b[idx]
            self.bias_packed[n] = b.float().contiguous()

    def _build_bwd_packs(self) -> None:
        """dgrad packs.  Dense block: for output slot jj (0..4) one pack with K-blocks ordered like the gradient
        growth buffer [g_out(conv5), gz4, ..., gz_{jj+1}]; conv5's block is pre-scaled by the residual scale."""
        dev = self.p["conv1.weight"].device
        F_, kbF = self.F, self.F // 64
        key = self._ptr_key()
        if getattr(self, "_bwd_tab_key", None) != key:
            self.wd = {}
            items = []

            def item(name, ci_off, n_out, pack, kb_dst, kb_total, unshuffle=False, scale=1.0):
                w = self._w(name)
                items.append(L.PackItem(w.data_ptr(), pack.data_ptr(), w.shape[0], w.shape[1], ci_off, n_out,
                                        int(unshuffle), 1, kb_dst, kb_total, float(scale), 0))

            for prefix, nb in (("res_blocks", self.nb), ("res_blocks_final", self.nfinal)):
                for i in range(nb):
                    for j in range(3):
                        s5 = DRB_SCALE * (self.rs if j == 2 else 1.0)
                        base = f"{prefix}.{i}.dense_blocks.{j}."
                        for jj in range(5):
                            nsrc = 5 - jj
                            pack = torch.empty(9 * nsrc * F_ * F_, dtype=torch.bfloat16, device=dev)
                            for kk in range(5, jj, -1):           # forward conv kk feeds slot jj
                                item(base + f"b{kk}.0", jj * F_, F_, pack, (5 - kk) * kbF, nsrc * kbF,
                                     scale=s5 if kk == 5 else 1.0)
                            self.wd[base + f"slot{jj}"] = pack
            for n in ["conv2", "conv3.0"]:
                pack = torch.empty(9 * F_ * F_, dtype=torch.bfloat16, device=dev)
                item(n, 0, F_, pack, 0, kbF)
                self.wd[n] = pack
            for _, n in self.ups:
                pack = torch.empty(9 * 4 * F_ * F_, dtype=torch.bfloat16, device=dev)
                item(n, 0, F_, pack, 0, 4 * kbF, unshuffle=True)
                self.wd[n] = pack
            self._bwd_tab = self._upload_table(items, dev)
            self._bwd_tab_key = key
        self._run_table(self._bwd_tab)

    # ------------------------------------------------------------------ buffers
    def _trunk_bufs(self, T: _Bufs, nb: int, B: int, h: int, w: int, training: bool, dev) -> None:
        """Buffers of one RRDB trunk (the LR trunk or the num_final_layer_res trunk) at geometry h x w."""
        F_ = self.F
        # training keeps every dense block's buffer (the saved activations); inference cycles through four.  keep0: the
        # final trunk's input (growth[0] slot 0) is also the `+ out` residual of its last convolution, so buffer 0 stays
        # out of the ring there
        T.keep0 = (not training) and T is not self._main_marker
        n_growth = 3 * nb if training else (5 if T.keep0 else 4)
        T.growth = [ops.pnhwc_zeros(B, h, w, 5 * F_, dev) for _ in range(max(n_growth, 1))]
        T.xs = [torch.zeros(B, h + 1, w + 1, F_, dtype=torch.float32, device=dev) for _ in range(2)] \
            if self.fp32_stream else None
        n_flag = 15 * nb * ops.chain_flag_ints(B, h, w, 1)
        T.flags_f = _FlagArena(n_flag, dev)
        T.flags_b = _FlagArena(n_flag if training else 0, dev)
        T.sign = T.G = T.csum = T.Gu = None
        if training:
            T.G = [ops.pnhwc_zeros(B, h, w, 5 * F_, dev) for _ in range(3)]
            T.csum = torch.zeros(3, 5 * F_, device=dev)      # per-channel sums of the G buffers (bias gradients)
            # LeakyReLU sign bits of slots 1..4 of every dense block (F = 64: two 32-bit words per slot and pixel),
            # written by the forward convolutions and read by the input-gradient pass instead of the bf16 activations
            # plane-major [word][pixel]: a warp's 32 rows of one word are one 128-byte access (SRB_SIGN_ROWMAJOR=1: [pixel][word])
            shape = (B, h + 1, w + 1, 8) if os.environ.get("SRB_SIGN_ROWMAJOR") == "1" else (8, B * (h + 1) * (w + 1))
            T.sign = [torch.zeros(*shape, dtype=torch.int32, device=dev) for _ in range(3 * nb)] if F_ == 64 else None
            if self.drop_rate > 0:                           # unmasked copies of the dense blocks' output gradients
                T.Gu = [ops.pnhwc_zeros(B, h, w, F_, dev) for _ in range(3)]

    def _get_bufs(self, B: int, h: int, w: int, training: bool, dev) -> _Bufs:
        key = (B, h, w, training, str(dev), self.fp32_stream)
        bf = self._bufs.pop(key, None)
        if bf is not None:
            self._bufs[key] = bf            # re-insert: dict order = recency
            return bf
        if len(self._bufs) >= 6:
            # evict the least recently used shape only (a hyper-parameter sweep or a partial last batch must not drop
            # the steady-state buffers and force a re-allocation of GBs)
            del self._bufs[next(iter(self._bufs))]
        F_ = self.F
        bf = _Bufs()
        self._main_marker = bf
        self._trunk_bufs(bf, self.nb, B, h, w, training, dev)
        bf.out1 = None if training else ops.pnhwc_zeros(B, h, w, F_, dev)
        bf.trunk = ops.pnhwc_zeros(B, h, w, F_, dev)
        bf.feat = ops.pnhwc_zeros(B, h, w, F_, dev)
        H, W = h << self.nup, w << self.nup
        bf.fin = None
        if self.nfinal:
            bf.fin = _Bufs()
            self._trunk_bufs(bf.fin, self.nfinal, B, H, W, training, dev)
            bf.hrx = ops.pnhwc_zeros(B, H, W, F_, dev)       # res_blocks_final(out) + out: input of conv3.0
        # output of upsampling stage u; with final blocks the last stage stores into their first growth buffer
        bf.up = [ops.pnhwc_zeros(B, h << (u + 1), w << (u + 1), F_, dev) for u in range(self.nup - (1 if self.nfinal else 0))]
        bf.hr = ops.pnhwc_zeros(B, H, W, F_, dev)
        bf.z = torch.empty(B, self.ch, H, W, device=dev)
        if training:
            bf.g_s = ops.pnhwc_zeros(B, h, w, F_, dev)
            bf.g_out1 = ops.pnhwc_zeros(B, h, w, F_, dev)
            bf.g_hr = ops.pnhwc_zeros(B, H, W, F_, dev)
            bf.g_up = [ops.pnhwc_zeros(B, h << (u + 1), w << (u + 1), F_, dev) for u in range(self.nup)]
            bf.g_uns = [ops.pnhwc_zeros(B, h << u, w << u, 4 * F_, dev) for u in range(self.nup)]
            bf.g_top = ops.pnhwc_zeros(B, H, W, F_, dev) if (self.nfinal > 1 or (self.nfinal and self.drop_rate > 0)) \
                else None
            bf.gz_scratch = torch.empty(B, self.ch, H, W, device=dev)
            ws = 0
            for (hh, ww, cin, cout) in ([(h, w, k * F_, F_) for k in range(1, 6)]
                                        + [(h << u, w << u, F_, 4 * F_) for u in range(self.nup)] + [(H, W, F_, F_)]):
                ws = max(ws, ops.wgrad_workspace_bytes(B, hh, ww, cin, cout))
            ws = max(ws, ops.wgrad_group_workspace_bytes(B, h, w, [k * F_ for k in range(1, 6)], [F_] * 5))
            if self.nfinal:
                ws = max(ws, ops.wgrad_group_workspace_bytes(B, H, W, [k * F_ for k in range(1, 6)], [F_] * 5))
            bf.ws = torch.empty(ws, dtype=torch.uint8, device=dev)
            gp = [(n, p) for n, p in self.p.items() if n not in ("power", "multiplier")]
            bf.gflat = torch.zeros(sum(p.numel() for _, p in gp), device=dev)
            bf.grads = {}
            off = 0
            for n, p in gp:
                bf.grads[n] = bf.gflat[off:off + p.numel()].view(p.shape)
                off += p.numel()
        self._bufs[key] = bf
        return bf

    # ------------------------------------------------------------------ forward
    def _conv(self, x, x_coff, cin, name, y, y_coff, cout, **kw):
        ops.conv3x3(x, x_coff, cin, self.wp[name], self.bias_packed[name], y, y_coff, cout, impl=self.impl,
                    flags=self.conv_flags, **kw)

    @staticmethod
    def _flush(pend: list, arena: _FlagArena, B: int, h: int, w: int) -> None:
        """Launch the pending convolutions: one chained launch, or a plain one if there is only one."""
        if not pend:
            return
        if len(pend) == 1:
            a, flops = pend[0]
            with ops._Prof("conv3x3", flops):
                L.check(L.lib().srb_conv3x3(C.byref(a), _st()), "conv3x3")
        else:
            ops.conv3x3_chain(pend, arena.take(ops.chain_flag_ints(B, h, w, len(pend))))
        pend.clear()

    def _trunk_forward(self, T: _Bufs, prefix: str, nb: int, B: int, h: int, w: int, save: bool, dst: Tensor,
                       masks: Optional[List[Tensor]], res3=None) -> None:
        """models.py:124 / 129: ``nb`` RRDBs over T.growth, input already in slot 0 of T.growth[0], result into ``dst``
        (64-channel buffer).  With self.chain > 0 the convolutions of `chain // 5` consecutive dense blocks go into one
        persistent launch (csrc/conv_chain.cu) instead of one launch each.  ``masks``: Dropout2d factors per dense block
        ([B, F] fp32) or None.  ``res3``: extra residual of the very last convolution (``out3 + out``)."""
        F_ = self.F
        n_buf = len(T.growth)

        class _Ring:
            """growth buffer of dense block d: training keeps one per block; inference cycles (buffer 0 protected when
            the trunk's input is needed again at its end, see _trunk_bufs)"""
            def __getitem__(_, d):
                if T.keep0:
                    return T.growth[0 if d == 0 else 1 + (d - 1) % (n_buf - 1)]
                return T.growth[d % n_buf]
        g = _Ring()
        ng = 1 << 30                    # indices below are block numbers; the ring maps them
        chain = self.chain if (self.chain >= 5 and self.impl == L.IMPL_TCGEN05) else 0
        pend: list = []
        if chain:
            T.flags_f.reset()

        def conv(x, x_coff, cin, name, y, y_coff, cout, **kw):
            if chain:
                pend.append(ops.conv3x3_args(x, x_coff, cin, self.wp[name], self.bias_packed[name], y, y_coff, cout,
                                             impl=self.impl, flags=self.conv_flags, **kw))
            else:
                self._conv(x, x_coff, cin, name, y, y_coff, cout, **kw)

        d = 0
        for i in range(nb):
            first = d
            for j in range(3):
                cur = g[d % ng]
                base = f"{prefix}.{i}.dense_blocks.{j}."
                sign = T.sign[d] if (save and self.impl == L.IMPL_TCGEN05 and T.sign is not None) else None
                for k in range(1, 5):
                    conv(cur, 0, k * F_, base + f"b{k}.0", cur, k * F_, F_, act=L.ACT_LRELU, slope=G_SLOPE,
                         sign_out=(sign, 2 * (k - 1)) if sign is not None else None)
                last = (i == nb - 1 and j == 2)
                out_buf = dst if last else g[(d + 1) % ng]
                cs = masks[d] if masks is not None else None
                if j < 2:
                    conv(cur, 0, 5 * F_, base + "b5.0", out_buf, 0, F_, alpha=DRB_SCALE, res1=(cur, 0, 1.0),
                         chan_scale=cs)
                else:
                    # outer skip of the RRDB: its input is g[first] slot 0 (bf16) or, with the fp32 stream, xs[i % 2]
                    # (RRDB 0 reads the trunk input's bf16 copy); the result also goes to xs[(i + 1) % 2] unrounded
                    fs = self.fp32_stream and T.xs is not None
                    skip = T.xs[i % 2] if (fs and i > 0) else g[first % ng]
                    y32 = (T.xs[(i + 1) % 2], 0) if (fs and not last) else None
                    conv(cur, 0, 5 * F_, base + "b5.0", out_buf, 0, F_, alpha=DRB_SCALE * self.rs,
                         res1=(cur, 0, self.rs), res2=(skip, 0, 1.0), y32=y32, chan_scale=cs,
                         res3=res3 if last else None)
                d += 1
                if chain and len(pend) + 5 > chain:
                    self._flush(pend, T.flags_f, B, h, w)
        self._flush(pend, T.flags_f, B, h, w)

    def draw_drop_masks(self, B: int, dev) -> Optional[Dict[str, List[Tensor]]]:
        """Dropout2d factors (0 or 1/(1-p) per image and channel, models.py:29-31) for every dense block of both
        trunks, drawn with torch's CUDA generator as nn.Dropout2d would."""
        if self.drop_rate <= 0:
            return None
        p, F_ = self.drop_rate, self.F
        n = 3 * (self.nb + self.nfinal)
        m = (torch.rand(n, B, F_, device=dev) >= p).float() / (1.0 - p)
        return {"main": list(m[:3 * self.nb]), "final": list(m[3 * self.nb:])}

    def forward(self, img: Tensor, *, power: float, multiplier: float, thres: float, training: bool,
                save_for_backward: bool, drop_masks=None) -> Tuple[Tensor, Tensor]:
        """models.py:120-135.  ``img`` fp32 [B, channels, h, w] on the GPU.  Returns (out, srs) as fresh tensors.
        ``drop_masks``: {"main": [...], "final": [...]} of [B, F] fp32 factors (tests); drawn here otherwise."""
        if not img.is_cuda:
            raise RuntimeError("CUDA tensor required: super-resolution_b200 has no CPU path for GeneratorRRDB")
        img = img.detach().float().contiguous()
        B, ch, h, w = img.shape
        if ch != self.ch:
            raise RuntimeError(f"CUDA srb200: expected {self.ch} image channels, got {ch}")
        if self.nb == 0:
            raise RuntimeError("CUDA srb200: num_res_blocks must be >= 1")
        self._ensure_packs(save_for_backward)
        bf = self._get_bufs(B, h, w, save_for_backward, img.device)
        F_, lib = self.F, L.lib()
        g = bf.growth
        masks = None
        if training and self.drop_rate > 0:
            masks = drop_masks if drop_masks is not None else self.draw_drop_masks(B, img.device)
            masks = {k: [m.detach().float().contiguous() for m in v] for k, v in masks.items()}
        # conv1 (models.py:122-123) -> slot 0 of the first growth buffer (+ a copy for the long skip at inference)
        y2 = bf.out1
        L.check(lib.srb_conv_in_fwd(img.data_ptr(), B, ch, h, w, float(multiplier), float(power),
                                    self.p["conv1.weight"].data_ptr(), self.p["conv1.bias"].data_ptr(), F_,
                                    g[0].data_ptr(), 5 * F_, 0, y2.data_ptr() if y2 is not None else None, F_, 0,
                                    _st()), "conv_in_fwd")
        # RRDB trunk (models.py:124)
        self._trunk_forward(bf, "res_blocks", self.nb, B, h, w, save_for_backward, bf.trunk,
                            masks["main"] if masks else None)
        # conv2 + long skip (models.py:125-126)
        out1 = (g[0], 0, 1.0) if bf.out1 is None else (bf.out1, 0, 1.0)
        self._conv(bf.trunk, 0, F_, "conv2", bf.feat, 0, F_, res1=out1)
        # upsampling (models.py:127 / 70-91): conv F->4F (tconv: its 1x1 equivalent), LeakyReLU, PixelShuffle(2) fused in
        # the epilogue store; the last stage feeds the final RRDB trunk directly when there is one
        cur = bf.feat
        for u, (_, name) in enumerate(self.ups):
            to_final = self.nfinal and u == self.nup - 1
            dst = bf.fin.growth[0] if to_final else bf.up[u]
            self._conv(cur, 0, F_, name, dst, 0, 4 * F_, act=L.ACT_LRELU, slope=G_SLOPE, shuffle=True)
            cur = dst
        H, W = h << self.nup, w << self.nup
        if self.nfinal:
            # out3 = res_blocks_final(out); out = out3 + out (models.py:128-130): `+ out` is res3 of the last convolution
            self._trunk_forward(bf.fin, "res_blocks_final", self.nfinal, B, H, W, save_for_backward, bf.hrx,
                                masks["final"] if masks else None, res3=(bf.fin.growth[0], 0, 1.0))
            cur = bf.hrx
        # conv3.0 + LeakyReLU (models.py:97-98)
        self._conv(cur, 0, F_, "conv3.0", bf.hr, 0, F_, act=L.ACT_LRELU, slope=G_SLOPE)
        # conv3.2, /multiplier and the output head (models.py:99,131-135,114-118)
        out = torch.empty(B, ch, H, W, device=img.device)
        srs = torch.empty(B, ch, H, W, device=img.device)
        L.check(lib.srb_conv_out_fwd(bf.hr.data_ptr(), F_, 0, B, H, W, F_, self.p["conv3.2.weight"].data_ptr(),
                                     self.p["conv3.2.bias"].data_ptr(), ch, float(multiplier), float(power), float(thres),
                                     int(training), out.data_ptr(), srs.data_ptr(), bf.z.data_ptr(), _st()),
                "conv_out_fwd")
        if save_for_backward:
            self._fwd_gen += 1
            self._saved = dict(img=img, bf=bf, power=float(power), multiplier=float(multiplier), thres=float(thres),
                               training=bool(training), key=self._pack_key, gen=self._fwd_gen, masks=masks)
        return out, srs

    # ------------------------------------------------------------------ backward
    def _wgrad(self, bf, x, cin, gy, gy_coff, cout, name, alpha=1.0, unshuffle=False):
        v = self._virt.get(name)
        dw, db = (v["gw3"], v["gb3"]) if v is not None else (bf.grads[name + ".weight"], bf.grads[name + ".bias"])
        ops.wgrad3x3(x, 0, cin, gy, gy_coff, cout, dw, db, alpha=alpha, unshuffle=unshuffle, workspace=bf.ws,
                     impl=self.impl)
        if v is not None:
            # ConvTranspose2d parameter layout [ci][co][i][j] <- centre tap of the virtual weight's gradient
            F_ = self.F
            bf.grads[name + ".weight"].copy_(dw[:, :, 1, 1].reshape(F_, 2, 2, F_).permute(3, 0, 1, 2))
            bf.grads[name + ".bias"].copy_(db.reshape(F_, 4).sum(1))

    def _dconv(self, x, cin, pack, y, y_coff, cout, **kw):
        ops.conv3x3(x, 0, cin, pack, None, y, y_coff, cout, impl=self.impl, flags=self.conv_flags, **kw)

    def _trunk_backward(self, bf: _Bufs, T: _Bufs, prefix: str, nb: int, B: int, h: int, w: int, dst: Tensor,
                        masks: Optional[List[Tensor]], hook, final_kw: Optional[dict] = None) -> None:
        """Input-gradient chains and weight gradients of ``nb`` RRDBs, last to first.  On entry T.G[2] slot 0 holds the
        gradient w.r.t. the trunk's output (and, with fused bias sums, its column sums in T.csum[2, :F]); the gradient
        w.r.t. the trunk's input goes to ``dst`` (64 channels), with ``final_kw`` merged into that last convolution.
        The five input-gradient convolutions of a dense block are one chained launch (the block's weight gradients
        need its complete gradient buffer, so chains stop there)."""
        F_, lib = self.F, L.lib()
        drop = masks is not None
        fused_bias = self.impl == L.IMPL_TCGEN05 and F_ == 64 and not drop
        cs = T.csum if fused_bias else None

        def col(j, slot):
            return cs[j, slot * F_:(slot + 1) * F_] if fused_bias else None

        chain = self.chain >= 5 and self.impl == L.IMPL_TCGEN05
        pend: list = []
        if chain:
            T.flags_b.reset()

        def dconv(x, cin, pack, y, y_coff, cout, **kw):
            if chain:
                pend.append(ops.conv3x3_args(x, 0, cin, pack, None, y, y_coff, cout, impl=self.impl,
                                             flags=self.conv_flags, **kw))
            else:
                self._dconv(x, cin, pack, y, y_coff, cout, **kw)

        for i in range(nb - 1, -1, -1):
            for j in (2, 1, 0):
                G = T.G[j]
                d = 3 * i + j
                gb = T.growth[d]
                base = f"{prefix}.{i}.dense_blocks.{j}."
                s5 = DRB_SCALE * (self.rs if j == 2 else 1.0)
                sid = self.rs if j == 2 else 1.0
                sign = T.sign[d] if (self.impl == L.IMPL_TCGEN05 and T.sign is not None) else None
                if drop:
                    # Dropout2d: conv 5 sees the masked output gradient (in place in G slot 0), the identity paths the
                    # unmasked copy
                    L.check(lib.srb_chan_mask_split(G.data_ptr(), 5 * F_, 0, T.Gu[j].data_ptr(), F_, 0,
                                                    masks[d].data_ptr(), B, h, w, F_, _st()), "chan_mask_split")
                ident = T.Gu[j] if drop else G            # gradient w.r.t. this block's output, unmasked
                outer = T.Gu[2] if drop else T.G[2]       # gradient w.r.t. the RRDB's output, unmasked
                for jj in (4, 3, 2, 1):
                    mk = dict(mask_bits=(sign, 2 * (jj - 1))) if sign is not None else dict(mask=(gb, jj * F_))
                    dconv(G, (5 - jj) * F_, self.wd[base + f"slot{jj}"], G, (5 - jj) * F_, F_,
                          act=L.ACT_LRELU_MASK, slope=G_SLOPE, colsum=col(j, 5 - jj), **mk)
                if j > 0:
                    dconv(G, 5 * F_, self.wd[base + "slot0"], T.G[j - 1], 0, F_, res1=(ident, 0, sid),
                          colsum=col(j - 1, 0))
                elif i > 0:
                    dconv(G, 5 * F_, self.wd[base + "slot0"], T.G[2], 0, F_, res1=(ident, 0, sid),
                          res2=(outer, 0, 1.0), colsum=col(2, 0))
                else:
                    dconv(G, 5 * F_, self.wd[base + "slot0"], dst, 0, F_, res1=(ident, 0, sid),
                          res2=(outer, 0, 1.0), **(final_kw or {}))
                self._flush(pend, T.flags_b, B, h, w)
                if self.impl == L.IMPL_TCGEN05:      # the block's five weight gradients in one launch
                    ops.wgrad3x3_group(gb, G, [(0, k * F_, (5 - k) * F_ if k < 5 else 0, F_,
                                                bf.grads[base + f"b{k}.0.weight"], bf.grads[base + f"b{k}.0.bias"],
                                                s5 if k == 5 else 1.0) for k in range(1, 6)], bf.ws,
                                       gy_colsum=cs[j] if fused_bias else None)
                else:
                    for k in range(1, 6):
                        self._wgrad(bf, gb, k * F_, G, (5 - k) * F_ if k < 5 else 0, F_, base + f"b{k}.0",
                                    alpha=s5 if k == 5 else 1.0)
            hook(f"{prefix}.{i}.dense_blocks.0.b1.0.weight")

    def backward(self, g_out: Optional[Tensor], g_srs: Optional[Tensor], fwd_gen: Optional[int] = None
                 ) -> Dict[str, Tensor]:
        """Gradients of every parameter for the last ``forward(save_for_backward=True)``, given d loss / d out and
        d loss / d srs (fp32 [B, ch, H, W] or None).  Returns name -> fp32 tensor views of one flat buffer (valid
        until the next backward of the same shape).  ``fwd_gen``: the forward this backward belongs to (autograd
        passes it); the activation buffers hold ONE forward, so a backward for an older one is refused."""
        sv = self._saved
        if sv is None:
            raise RuntimeError("CUDA srb200: backward called without a saved forward")
        if fwd_gen is not None and fwd_gen != sv["gen"]:
            raise RuntimeError("CUDA srb200: GeneratorRRDB keeps the activations of its most recent grad-enabled "
                               "forward only; backward was asked for an earlier forward (two forwards before one "
                               "backward). Run forward/backward pairs one after the other, as esrgan.py:468,553 does")
        if sv["key"] != self._params_key():
            raise RuntimeError("CUDA srb200: parameters changed between forward and backward")
        bf, img, masks = sv["bf"], sv["img"], sv["masks"]
        if self.grad_views is not None:
            bf.grads = self.grad_views
        B, ch, h, w = img.shape
        F_, lib = self.F, L.lib()
        H, W = h << self.nup, w << self.nup
        go = g_out.detach().float().contiguous() if g_out is not None else None
        gs = g_srs.detach().float().contiguous() if g_srs is not None else None
        fin = bf.fin
        up_out = [fin.growth[0] if (fin is not None and u == self.nup - 1) else bf.up[u] for u in range(self.nup)]
        top_in = bf.hrx if fin is not None else (up_out[-1] if self.nup else bf.feat)      # input of conv3.0
        # conv3.2 + head (models.py:99,131-135): gradient of conv3.0's activation, masked by LeakyReLU'
        L.check(lib.srb_conv_out_bwd(bf.z.data_ptr(), go.data_ptr() if go is not None else None,
                                     gs.data_ptr() if gs is not None else None, B, ch, H, W, sv["multiplier"],
                                     sv["power"], sv["thres"], int(sv["training"]), self.p["conv3.2.weight"].data_ptr(),
                                     F_, bf.hr.data_ptr(), F_, 0, G_SLOPE, bf.gz_scratch.data_ptr(), bf.g_hr.data_ptr(),
                                     F_, 0, bf.grads["conv3.2.weight"].data_ptr(), bf.grads["conv3.2.bias"].data_ptr(),
                                     _st()), "conv_out_bwd")
        hook = self.ready_hook or (lambda name: None)
        # conv3.0
        self._wgrad(bf, top_in, F_, bf.g_hr, 0, F_, "conv3.0")
        hook("conv3.0.weight")
        if fin is not None:
            # gradient w.r.t. `out3 + out` -> output gradient of the final trunk (fin.G[2] slot 0, no activation in
            # between); the skip's share is res3 of the trunk's last input-gradient convolution, whose result - the
            # gradient w.r.t. the last upsampling stage's activation - carries that stage's LeakyReLU' mask
            fused = self.impl == L.IMPL_TCGEN05 and F_ == 64 and masks is None
            if fused:
                fin.csum.zero_()
            self._dconv(bf.g_hr, F_, self.wd["conv3.0"], fin.G[2], 0, F_, colsum=fin.csum[2, :F_] if fused else None)
            g_top = fin.G[2]
            if bf.g_top is not None:       # fin.G[2] is recycled per RRDB (and masked in place under dropout): keep the
                                           # top gradient for the skip
                bf.g_top.copy_(fin.G[2][..., :F_])
                g_top = bf.g_top
            self._trunk_backward(bf, fin, "res_blocks_final", self.nfinal, B, H, W, bf.g_up[-1],
                                 masks["final"] if masks else None, hook,
                                 final_kw=dict(res3=(g_top, 0, 1.0), act=L.ACT_LRELU_MASK, slope=G_SLOPE,
                                               mask=(up_out[-1], 0)))
        elif self.nup:
            self._dconv(bf.g_hr, F_, self.wd["conv3.0"], bf.g_up[-1], 0, F_, act=L.ACT_LRELU_MASK, slope=G_SLOPE,
                        mask=(up_out[-1], 0))
        else:
            self._dconv(bf.g_hr, F_, self.wd["conv3.0"], bf.g_s, 0, F_)
        # upsampling stages, last to first
        for u in range(self.nup - 1, -1, -1):
            hh, ww = h << u, w << u
            L.check(lib.srb_unshuffle2(bf.g_up[u].data_ptr(), F_, 0, B, hh, ww, F_, bf.g_uns[u].data_ptr(), 4 * F_, 0,
                                       _st()), "unshuffle")
            name = self.ups[u][1]
            x_in = bf.feat if u == 0 else up_out[u - 1]
            self._wgrad(bf, x_in, F_, bf.g_uns[u], 0, 4 * F_, name, unshuffle=True)
            hook(name + ".weight")
            if u == 0:
                self._dconv(bf.g_uns[u], 4 * F_, self.wd[name], bf.g_s, 0, F_)
            else:
                self._dconv(bf.g_uns[u], 4 * F_, self.wd[name], bf.g_up[u - 1], 0, F_, act=L.ACT_LRELU_MASK,
                            slope=G_SLOPE, mask=(up_out[u - 1], 0))
        # conv2 (+ the long skip: g_s also flows straight into conv1's output, added in conv_in_bwd)
        self._wgrad(bf, bf.trunk, F_, bf.g_s, 0, F_, "conv2")
        hook("conv2.weight")
        # Bias gradients of the dense blocks: the kernels that write a gradient slot also add up its columns
        # (srb_conv3x3_args.colsum), so the weight-gradient launch needs no pass over the gradient buffer.
        fused = self.impl == L.IMPL_TCGEN05 and F_ == 64 and masks is None
        if fused:
            bf.csum.zero_()
        self._dconv(bf.g_s, F_, self.wd["conv2"], bf.G[2], 0, F_, colsum=bf.csum[2, :F_] if fused else None)
        # RRDB trunk, last to first
        self._trunk_backward(bf, bf, "res_blocks", self.nb, B, h, w, bf.g_out1, masks["main"] if masks else None, hook)
        # conv1 (models.py:63,122-123): weight gradient from g_out1 + g_s
        L.check(lib.srb_conv_in_bwd(img.data_ptr(), B, ch, h, w, sv["multiplier"], sv["power"], bf.g_out1.data_ptr(),
                                    F_, 0, bf.g_s.data_ptr(), F_, 0, F_, bf.grads["conv1.weight"].data_ptr(),
                                    bf.grads["conv1.bias"].data_ptr(), _st()), "conv_in_bwd")
        return bf.grads
