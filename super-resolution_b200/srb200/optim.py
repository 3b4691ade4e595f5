"""FusedAdam: torch.optim.Adam semantics (esrgan.py:299,305: betas, eps 1e-8, coupled L2 weight_decay, no amsgrad)
as ONE kernel over a flat fp32 buffer.  Parameters are re-pointed at views of that buffer (values preserved), so
state_dict()/load_state_dict()/checkpoints keep working; m, v and the gradient staging buffer share the layout, which
is also what the data-parallel allreduce operates on (one contiguous message per optimizer).

`step(gate=(loss, threshold))` applies the update only where loss > threshold, decided on the device - the
reference's `if loss_D.item() > opt.d_threshold` (esrgan.py:623) without the host round trip.

It IS a torch.optim.Optimizer, so torch.optim.lr_scheduler.ReduceLROnPlateau (esrgan.py:308-310, stepped at 666-667)
wraps it unchanged.  The learning rate the kernel uses lives on the DEVICE (dev_state[2]); `sync_lr()` mirrors
param_groups[0]['lr'] into it with one tiny fill whenever it has changed.  A captured CUDA graph (GraphedStep) therefore
follows the scheduler: it calls sync_lr() before every replay instead of baking lr into the capture."""
from __future__ import annotations

from typing import Iterable, Optional, Tuple

import torch

from . import _lib as L
from . import generator as _gen


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params: Iterable[torch.nn.Parameter], lr=2e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        self.all_params = list(params)
        trainable = [p for p in self.all_params if p.requires_grad]
        if not trainable:
            raise ValueError("FusedAdam got no trainable parameters")
        dev = trainable[0].device
        if dev.type != "cuda":
            raise RuntimeError("CUDA parameters required: FusedAdam has no CPU path")
        super().__init__(trainable, dict(lr=float(lr), betas=tuple(betas), eps=float(eps),
                                         weight_decay=float(weight_decay)))
        self.params = trainable
        self.betas, self.eps, self.weight_decay = tuple(betas), float(eps), float(weight_decay)
        # 16-byte aligned slots so every view can also be used as a TMA / float4 base
        offs, n = [], 0
        for p in self.params:
            offs.append(n)
            n += (p.numel() + 3) // 4 * 4
        self.n = n
        self.offsets = offs                      # start of each trainable parameter in the flat buffers
        self.flat_p = torch.zeros(n, device=dev)
        self.flat_g = torch.zeros(n, device=dev)
        self.flat_m = torch.zeros(n, device=dev)
        self.flat_v = torch.zeros(n, device=dev)
        self.dev_state = torch.zeros(4, device=dev)      # {step, skip, lr, reserved} (include/srb200.h)
        self._lr_on_device = None
        self.grad_views = []
        with torch.no_grad():
            for p, o in zip(self.params, offs):
                view = self.flat_p[o:o + p.numel()].view(p.shape)
                view.copy_(p.data)
                p.data = view
                self.grad_views.append(self.flat_g[o:o + p.numel()].view(p.shape))
        self.sync_lr()
        _gen.PARAM_EPOCH[0] += 1

    @property
    def lr(self) -> float:
        return float(self.param_groups[0]["lr"])

    def sync_lr(self) -> None:
        """Mirror param_groups[0]['lr'] into the device state (no-op while it is unchanged).  Must run OUTSIDE a graph
        capture; GraphedStep calls it before every replay."""
        lr = self.lr
        if lr != self._lr_on_device:
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError("CUDA srb200: the learning rate changed inside a CUDA graph capture")
            self.dev_state[2:3].fill_(lr)
            self._lr_on_device = lr

    def zero_grad(self, set_to_none: bool = True):
        for p in self.params:
            p.grad = None

    def gather_grads(self) -> torch.Tensor:
        """Stage p.grad into the flat gradient buffer (no copy for gradients that already live there)."""
        src, dst = [], []
        for p, gv in zip(self.params, self.grad_views):
            if p.grad is None:
                gv.zero_()
            elif p.grad.data_ptr() != gv.data_ptr():
                src.append(p.grad)
                dst.append(gv)
        if src:
            torch._foreach_copy_(dst, src)
        return self.flat_g

    @torch.no_grad()
    def step(self, closure=None, *, gate: Optional[Tuple[torch.Tensor, float]] = None, grad_scale: float = 1.0,
             gathered: bool = False):
        if closure is not None:
            raise RuntimeError("CUDA srb200: FusedAdam takes no closure")
        if not gathered:
            self.gather_grads()
        if not torch.cuda.is_current_stream_capturing():
            self.sync_lr()
        gv, thr = (gate[0].detach().float().reshape(-1), float(gate[1])) if gate is not None else (None, 0.0)
        L.check(L.lib().srb_adam_step(self.flat_p.data_ptr(), self.flat_g.data_ptr(), self.flat_m.data_ptr(),
                                      self.flat_v.data_ptr(), self.n, -1.0, float(self.betas[0]),
                                      float(self.betas[1]), self.eps, self.weight_decay, float(grad_scale),
                                      self.dev_state.data_ptr(), gv.data_ptr() if gv is not None else None, thr,
                                      torch.cuda.current_stream().cuda_stream), "adam_step")
        _gen.PARAM_EPOCH[0] += 1

    def state_dict(self):
        return dict(m=self.flat_m.clone(), v=self.flat_v.clone(), state=self.dev_state.clone(), lr=self.lr)

    def load_state_dict(self, sd):
        self.flat_m.copy_(sd["m"]); self.flat_v.copy_(sd["v"])
        st = sd["state"]
        self.dev_state[:st.numel()].copy_(st)
        if "lr" in sd:
            self.param_groups[0]["lr"] = float(sd["lr"])
        self._lr_on_device = None
        self.sync_lr()
