"""FusedAdam: torch.optim.Adam semantics (esrgan.py:299,305: betas, eps 1e-8, coupled L2 weight_decay, no amsgrad)
as ONE kernel over a flat fp32 buffer.  Parameters are re-pointed at views of that buffer (values preserved), so
state_dict()/load_state_dict()/checkpoints keep working; m, v and the gradient staging buffer share the layout, which
is also what the data-parallel allreduce operates on (one contiguous message per optimizer).

`step(gate=(loss, threshold))` applies the update only where loss > threshold, decided on the device - the
reference's `if loss_D.item() > opt.d_threshold` (esrgan.py:623) without the host round trip."""
from __future__ import annotations

from typing import Iterable, Optional, Tuple

import torch

from . import _lib as L
from . import generator as _gen


class FusedAdam:
    def __init__(self, params: Iterable[torch.nn.Parameter], lr=2e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        self.all_params = list(params)
        self.params = [p for p in self.all_params if p.requires_grad]
        if not self.params:
            raise ValueError("FusedAdam got no trainable parameters")
        dev = self.params[0].device
        if dev.type != "cuda":
            raise RuntimeError("CUDA parameters required: FusedAdam has no CPU path")
        self.lr, self.betas, self.eps, self.weight_decay = float(lr), betas, float(eps), float(weight_decay)
        self.param_groups = [dict(params=self.params, lr=self.lr, betas=betas, eps=eps, weight_decay=weight_decay)]
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
        self.state = torch.zeros(2, device=dev)
        self.grad_views = []
        with torch.no_grad():
            for p, o in zip(self.params, offs):
                view = self.flat_p[o:o + p.numel()].view(p.shape)
                view.copy_(p.data)
                p.data = view
                self.grad_views.append(self.flat_g[o:o + p.numel()].view(p.shape))
        _gen.PARAM_EPOCH[0] += 1

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

    def step(self, gate: Optional[Tuple[torch.Tensor, float]] = None, grad_scale: float = 1.0, gathered: bool = False):
        if not gathered:
            self.gather_grads()
        lr = self.param_groups[0]["lr"]
        gv, thr = (gate[0].detach().float().reshape(-1), float(gate[1])) if gate is not None else (None, 0.0)
        L.check(L.lib().srb_adam_step(self.flat_p.data_ptr(), self.flat_g.data_ptr(), self.flat_m.data_ptr(),
                                      self.flat_v.data_ptr(), self.n, float(lr), float(self.betas[0]),
                                      float(self.betas[1]), self.eps, self.weight_decay, float(grad_scale),
                                      self.state.data_ptr(), gv.data_ptr() if gv is not None else None, thr,
                                      torch.cuda.current_stream().cuda_stream), "adam_step")
        _gen.PARAM_EPOCH[0] += 1

    def state_dict(self):
        return dict(m=self.flat_m.clone(), v=self.flat_v.clone(), state=self.state.clone(), lr=self.lr)

    def load_state_dict(self, sd):
        self.flat_m.copy_(sd["m"]); self.flat_v.copy_(sd["v"]); self.state.copy_(sd["state"])
