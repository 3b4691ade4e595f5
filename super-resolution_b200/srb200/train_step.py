"""One ESRGAN training iteration (the tensor math of esrgan.py:399-626, default-active terms) on the sm_100a
kernels, numerically the same schedule as the reference:

  G step  (esrgan.py:457-555): forward, per-k losses on (out, srs): L1 of batch-mean image (487), L1 of the pooled LR
          image (471,490), relativistic-average adversarial loss through D_k (493-508); backward; Adam.
  D step  (esrgan.py:561-626), per k: RaGAN loss (569-588), gradient penalty via a differentiable input gradient
          (596-606), backward, Adam gated on loss_D > d_threshold (623-624) - the gate is evaluated on the device.

The reference's in-step host syncs (.item() at models.py:116,133 and esrgan.py:623,645) are gone; losses are
returned as device scalars.  With a process group, gradients (and the gate value) are all-reduced over NCCL:
one contiguous message per optimizer, launched as soon as that optimizer's backward has finished."""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence

import torch

from . import disc_engine as DE
from . import dp as DP
from . import discriminator as D
from . import losses as LS
from .optim import FusedAdam


class GANTrainStep:
    def __init__(self, generator, discriminators: Sequence, *, factor: int, lambdas=(0.2, 1.0), lambda_hr=1.0,
                 lambda_lr=0.1, lambda_adv=0.01, lambda_reg=0.01, d_threshold=0.001, lr=2e-4, lr_g=0.0, lr_d=0.0,
                 betas=(0.9, 0.999), l2decay=0.0, scaling_power=1.0, process_group=None, use_disc_engine=True):
        if scaling_power != 1:
            raise RuntimeError("CUDA srb200: the fused train step covers scaling_power == 1 (options/default.json); "
                               "other powers train through the module API + autograd")
        self.G, self.Ds = generator, list(discriminators)
        self.factor, self.lambdas = int(factor), tuple(lambdas)
        self.l_hr, self.l_lr, self.l_adv, self.l_reg = lambda_hr, lambda_lr, lambda_adv, lambda_reg
        self.d_threshold = d_threshold
        self.opt_G = FusedAdam(generator.parameters(), lr=lr_g if lr_g > 0 else lr, betas=betas, weight_decay=l2decay)
        self.opt_D = [FusedAdam(d.parameters(), lr=lr_d if lr_d > 0 else lr, betas=betas) if d is not None else None
                      for d in self.Ds]
        self.pg = process_group
        self.world = torch.distributed.get_world_size(process_group) if process_group is not None else 1
        self._work = {}
        # The discriminator steps need nothing from the generator's backward pass (only the detached generated images
        # and the discriminator forward already computed in the generator step), so they can run on a second stream
        # beside it: bandwidth / latency-bound kernels next to tensor-bound ones.  SRB_OVERLAP_D=0/1.
        self.overlap_d = os.environ.get("SRB_OVERLAP_D", "0") == "1"
        self._d_stream = None
        # bucketed gradient all-reduce overlapped with the generator's backward (srb200/dp.py)
        self._g_sync = DP.BucketedAllReduce(self.opt_G.flat_g, process_group) if self.world > 1 else None
        self._g_off = {n: o for (n, _), o in zip(((n, p) for n, p in generator.named_parameters() if p.requires_grad),
                                                 self.opt_G.offsets)}
        # tensor-core discriminator engine where the configuration allows it; the general fp32 autograd path otherwise
        self.d_eng = []
        for d in self.Ds:
            ok = (use_disc_engine and d is not None and type(d).__name__ == "Markovian_Discriminator"
                  and DE.supported(d.model, d.input_shape))
            self.d_eng.append(DE.DiscEngine(d.model, d.input_shape) if ok else None)

    # ------------------------------------------------------------------ helpers
    def _pix_work(self, B, H, W, dev):
        key = (B, H, W)
        w = self._work.get(key)
        if w is None:
            w = dict(w0=LS.PixelLossWork(B, H, W, self.factor, dev), w1=LS.PixelLossWork(B, H, W, self.factor, dev),
                     g0=torch.empty(B, 1, H, W, device=dev), g1=torch.empty(B, 1, H, W, device=dev))
            self._work[key] = w
        return w

    def _allreduce(self, flat: torch.Tensor):
        if self.world > 1:
            torch.distributed.all_reduce(flat, group=self.pg)

    # ------------------------------------------------------------------ the step
    def step(self, imgs_lr: torch.Tensor, imgs_hr: torch.Tensor, epsilons=None) -> Dict[str, torch.Tensor]:
        """``epsilons``: optional per-discriminator (B,1,1,1) interpolation draws (esrgan.py:598) for reproducible
        tests; drawn with torch.rand on the device otherwise."""
        G, eng = self.G, self.G._engine()
        eng.grad_views = {n: gv for (n, p), gv in zip(((n, p) for n, p in G.named_parameters() if p.requires_grad),
                                                     self.opt_G.grad_views)}
        B, _, H, W = imgs_hr.shape
        dev = imgs_hr.device
        wk = self._pix_work(B, H, W, dev)
        lam0, lam1 = self.lambdas
        out_losses: Dict[str, torch.Tensor] = {}

        # ---------------- generator ----------------
        out, srs = eng.forward(imgs_lr, power=G._host_scalar("power"), multiplier=G._host_scalar("multiplier"),
                               thres=float(G.thres), training=G.training, save_for_backward=True)
        G.srs = srs
        generated = (out, srs)
        g = (wk["g0"], wk["g1"])
        # pooled-LR term always differentiates generated[0] (esrgan.py:471-472), for both k
        w_lr0 = (lam0 if lam0 > 0 else 0.0) * self.l_lr + (lam1 if lam1 > 0 else 0.0) * self.l_lr
        l_hr0, l_lr0 = LS.pixel_losses(out, imgs_hr, imgs_lr, self.factor, max(lam0, 0.0) * self.l_hr, w_lr0, g[0],
                                       False, wk["w0"])
        l_hr1, _ = LS.pixel_losses(srs, imgs_hr, imgs_lr, self.factor, max(lam1, 0.0) * self.l_hr, 0.0, g[1], False,
                                   wk["w1"])
        out_losses.update(hr0=l_hr0, lr0=l_lr0, hr1=l_hr1, lr1=l_lr0)
        loss_G = lam0 * (self.l_hr * l_hr0 + self.l_lr * l_lr0) + lam1 * (self.l_hr * l_hr1 + self.l_lr * l_lr0)
        for k in range(2):
            if self.lambdas[k] <= 0 or self.Ds[k] is None or self.l_adv <= 0:
                continue
            d = self.Ds[k]
            w_adv = self.lambdas[k] * self.l_adv
            if self.d_eng[k] is not None:
                eng_d = self.d_eng[k]
                A = eng_d.forward([imgs_hr, generated[k]], tag="rf")                       # esrgan.py:493-494 (one 2B batch)
                l_adv, gu, _ = LS.ragan_raw(A.pred[B:], A.pred[:B])                        # esrgan.py:507-508
                g_adv, _ = eng_d.input_grad(A, gu * w_adv, B, B, tag="ig")
            else:
                with torch.no_grad():
                    pred_real = D.sequential_forward(d.model, imgs_hr)                    # esrgan.py:493 (.detach())
                x = generated[k].detach().requires_grad_(True)
                with torch.enable_grad():
                    pred_fake = D.sequential_forward(d.model, x, detach_params=True)      # esrgan.py:494
                    l_adv = LS.RaGAN.apply(pred_fake, pred_real)                          # esrgan.py:507-508
                    (g_adv,) = torch.autograd.grad(l_adv * w_adv, x)
            g[k].add_(g_adv)
            out_losses[f"adv{k}"] = l_adv.detach()
            loss_G = loss_G + self.lambdas[k] * self.l_adv * l_adv.detach()
        main = torch.cuda.current_stream()
        if self.overlap_d:
            if self._d_stream is None:
                self._d_stream = torch.cuda.Stream()
            fork = torch.cuda.Event()
            fork.record(main)          # the discriminator steps depend on the work up to here only
        if self._g_sync is not None:
            self._g_sync.reset()
            eng.ready_hook = lambda name: self._g_sync.mark_ready(self._g_off[name])
        grads = eng.backward(g[0], g[1])                                                   # esrgan.py:553
        eng.ready_hook = None
        for n, p in G.named_parameters():
            if p.requires_grad:
                p.grad = grads[n]                      # views of the optimizer's flat gradient buffer: no copy
        self.opt_G.gather_grads()
        if self._g_sync is not None:
            self._g_sync.finish()
        self.opt_G.step(grad_scale=1.0 / self.world, gathered=True)                        # esrgan.py:555
        out_losses["g_loss"] = loss_G
        if self.overlap_d:
            # issued after the generator's backward on the host (FusedAdam.step bumps the global parameter epoch the
            # generator engine checks), but ordered after `fork` only: inside the CUDA graph they are a parallel branch
            self._d_stream.wait_event(fork)
            with torch.cuda.stream(self._d_stream):
                self._discriminator_steps(imgs_hr, generated, epsilons, out_losses, B, dev)
            join = torch.cuda.Event()
            join.record(self._d_stream)
            main.wait_event(join)
        else:
            self._discriminator_steps(imgs_hr, generated, epsilons, out_losses, B, dev)
        return out_losses

    def _discriminator_steps(self, imgs_hr, generated, epsilons, out_losses, B, dev) -> None:
        """esrgan.py:561-626 for every active discriminator (on the current stream)."""
        for k in range(2):
            d = self.Ds[k]
            if self.lambdas[k] <= 0 or d is None:
                continue
            opt = self.opt_D[k]
            opt.zero_grad()
            real, fake = imgs_hr, generated[k].detach()
            if self.d_eng[k] is not None:
                eng_d = self.d_eng[k]
                # D_k's weights have not moved since the generator step: its forward over (real, fake) is reused
                A = eng_d.acts(2 * B, dev, "rf") if (self.lambdas[k] > 0 and self.l_adv > 0) else \
                    eng_d.forward([real, fake], tag="rf")
                loss_D, gu, gv = LS.ragan_raw(A.pred[:B], A.pred[B:])                      # esrgan.py:578-588
                opt.flat_g.zero_()
                grads = [(opt.grad_views[2 * l], opt.grad_views[2 * l + 1]) for l in range(eng_d.L)]
                eng_d.backward(A, torch.cat([gu, gv]), grads, tag="bw")
                if self.l_reg > 0:
                    eps = epsilons[k].to(dev) if epsilons is not None else torch.rand(B, 1, 1, 1, device=dev)   # 598
                    interp = eps * real + (1 - eps) * fake                                 # 599-600
                    Ai = eng_d.forward([interp], tag="gp")                                 # 601
                    gi, hz = eng_d.input_grad(Ai, torch.ones_like(Ai.pred), 0, B, tag="gph")   # 602
                    gp, q0 = LS.gradient_penalty_raw(gi.reshape(B, -1), self.l_reg / 2)    # 604-605
                    eng_d.penalty_backward(Ai, hz, q0.view_as(gi), grads)
                    loss_D = loss_D + gp                                                   # 606
                    out_losses[f"gp{k}"] = gp
                for p_, gv_ in zip(opt.params, opt.grad_views):
                    p_.grad = gv_
            else:
                pred_real = D.sequential_forward(d.model, real)                            # esrgan.py:569
                pred_fake = D.sequential_forward(d.model, fake)                            # esrgan.py:570
                loss_D = LS.RaGAN.apply(pred_real, pred_fake)                              # esrgan.py:578-588
                if self.l_reg > 0:
                    eps = epsilons[k].to(dev) if epsilons is not None else torch.rand(B, 1, 1, 1, device=dev)   # 598
                    interp = eps * real + (1 - eps) * fake                                 # 599-600
                    _, gi = D.input_gradient(d.model, interp)                              # 601-602
                    gp = LS.GradientPenalty.apply(gi.reshape(B, -1), self.l_reg / 2)       # 604-605
                    loss_D = loss_D + gp                                                   # 606
                    out_losses[f"gp{k}"] = gp.detach()
                loss_D.backward()                                                          # esrgan.py:616
            gate = loss_D.detach().clone()
            opt.gather_grads()
            if self.world > 1:
                self._allreduce(opt.flat_g)
                DP.sync_gate(gate, self.pg)
            opt.step(gate=(gate, self.d_threshold), grad_scale=1.0 / self.world, gathered=True)   # 623-624
            out_losses[f"d_loss{k}"] = loss_D.detach()


class GraphedStep:
    """One CUDA graph for the whole iteration: ~1500 kernel launches per step become a single graph launch, which
    removes the host-side launch gaps (about 4 ms of a 31 ms step at B=256) and the Python/ctypes cost per rank.
    Inputs are copied into static buffers; the returned losses are static device tensors overwritten by every replay.
    The interpolation draws of the gradient penalty (esrgan.py:598) come from the CUDA generator, whose Philox
    offset torch advances per replay.  NCCL all-reduces issued inside the step are captured with it."""

    def __init__(self, trainer, imgs_lr: torch.Tensor, imgs_hr: torch.Tensor, warmup: int = 3, epsilons=None):
        """``epsilons``: optional per-discriminator (B,1,1,1) device tensors - the graph then reads the interpolation
        draws of esrgan.py:598 from static buffers that step(..., epsilons=...) refills (reproducible tests)."""
        self.tr = trainer
        self.lr = imgs_lr.detach().clone()
        self.hr = imgs_hr.detach().clone()
        self.eps = [e.detach().clone() for e in epsilons] if epsilons is not None else None
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(max(warmup, 1)):          # allocates every cached buffer, sets kernel attributes
                trainer.step(self.lr, self.hr, epsilons=self.eps)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.losses = trainer.step(self.lr, self.hr, epsilons=self.eps)

    def step(self, imgs_lr: torch.Tensor, imgs_hr: torch.Tensor, epsilons=None):
        if imgs_lr.shape != self.lr.shape or imgs_hr.shape != self.hr.shape:
            # the reference's DataLoader has no drop_last (esrgan.py:311-317): the last batch of an epoch is partial.
            # The graph is shape-specific, so such a batch takes the eager path (same kernels, same optimizers).
            return self.tr.step(imgs_lr, imgs_hr, epsilons=epsilons)
        self.lr.copy_(imgs_lr, non_blocking=True)
        self.hr.copy_(imgs_hr, non_blocking=True)
        if self.eps is not None and epsilons is not None:
            for dst, src in zip(self.eps, epsilons):
                dst.copy_(src, non_blocking=True)
        for opt in [self.tr.opt_G] + [o for o in self.tr.opt_D if o is not None]:
            opt.sync_lr()                        # ReduceLROnPlateau may have moved it since the last replay
        self.graph.replay()
        return self.losses


class GeneratorTrainStep:
    """The warm-up iteration (esrgan.py:418-427): generator forward, L1 against the HR image, backward, Adam."""

    def __init__(self, generator, lr=2e-4, betas=(0.9, 0.999), l2decay=0.0, process_group=None):
        self.G = generator
        self.opt_G = FusedAdam(generator.parameters(), lr=lr, betas=betas, weight_decay=l2decay)
        self.pg = process_group
        self.world = torch.distributed.get_world_size(process_group) if process_group is not None else 1

    def step(self, imgs_lr, imgs_hr):
        G, eng = self.G, self.G._engine()
        eng.grad_views = {n: gv for (n, p), gv in zip(((n, p) for n, p in G.named_parameters() if p.requires_grad),
                                                     self.opt_G.grad_views)}
        out, srs = eng.forward(imgs_lr, power=G._host_scalar("power"), multiplier=G._host_scalar("multiplier"),
                               thres=float(G.thres), training=G.training, save_for_backward=True)
        G.srs = srs
        diff = out - imgs_hr
        loss = diff.abs().mean()                                   # nn.L1Loss, esrgan.py:424
        g_out = torch.sign(diff) / diff.numel()
        grads = eng.backward(g_out, None)
        for n, p in G.named_parameters():
            if p.requires_grad:
                p.grad = grads[n]
        self.opt_G.gather_grads()
        if self.world > 1:
            torch.distributed.all_reduce(self.opt_G.flat_g, group=self.pg)
        self.opt_G.step(grad_scale=1.0 / self.world, gathered=True)
        return loss
