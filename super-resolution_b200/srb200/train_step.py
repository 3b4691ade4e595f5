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


class _NullCtx:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


class GANTrainStep:
    def __init__(self, generator, discriminators: Sequence, *, factor: int, lambdas=(0.2, 1.0), lambda_hr=1.0,
                 lambda_lr=0.1, lambda_adv=0.01, lambda_reg=0.01, d_threshold=0.001, lr=2e-4, lr_g=0.0, lr_d=0.0,
                 betas=(0.9, 0.999), l2decay=0.0, scaling_power=1.0, process_group=None, use_disc_engine=True,
                 second_discriminators: Optional[Sequence] = None, second_discr_reset_interval: int = 0,
                 uniform_reset: bool = False, lambda_nnz=0.0, lambda_mask=0.0, lambda_hist=0.0, lambda_hit=0.0,
                 binedges=None, sigma=500.0, hit_threshold=0.5, lambda_wasser=0.0, sinkhorn_eps=0.1):
        """``scaling_power``: opt.scaling_power (esrgan.py:470-473): the "pow" branch (k = 1) compares generator.srs with
        hr ** p, the pooled image ** p with lr ** p, and shows D_pow the powered images; the generator must have been
        built with power=p.  ``second_discriminators``: the SecondDiscriminators ensemble of esrgan.py:221-237 (one per
        branch): the adversarial term averages both discriminators (495-505), both are trained every iteration
        (565-620, stepping together on the first one's d_threshold test, 623-626) and the second ones are
        re-initialised every ``second_discr_reset_interval`` iterations (789-795).
        ``lambda_nnz / lambda_mask / lambda_hist / lambda_hit``: the optional terms of esrgan.py:530-550 on fused kernels
        (csrc/opt_losses.cu); ``binedges`` = per branch the histogram bin edges esrgan.py:440-450 derives during warm-up
        (info['binedges0'], info['binedges1']), ``sigma`` = opt.sigma, ``hit_threshold`` = opt.hit_threshold.
        ``lambda_wasser`` / ``sinkhorn_eps``: the sorted-Wasserstein / Sinkhorn term of esrgan.py:539-542 (csrc/sinkhorn.cu)."""
        self.l_nnz, self.l_mask, self.l_hist, self.l_hit = lambda_nnz, lambda_mask, lambda_hist, lambda_hit
        self.l_wasser, self.sinkhorn_eps = float(lambda_wasser), float(sinkhorn_eps)
        self.binedges, self.sigma, self.hit_threshold = binedges, float(sigma), float(hit_threshold)
        if lambda_hist > 0 and binedges is None:
            raise RuntimeError("CUDA srb200: lambda_hist > 0 needs the bin edges (esrgan.py:440-450)")
        self.power = float(scaling_power)
        self.G, self.Ds = generator, list(discriminators)
        self.D2s = list(second_discriminators) if second_discriminators is not None else [None] * len(self.Ds)
        self.reset_interval, self.uniform_reset = int(second_discr_reset_interval), bool(uniform_reset)
        self.batches_done = -1
        self.factor, self.lambdas = int(factor), tuple(lambdas)
        self.l_hr, self.l_lr, self.l_adv, self.l_reg = lambda_hr, lambda_lr, lambda_adv, lambda_reg
        self.d_threshold = d_threshold
        self.opt_G = FusedAdam(generator.parameters(), lr=lr_g if lr_g > 0 else lr, betas=betas, weight_decay=l2decay)
        self.opt_D = [FusedAdam(d.parameters(), lr=lr_d if lr_d > 0 else lr, betas=betas) if d is not None else None
                      for d in self.Ds]
        self.opt_D2 = [FusedAdam(d.parameters(), lr=lr_d if lr_d > 0 else lr, betas=betas) if d is not None else None
                       for d in self.D2s]
        self.pg = process_group
        self.world = torch.distributed.get_world_size(process_group) if process_group is not None else 1
        self._work = {}
        # The discriminator steps need nothing from the generator's backward pass (only the detached generated images
        # and the discriminator forward already computed in the generator step), so they can run on a second stream
        # beside it: bandwidth / latency-bound kernels next to tensor-bound ones.  SRB_OVERLAP_D=0/1.
        self.overlap_d = os.environ.get("SRB_OVERLAP_D", "0") == "1"
        self._d_stream = None
        # The two discriminators never exchange anything inside an iteration, and their kernels are small and
        # latency-bound (15-45 % of HBM each, profiles/r02_s3_ncu_step_kernels_dram.txt).  SRB_D_STREAMS=1 (default) runs
        # discriminator k's work - forward over (real, fake), input gradient for the generator's adversarial term, and
        # later its own training step - on its own stream, so D_def and D_pow overlap each other (and, for the training
        # steps, the generator's backward pass); inside the CUDA graph these are parallel branches.
        self.d_streams = os.environ.get("SRB_D_STREAMS", "1") == "1"
        self._k_streams = None
        self.marks = None            # tools/phase_breakdown.py: list that receives (phase name, CUDA event) on the main stream
        # bucketed gradient all-reduce overlapped with the generator's backward (srb200/dp.py)
        self._g_sync = DP.BucketedAllReduce(self.opt_G.flat_g, process_group) if self.world > 1 else None
        self._g_off = {n: o for (n, _), o in zip(((n, p) for n, p in generator.named_parameters() if p.requires_grad),
                                                 self.opt_G.offsets)}
        # tensor-core discriminator engine where the configuration allows it; the general fp32 autograd path otherwise
        def engine_for(d):
            ok = (use_disc_engine and d is not None and type(d).__name__ == "Markovian_Discriminator"
                  and DE.supported(d.model, d.input_shape))
            return DE.DiscEngine(d.model, d.input_shape) if ok else None
        self.d_eng = [engine_for(d) for d in self.Ds]
        self.d2_eng = [engine_for(d) for d in self.D2s]

    def _mark(self, name: str) -> None:
        if self.marks is not None:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self.marks.append((name, e))

    def _members(self, k):
        """(discriminator, engine, optimizer, key suffix) of branch k: the discriminator and, if configured, its
        SecondDiscriminators twin."""
        m = [(self.Ds[k], self.d_eng[k], self.opt_D[k], "")]
        if self.D2s[k] is not None:
            m.append((self.D2s[k], self.d2_eng[k], self.opt_D2[k], "2"))
        return m

    def optimizers(self):
        return [self.opt_G] + [o for o in self.opt_D + self.opt_D2 if o is not None]

    def end_of_batch(self) -> None:
        """esrgan.py:789-795: re-initialise the second discriminators every second_discr_reset_interval iterations
        (host-side, outside any CUDA graph; GraphedStep calls it after each replay)."""
        self.batches_done += 1
        if self.reset_interval > 0 and (self.batches_done + 1) % self.reset_interval == 0:
            from .generator import PARAM_EPOCH
            with torch.no_grad():
                for d2 in self.D2s:
                    if d2 is None:
                        continue
                    for m in d2.modules():
                        if isinstance(m, torch.nn.Conv2d) and self.uniform_reset:       # models.py:380-384
                            torch.nn.init.xavier_uniform_(m.weight)
                            torch.nn.init.constant_(m.bias, 0.)
                        elif isinstance(m, (torch.nn.Conv2d, torch.nn.Linear)) and not self.uniform_reset:
                            m.reset_parameters()                                        # models.py:375-378
            PARAM_EPOCH[0] += 1

    # ------------------------------------------------------------------ helpers
    def _pix_work(self, B, H, W, dev):
        key = (B, H, W)
        w = self._work.get(key)
        if w is None:
            w = dict(w0=LS.PixelLossWork(B, H, W, self.factor, dev), w1=LS.PixelLossWork(B, H, W, self.factor, dev),
                     g0=torch.empty(B, 1, H, W, device=dev), g1=torch.empty(B, 1, H, W, device=dev),
                     opt=LS.OptionalLossWork(B, H, W, self.factor, dev, self.binedges),
                     sink=LS.SinkhornWork(B, H * W, dev) if self.l_wasser > 0 else None)
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
        self._mark("start")
        out, srs = eng.forward(imgs_lr, power=G._host_scalar("power"), multiplier=G._host_scalar("multiplier"),
                               thres=float(G.thres), training=G.training, save_for_backward=True)
        G.srs = srs
        self._mark("generator forward")
        generated = (out, srs)
        g = (wk["g0"], wk["g1"])
        p = self.power
        truth = (imgs_hr, imgs_hr if p == 1 else imgs_hr.pow(p))                           # esrgan.py:470
        # pooled-LR term always differentiates generated[0] (esrgan.py:471-472), for both k
        if p == 1:
            w_lr0 = (lam0 if lam0 > 0 else 0.0) * self.l_lr + (lam1 if lam1 > 0 else 0.0) * self.l_lr
            l_hr0, l_lr0 = LS.pixel_losses(out, imgs_hr, imgs_lr, self.factor, max(lam0, 0.0) * self.l_hr, w_lr0, g[0],
                                           False, wk["w0"])
            l_lr1 = l_lr0
        else:
            l_hr0, l_lr0 = LS.pixel_losses(out, imgs_hr, imgs_lr, self.factor, max(lam0, 0.0) * self.l_hr,
                                           max(lam0, 0.0) * self.l_lr, g[0], False, wk["w0"])
            # generated_lr[1] = SumPool(out) ** p against lr ** p (472-473,490): its gradient also lands on `out`
            _, l_lr1 = LS.pixel_losses(out, imgs_hr, imgs_lr, self.factor, 0.0, max(lam1, 0.0) * self.l_lr, g[0], True,
                                       wk["w1"], lr_power=p)
        l_hr1, _ = LS.pixel_losses(srs, truth[1], imgs_lr, self.factor, max(lam1, 0.0) * self.l_hr, 0.0, g[1], False,
                                   wk["w1"])
        out_losses.update(hr0=l_hr0, lr0=l_lr0, hr1=l_hr1, lr1=l_lr1)
        loss_G = lam0 * (self.l_hr * l_hr0 + self.l_lr * l_lr0) + lam1 * (self.l_hr * l_hr1 + self.l_lr * l_lr1)
        # optional terms (esrgan.py:530-550), each fused with its gradient w.r.t. generated[k]
        for k in range(2):
            lam = self.lambdas[k]
            if lam <= 0:
                continue
            for name, coef, fn in (
                    ("nnz", self.l_nnz, lambda c: LS.nnz_loss(generated[k], truth[k], c, g[k], wk["opt"])),
                    ("mask", self.l_mask, lambda c: LS.mask_loss(generated[k], truth[k], c, g[k])),
                    ("hist", self.l_hist, lambda c: LS.hist_loss(generated[k], truth[k], k, self.sigma, c, g[k], wk["opt"])),
                    ("wasser", self.l_wasser, lambda c: LS.wasserstein_loss(generated[k], truth[k], self.sinkhorn_eps, c,
                                                                            g[k], wk["sink"])),
                    ("hit", self.l_hit, lambda c: LS.hit_loss(generated[k], truth[k], self.factor, self.hit_threshold,
                                                              self.sigma, c, g[k], wk["opt"]))):
                if coef > 0:
                    l_ = fn(lam * coef)
                    out_losses[f"{name}{k}"] = l_
                    loss_G = loss_G + lam * coef * l_
        main0 = torch.cuda.current_stream()
        active = [k for k in range(2) if not (self.lambdas[k] <= 0 or self.Ds[k] is None or self.l_adv <= 0)]
        use_streams = self.d_streams and len(active) == 2 and \
            all(e is not None for k in active for (_, e, _, _) in self._members(k))
        if use_streams:
            if self._k_streams is None:
                self._k_streams = [torch.cuda.Stream(), torch.cuda.Stream()]
            fork0 = torch.cuda.Event()
            fork0.record(main0)                 # generator forward and pixel-loss gradients are complete here
        adv_done = []
        for k in active:
            members = self._members(k)
            w_adv = self.lambdas[k] * self.l_adv / len(members)          # esrgan.py:498-502: mean over the ensemble
            ctx = torch.cuda.stream(self._k_streams[k]) if use_streams else _NullCtx()
            if use_streams:
                self._k_streams[k].wait_event(fork0)
            l_adv = None
            with ctx:
                for d, eng_d, _, sfx in members:
                    if eng_d is not None:
                        A = eng_d.forward([truth[k], generated[k]], tag="rf")              # esrgan.py:493-494 (one 2B batch)
                        l_m, gu, _ = LS.ragan_raw(A.pred[B:], A.pred[:B])                  # esrgan.py:507-508
                        g_adv, _ = eng_d.input_grad(A, gu * w_adv, B, B, tag="ig")
                    else:
                        with torch.no_grad():
                            pred_real = D.sequential_forward(d.model, truth[k])           # esrgan.py:493 (.detach())
                        x = generated[k].detach().requires_grad_(True)
                        with torch.enable_grad():
                            pred_fake = D.sequential_forward(d.model, x, detach_params=True)   # esrgan.py:494
                            l_m = LS.RaGAN.apply(pred_fake, pred_real)                    # esrgan.py:507-508
                            (g_adv,) = torch.autograd.grad(l_m * w_adv, x)
                        l_m = l_m.detach()
                    g[k].add_(g_adv)
                    l_adv = l_m if l_adv is None else l_adv + l_m
                l_adv = l_adv / len(members)
                if use_streams:
                    ev = torch.cuda.Event()
                    ev.record(self._k_streams[k])
                    adv_done.append(ev)
            out_losses[f"adv{k}"] = l_adv.detach()
        for ev in adv_done:
            main0.wait_event(ev)                # the generator's backward pass needs both adversarial gradients
        self._mark("losses + discriminator forward / input gradient")
        for k in active:
            loss_G = loss_G + self.lambdas[k] * self.l_adv * out_losses[f"adv{k}"]
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
        self._mark("generator backward (buckets all-reduced beside it)")
        for n, p in G.named_parameters():
            if p.requires_grad:
                p.grad = grads[n]                      # views of the optimizer's flat gradient buffer: no copy
        self.opt_G.gather_grads()
        if self._g_sync is not None:
            self._g_sync.finish()
        self._mark("wait for the last gradient bucket")
        self.opt_G.step(grad_scale=1.0 / self.world, gathered=True)                        # esrgan.py:555
        self._mark("generator Adam")
        out_losses["g_loss"] = loss_G
        if use_streams:
            # each discriminator's training step continues on ITS stream (ordered after its own forward / input
            # gradient above, independent of the generator's backward pass and of the other discriminator); the host
            # issues them here, after FusedAdam.step of the generator, the streams join at the end of the iteration
            for k in active:
                with torch.cuda.stream(self._k_streams[k]):
                    self._discriminator_steps(truth, generated, epsilons, out_losses, B, dev, only=k)
                join = torch.cuda.Event()
                join.record(self._k_streams[k])
                main.wait_event(join)
            for k in range(2):
                if k not in active:
                    self._discriminator_steps(truth, generated, epsilons, out_losses, B, dev, only=k)
        elif self.overlap_d:
            # issued after the generator's backward on the host (FusedAdam.step bumps the global parameter epoch the
            # generator engine checks), but ordered after `fork` only: inside the CUDA graph they are a parallel branch
            self._d_stream.wait_event(fork)
            with torch.cuda.stream(self._d_stream):
                self._discriminator_steps(truth, generated, epsilons, out_losses, B, dev)
            join = torch.cuda.Event()
            join.record(self._d_stream)
            main.wait_event(join)
        else:
            self._discriminator_steps(truth, generated, epsilons, out_losses, B, dev)
        self._mark("discriminator steps (joined)")
        if not torch.cuda.is_current_stream_capturing():
            self.end_of_batch()
        return out_losses

    def _discriminator_steps(self, truth, generated, epsilons, out_losses, B, dev, only=None) -> None:
        """esrgan.py:561-626 for every active discriminator (or branch ``only``) on the current stream."""
        for k in range(2):
            if self.lambdas[k] <= 0 or self.Ds[k] is None or (only is not None and k != only):
                continue
            real, fake = truth[k], generated[k].detach()
            eps = None
            if self.l_reg > 0:                                                             # esrgan.py:598: one draw per
                eps = epsilons[k].to(dev) if epsilons is not None else torch.rand(B, 1, 1, 1, device=dev)   # branch
            gate = None
            for d, eng_d, opt, sfx in self._members(k):
                opt.zero_grad()
                if eng_d is not None:
                    # the weights have not moved since the generator step: the forward over (real, fake) is reused
                    A = eng_d.acts(2 * B, dev, "rf") if (self.lambdas[k] > 0 and self.l_adv > 0) else \
                        eng_d.forward([real, fake], tag="rf")
                    loss_D, gu, gv = LS.ragan_raw(A.pred[:B], A.pred[B:])                  # esrgan.py:578-588
                    opt.flat_g.zero_()
                    grads = [(opt.grad_views[2 * l], opt.grad_views[2 * l + 1]) for l in range(eng_d.L)]
                    eng_d.backward(A, torch.cat([gu, gv]), grads, tag="bw")
                    if self.l_reg > 0:
                        interp = eps * real + (1 - eps) * fake                             # 599-600
                        Ai = eng_d.forward([interp], tag="gp")                             # 601
                        gi, hz = eng_d.input_grad(Ai, torch.ones_like(Ai.pred), 0, B, tag="gph")   # 602
                        gp, q0 = LS.gradient_penalty_raw(gi.reshape(B, -1), self.l_reg / 2)    # 604-605
                        eng_d.penalty_backward(Ai, hz, q0.view_as(gi), grads)
                        loss_D = loss_D + gp                                               # 606
                        out_losses[f"gp{sfx}{k}"] = gp
                    for p_, gv_ in zip(opt.params, opt.grad_views):
                        p_.grad = gv_
                else:
                    pred_real = D.sequential_forward(d.model, real)                        # esrgan.py:569
                    pred_fake = D.sequential_forward(d.model, fake)                        # esrgan.py:570
                    loss_D = LS.RaGAN.apply(pred_real, pred_fake)                          # esrgan.py:578-588
                    if self.l_reg > 0:
                        interp = eps * real + (1 - eps) * fake                             # 599-600
                        _, gi = D.input_gradient(d.model, interp)                          # 601-602
                        gp = LS.GradientPenalty.apply(gi.reshape(B, -1), self.l_reg / 2)   # 604-605
                        loss_D = loss_D + gp                                               # 606
                        out_losses[f"gp{sfx}{k}"] = gp.detach()
                    loss_D.backward()                                                      # esrgan.py:616
                if gate is None:         # both ensemble members step on the FIRST one's test (esrgan.py:623-626)
                    gate = loss_D.detach().clone()
                    if self.world > 1:
                        DP.sync_gate(gate, self.pg)
                opt.gather_grads()
                if self.world > 1:
                    self._allreduce(opt.flat_g)
                opt.step(gate=(gate, self.d_threshold), grad_scale=1.0 / self.world, gathered=True)   # 623-624
                out_losses[f"d{sfx}_loss{k}"] = loss_D.detach()


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
        for opt in self.tr.optimizers():
            opt.sync_lr()                        # ReduceLROnPlateau may have moved it since the last replay
        self.graph.replay()
        self.tr.end_of_batch()                   # second-discriminator resets happen between replays
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
