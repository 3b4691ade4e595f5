"""CPU oracle for the ESRGAN hot path.  TEST INFRASTRUCTURE ONLY.

This file is a functional, fp32, CPU restatement of the reference's algorithm
(lukas-blecher/super-resolution).  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import it;
the product (``super-resolution_b200/``) never does.

Parity status: the reference has no tests or golden vectors of its own
(SURVEY.md section 4), so formally "parity unpinned by the reference"; the
oracle is pinned instead against outputs of the reference itself run in the
build container (``tests/golden/make_golden.py`` imports /root/reference and
writes ``tests/golden/*.pt``; ``tests/test_oracle_golden.py`` checks this file
against them bit-for-bit-ish, tolerance 1e-6).

Every function cites the reference file:line it restates.  The arithmetic is
torch CPU fp32 (the reference's own numeric backend: ATen/oneDNN), written as
explicit functional ops over a ``state_dict`` - no nn.Module from the
reference is reused.  ``conv3x3_numpy`` is an independent numpy restatement of
the convolution used to pin ``F.conv2d`` itself on small cases.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

Tensor = torch.Tensor
SD = Dict[str, Tensor]

G_LRELU_SLOPE = 0.01   # nn.LeakyReLU() default, models.py:21,88,98
D_LRELU_SLOPE = 0.2    # nn.LeakyReLU(0.2), models.py:143,145
DRB_RES_SCALE = 0.2    # DenseResidualBlock default; RRDB ctor never forwards res_scale (models.py:49)


# ----------------------------------------------------------------------------
# independent numpy convolution (pins F.conv2d on small cases)
# ----------------------------------------------------------------------------
def conv3x3_numpy(x: np.ndarray, w: np.ndarray, b: Optional[np.ndarray], stride: int = 1) -> np.ndarray:
    """3x3, padding 1 cross-correlation, NCHW / OIHW, float64 accumulate.
    Restates what nn.Conv2d(k=3, padding=1) computes (models.py:19,63,67,87,97,99,142,144)."""
    B, C, H, W = x.shape
    O = w.shape[0]
    xp = np.zeros((B, C, H + 2, W + 2), dtype=np.float64)
    xp[:, :, 1:-1, 1:-1] = x
    Ho = (H + 2 - 3) // stride + 1
    Wo = (W + 2 - 3) // stride + 1
    out = np.zeros((B, O, Ho, Wo), dtype=np.float64)
    for ky in range(3):
        for kx in range(3):
            patch = xp[:, :, ky:ky + stride * (Ho - 1) + 1:stride, kx:kx + stride * (Wo - 1) + 1:stride]
            out += np.einsum("bchw,oc->bohw", patch, w[:, :, ky, kx].astype(np.float64))
    if b is not None:
        out += b.astype(np.float64)[None, :, None, None]
    return out


# ----------------------------------------------------------------------------
# generator (models.py:9-135)
# ----------------------------------------------------------------------------
class DropMasks:
    """Dropout2d channel masks of the dense blocks (models.py:29-31,39-40), in forward order.  nn.Dropout2d draws
    ``empty(B, C, 1, 1).bernoulli_(1 - p) / (1 - p)`` from the default generator per call; drawing the same way in the
    same order under the same seed reproduces the reference's masks on the CPU.  ``masks`` keeps what was drawn (or is
    preset) so the device path can be fed exactly the same factors."""

    def __init__(self, drop_rate: float, masks: Optional[List[Tensor]] = None):
        self.p = float(drop_rate)
        self.masks: List[Tensor] = [] if masks is None else list(masks)
        self.preset = masks is not None
        self.i = 0

    def next(self, like: Tensor) -> Tensor:
        if self.preset:
            m = self.masks[self.i]
        else:
            m = torch.empty(like.shape[0], like.shape[1], 1, 1).bernoulli_(1 - self.p).div_(1 - self.p)
            self.masks.append(m)
        self.i += 1
        return m


def dense_residual_block(sd: SD, prefix: str, x: Tensor, drop: Optional[DropMasks] = None) -> Tensor:
    """models.py:34-41.  conv k sees concat(x, out_1..out_{k-1}); LeakyReLU(0.01)
    after convs 1-4; Dropout2d on out5 in training when drop_rate > 0; returns out5 * 0.2 + x."""
    inputs = x
    out = x
    for k in range(1, 6):
        out = F.conv2d(inputs, sd[f"{prefix}b{k}.0.weight"], sd[f"{prefix}b{k}.0.bias"], 1, 1)
        if k < 5:
            out = F.leaky_relu(out, G_LRELU_SLOPE)
        inputs = torch.cat([inputs, out], 1)
    if drop is not None:
        out = out * drop.next(out)                                           # models.py:39-40
    return out * DRB_RES_SCALE + x


def rrdb(sd: SD, prefix: str, x: Tensor, res_scale: float, drop: Optional[DropMasks] = None) -> Tensor:
    """models.py:44-53: three dense blocks, then * res_scale + x."""
    out = x
    for j in range(3):
        out = dense_residual_block(sd, f"{prefix}dense_blocks.{j}.", out, drop)
    return out * res_scale + x


def _count(sd: SD, pattern: str) -> int:
    n = 0
    while any(k.startswith(pattern % n) for k in sd):
        n += 1
    return n


def generator_forward(sd: SD, x: Tensor, *, res_scale: float = 0.2, thres: float = 0.0,
                      training: bool = True, drop: Optional[DropMasks] = None) -> Tuple[Tensor, Tensor]:
    """GeneratorRRDB.forward, models.py:120-135, every constructor variant: upsampling stages are recognised by their
    weight shape ([4F,F,3,3]: conv + LeakyReLU + PixelShuffle, models.py:85-91; [F,F,2,2]: ConvTranspose2d(k=2, s=2) +
    LeakyReLU, models.py:70-83), ``res_blocks_final.*`` entries are the num_final_layer_res blocks (models.py:93-94,
    128-130), ``drop`` carries the Dropout2d masks (training only).  Returns (output, srs)."""
    power = float(sd["power"].item())
    multiplier = float(sd["multiplier"].item())
    nb = _count(sd, "res_blocks.%d.")
    nfinal = _count(sd, "res_blocks_final.%d.")
    up_idx = sorted({int(k.split(".")[1]) for k in sd if k.startswith("upsampling.") and k.endswith("weight")})
    if not training:
        drop = None

    x = multiplier * (x ** power)                                            # models.py:122
    out1 = F.conv2d(x, sd["conv1.weight"], sd["conv1.bias"], 1, 1)           # 123
    out = out1
    for i in range(nb):                                                      # 124
        out = rrdb(sd, f"res_blocks.{i}.", out, res_scale, drop)
    out2 = F.conv2d(out, sd["conv2.weight"], sd["conv2.bias"], 1, 1)         # 125
    out = out1 + out2                                                        # 126
    for i in up_idx:                                                         # 127 / 70-91
        w = sd[f"upsampling.{i}.weight"]
        if w.shape[-1] == 2:
            out = F.conv_transpose2d(out, w, sd[f"upsampling.{i}.bias"], stride=2)
            out = F.leaky_relu(out, G_LRELU_SLOPE)
        else:
            out = F.conv2d(out, w, sd[f"upsampling.{i}.bias"], 1, 1)
            out = F.leaky_relu(out, G_LRELU_SLOPE)
            out = F.pixel_shuffle(out, 2)
    if nfinal:                                                               # 128-130
        out3 = out
        for i in range(nfinal):
            out3 = rrdb(sd, f"res_blocks_final.{i}.", out3, res_scale, drop)
        out = out3 + out
    out = F.conv2d(out, sd["conv3.0.weight"], sd["conv3.0.bias"], 1, 1)      # 131 / 96-100
    out = F.leaky_relu(out, G_LRELU_SLOPE)
    out = F.conv2d(out, sd["conv3.2.weight"], sd["conv3.2.bias"], 1, 1) / multiplier

    def head(t: Tensor, p: float) -> Tensor:                                 # models.py:114-118
        lambd = float(thres) ** p
        if training:
            return F.hardshrink(t, lambd)
        return F.hardshrink(F.relu(t), lambd)

    srs = head(out, power)                                                   # 132
    if power != 1:                                                           # 133-134
        out = F.relu(out) ** (1.0 / power)
    return head(out, 1.0), srs                                               # 135


# ----------------------------------------------------------------------------
# discriminator (models.py:140-174) and SumPool2d (models.py:297-305)
# ----------------------------------------------------------------------------
def discriminator_forward(sd: SD, img: Tensor) -> Tensor:
    """Markovian_Discriminator.forward, models.py:149-174: per channel entry
    [conv s1, LReLU .2, conv s2, LReLU .2], then conv -> 1."""
    idx = sorted({int(k.split(".")[1]) for k in sd if k.startswith("model.") and k.endswith("weight")})
    out = img
    for n, i in enumerate(idx):
        last = n == len(idx) - 1
        stride = 1 if (last or n % 2 == 0) else 2
        out = F.conv2d(out, sd[f"model.{i}.weight"], sd[f"model.{i}.bias"], stride, 1)
        if not last:
            out = F.leaky_relu(out, D_LRELU_SLOPE)
    return out


def sum_pool(x: Tensor, k: int) -> Tensor:
    """SumPool2d.forward, models.py:304-305: k*k*AvgPool2d(k)."""
    return (k * k) * F.avg_pool2d(x, k)


# ----------------------------------------------------------------------------
# data path: constituent lists -> dense images
# ----------------------------------------------------------------------------
def extract_numpy(data: np.ndarray, etaBins: int, phiBins: int) -> np.ndarray:
    """``extract``, datasets.py:136-145: data = [2, N] (row 0 flat index, row 1 energy).  fp32 adds in list order into
    img[0, idx % etaBins, idx // etaBins]; the loop ends after the first zero-energy entry (which adds nothing)."""
    img = np.zeros((1, etaBins, phiBins), dtype=np.float32)
    for i in range(data.shape[1]):
        pos = float(data[0, i])
        phi = int(pos // etaBins)
        eta = int(pos % etaBins)
        img[0, eta, phi] = np.float32(img[0, eta, phi] + np.float32(data[1, i]))
        if data[1, i] == 0:
            break
    return img


def extract_batch(pairs: Tensor, etaBins: int, phiBins: int, factor: int) -> Tuple[Tensor, Tensor]:
    """A batch of SparseJetDataset.__getitem__ (datasets.py:238-249) without cutter / noise: pairs [B, N, 2] ->
    (hr [B,1,etaBins,phiBins], lr = SumPool2d(factor)(hr))."""
    hr = torch.stack([torch.from_numpy(extract_numpy(p.t().numpy(), etaBins, phiBins)) for p in pairs.float()])
    return hr, sum_pool(hr, factor)


def synthetic_constituents(batch: int, n_max: int, etaBins: int, phiBins: int, seed: int, dup_every: int = 3) -> Tensor:
    """Zero-padded (flat index, energy) lists [batch, n_max, 2]: 5..n_max hits ~ Exp(5); every ``dup_every``-th image
    repeats some cells (the += of datasets.py:142 matters there) and image 1 uses the whole list (no terminator)."""
    g = torch.Generator().manual_seed(seed)
    out = torch.zeros(batch, n_max, 2)
    for b in range(batch):
        n = n_max if b == 1 else int(torch.randint(5, n_max, (1,), generator=g))
        idx = torch.randint(0, etaBins * phiBins, (n,), generator=g).float()
        if b % dup_every == 0 and n > 8:
            idx[n // 2:n // 2 + 4] = idx[:4]
            idx[-1] = idx[0]
        e = -5.0 * torch.log(1 - torch.rand(n, generator=g)) + 1e-3
        out[b, :n, 0], out[b, :n, 1] = idx, e
    return out


def nth_hardest_numpy(arr: np.ndarray, n: int = 1):
    """``get_nth_hardest``, evaluation/eval.py:98-105: the n-th largest pixel of every image (flattened sort)."""
    temp = np.squeeze(arr)
    return [np.sort(temp[i], axis=None)[-n] for i in range(temp.shape[0])]


def image_stats_numpy(Ln: np.ndarray, thres: float):
    """The collector modes of evaluation/eval.py:323-335 on one batch Ln [B,1,H,W]: max, min, mean, nnz, sum
    (lines 325-333) and meannnz (334-339: mean of the pixels > thres, 0 if there is none)."""
    meannnz = []
    for j in range(len(Ln)):
        nnz = Ln[j][Ln[j] > thres]
        # the reference passes copy=0 positionally (np.nan_to_num(nnz.mean(), 0)), which numpy >= 2 rejects for scalars;
        # the value is the same: NaN -> 0
        meannnz.append(0 if len(nnz) == 0 else np.nan_to_num(nnz.mean()))
    return {"max": Ln.max((1, 2, 3)), "min": Ln.min((1, 2, 3)), "mean": Ln.mean((1, 2, 3)),
            "nnz": (Ln > thres).sum((1, 2, 3)), "sum": Ln.sum((1, 2, 3)), "meannnz": np.asarray(meannnz)}


def sum_raster_numpy(imgs: np.ndarray, factor: int, threshold: float) -> np.ndarray:
    """What one SumRaster.add call adds for one image stack (utils.py:290-293): the number of pixels above the threshold
    at every position inside the factor x factor cells, summed over cells, channels and images."""
    B, C, H, W = imgs.shape
    cells = imgs.reshape(B, C, H // factor, factor, W // factor, factor)
    return (cells > threshold).sum((0, 1, 2, 4)).astype(np.float64)


def mean_image_numpy(imgs: np.ndarray, threshold: float, energy: bool) -> np.ndarray:
    """What one MeanImage.add call adds for one image stack [B,1,H,W] (utils.py:317-322)."""
    if energy or threshold <= 0:
        return (np.clip(imgs, threshold, None).sum((0, 1)) * (imgs > threshold).sum((0, 1))).astype(np.float64)
    return (imgs > threshold).sum((0, 1)).astype(np.float64)


def l1_metrics_numpy(gen_hr: Tensor, hr: Tensor, lr: Tensor, factor: int) -> Tuple[np.ndarray, np.ndarray]:
    """calculate_metrics, evaluation/eval.py:477-485: per-image (lr L1, hr L1)."""
    gen_lr = sum_pool(gen_hr, factor)
    return ((gen_lr - lr).abs().numpy().mean((1, 2, 3)), (gen_hr - hr).abs().numpy().mean((1, 2, 3)))


# ----------------------------------------------------------------------------
# losses (esrgan.py:457-626), default-active terms only
# ----------------------------------------------------------------------------
EPS = 1e-7  # esrgan.py:319


def bce_logits(x: Tensor, target_value: float) -> Tensor:
    """nn.BCEWithLogitsLoss (mean) against a constant target, esrgan.py:244."""
    t = torch.full_like(x, target_value)
    return F.binary_cross_entropy_with_logits(x, t)


def ragan_loss_g(pred_real: Tensor, pred_fake: Tensor) -> Tensor:
    """esrgan.py:507-508."""
    return 0.5 * (bce_logits(EPS + pred_fake - pred_real.mean(0, keepdim=True), 1.0)
                  + bce_logits(EPS + pred_real - pred_fake.mean(0, keepdim=True), 0.0))


def ragan_loss_d(pred_real: Tensor, pred_fake: Tensor) -> Tensor:
    """esrgan.py:578-579,588."""
    loss_real = bce_logits(EPS + pred_real - pred_fake.mean(0, keepdim=True), 1.0)
    loss_fake = bce_logits(EPS + pred_fake - pred_real.mean(0, keepdim=True), 0.0)
    return (loss_real + loss_fake) / 2


def softgreater(x: Tensor, val: float, sigma: float = 5000, delta: float = 0) -> Tensor:
    """utils.py:259-261."""
    return torch.sigmoid(sigma * (x - val + delta))


def nnz_mask(x: Tensor, sigma: float = 5e4) -> Tensor:
    """utils.py:271-272."""
    return torch.sigmoid(sigma * x)


def get_hitogram(t: Tensor, factor: int, threshold: float = .1, sig: float = 80) -> Tensor:
    """utils.py:264-268 for sig > 0: mean soft hit indicator per position inside the factor x factor cells."""
    B, C, H, W = t.shape
    cells = t.reshape(B, C, H // factor, factor, W // factor, factor)
    return torch.sigmoid(sig * (cells - threshold)).mean((0, 1, 2, 4))


def diffable_histogram(x: Tensor, binedges: Tensor, sigma: float) -> Tensor:
    """DiffableHistogram.forward (models.py:308-342) built from bin edges, not batchwise: [1, bins]."""
    delta = binedges[1:] - binedges[:-1]
    centers = binedges[:-1] + .5 * delta
    d = x.reshape(1, 1, -1) - centers.reshape(1, -1, 1)
    half = delta.reshape(1, -1, 1) / 2
    return (torch.sigmoid(sigma * (d + half)) - torch.sigmoid(sigma * (d - half))).sum(2)


def kld_hist(q_entries: Tensor, p_entries: Tensor, binedges: Tensor) -> Tensor:
    """KLD_hist.forward, utils.py:103-113 (q = generated histogram, p = ground-truth histogram)."""
    binsizes = (binedges[1:] - binedges[:-1]).float()
    n_p, n_q = p_entries.sum().float(), q_entries.sum().float()
    p = p_entries * binsizes / n_p
    q = ((q_entries + 1e-6) * binsizes / n_q).log()
    return F.kl_div(q, p, reduction="sum") / binsizes.mean()


def cut_smaller(t: Tensor, threshold: float = 0.0) -> Tensor:
    """utils.py:241-255: of a row-wise ascending (B, N) tensor keep the last n columns, n = the largest per-row count
    of entries above ``threshold`` (rows with fewer such entries keep some smaller ones; n == 0 keeps everything, as
    the reference's ``t[:, -0:]`` does)."""
    n = int((t > threshold).sum(1).max())
    return t[:, -n:]


def sinkhorn_distance(x: Tensor, y: Tensor, eps: float, max_iter: int = 100, thresh: float = 0.1
                      ) -> Tuple[Tensor, int]:
    """SinkhornDistance(eps, max_iter, 'sum').forward (utils.py:176-219) for point clouds x (B, P1, 1), y (B, P2, 1)
    with uniform marginals: cost C = |x_i - y_j|^2 (228-234), log-domain updates of the duals u, v (198-199) until
    the batch mean of sum_i |u - u_prev| drops below 0.1 (200-204), then sum_b sum_ij pi_ij C_ij with
    pi = exp((-C + u_i + v_j) / eps) (206-216).  Returns (cost, iterations run)."""
    C = ((x.unsqueeze(-2) - y.unsqueeze(-3)).abs() ** 2).sum(-1)
    B, P1, P2 = C.shape
    log_mu = torch.log(torch.full((B, P1), 1.0 / P1) + 1e-8)
    log_nu = torch.log(torch.full((B, P2), 1.0 / P2) + 1e-8)
    u, v = torch.zeros(B, P1), torch.zeros(B, P2)

    def m(u_, v_):                                                    # utils.py:221-224
        return (-C + u_.unsqueeze(-1) + v_.unsqueeze(-2)) / eps

    nits = 0
    for _ in range(max_iter):
        u_prev = u
        u = eps * (log_mu - torch.logsumexp(m(u, v), -1)) + u
        v = eps * (log_nu - torch.logsumexp(m(u, v).transpose(-2, -1), -1)) + v
        nits += 1
        if (u - u_prev).abs().sum(-1).mean().item() < thresh:
            break
    pi = torch.exp(m(u, v))
    return (pi * C).sum((-2, -1)).sum(), nits


def wasserstein_loss(gen: Tensor, truth: Tensor, eps: float) -> Tensor:
    """esrgan.py:539-542: Sinkhorn distance between the per-image sorted pixel values of the generated and the true
    image, each cut to the largest non-zero count of its batch."""
    B = gen.shape[0]
    gen_sort = torch.sort(gen.reshape(B, -1), 1)[0]
    real_sort = torch.sort(truth.reshape(B, -1), 1)[0]
    return sinkhorn_distance(cut_smaller(gen_sort)[..., None], cut_smaller(real_sort)[..., None], eps)[0]


def generator_loss(gen: Sequence[Tensor], hr: Tensor, lr: Tensor, d_sds: Sequence[SD], *, factor: int,
                   scaling_power: float = 1.0, lambdas=(0.2, 1.0), lambda_hr=1.0, lambda_lr=0.1,
                   lambda_adv=0.01, lambda_nnz=0.0, lambda_mask=0.0, lambda_hist=0.0, lambda_hit=0.0,
                   binedges: Optional[Sequence[Tensor]] = None, sigma: float = 500, hit_threshold: float = 0.5,
                   lambda_wasser: float = 0.0, sinkhorn_eps: float = 0.1) -> Tuple[Tensor, Dict[str, Tensor]]:
    """G-step loss assembly, esrgan.py:468-552 with the default.json weights (+ the optional nnz / mask / hist / hit
    / wasser terms of 530-550 when their lambda is > 0).  ``gen`` = (generator output, generator.srs)."""
    ground_truth = (hr, hr ** scaling_power)                                  # 470
    gen_lr = sum_pool(gen[0], factor)                                         # 471
    generated_lr = (gen_lr, gen_lr ** scaling_power)                          # 472
    ground_truth_lr = (lr, lr ** scaling_power)                               # 473
    parts: Dict[str, Tensor] = {}
    loss_g = torch.zeros(1)
    for k in range(2):
        if lambdas[k] <= 0:
            continue
        l_hr = F.l1_loss(gen[k].mean(0)[None], ground_truth[k].mean(0)[None])  # 487
        l_lr = F.l1_loss(generated_lr[k], ground_truth_lr[k])                   # 490
        pred_real = discriminator_forward(d_sds[k], ground_truth[k]).detach()   # 493
        pred_fake = discriminator_forward(d_sds[k], gen[k])                     # 494
        l_adv = ragan_loss_g(pred_real, pred_fake)                              # 507
        tot = lambda_hr * l_hr + lambda_adv * l_adv + lambda_lr * l_lr          # 549
        if lambda_nnz > 0:                                                      # 530-533
            gen_nnz = softgreater(gen[k], 0, 50000).sum(1).sum(1).sum(1)
            target = (ground_truth[k] > 0).sum(1).sum(1).sum(1).float()
            parts[f"nnz{k}"] = F.mse_loss(gen_nnz, target)
            tot = tot + lambda_nnz * parts[f"nnz{k}"]
        if lambda_mask > 0:                                                     # 534-537
            parts[f"mask{k}"] = F.l1_loss(nnz_mask(gen[k]), nnz_mask(ground_truth[k]))
            tot = tot + lambda_mask * parts[f"mask{k}"]
        if lambda_hist > 0:                                                     # 538-545
            gen_hist = diffable_histogram(gen[k][gen[k] > 0], binedges[k], sigma)
            real_hist = diffable_histogram(ground_truth[k][ground_truth[k] > 0], binedges[k], sigma)
            parts[f"hist{k}"] = kld_hist(gen_hist, real_hist, binedges[k])
            tot = tot + lambda_hist * parts[f"hist{k}"]
        if lambda_wasser > 0:                                                   # 539-542
            parts[f"wasser{k}"] = wasserstein_loss(gen[k], ground_truth[k], sinkhorn_eps)
            tot = tot + lambda_wasser * parts[f"wasser{k}"]
        if lambda_hit > 0:                                                      # 546-550
            parts[f"hit{k}"] = F.mse_loss(get_hitogram(gen[k], factor, hit_threshold, sigma),
                                          get_hitogram(ground_truth[k], factor, hit_threshold, sigma))
            tot = tot + lambda_hit * parts[f"hit{k}"]
        loss_g = loss_g + lambdas[k] * tot                                      # 552
        parts[f"hr{k}"], parts[f"lr{k}"], parts[f"adv{k}"] = l_hr, l_lr, l_adv
    return loss_g, parts


def discriminator_loss(d_sd: SD, real: Tensor, fake: Tensor, epsilon: Tensor, *, lambda_reg=0.01
                       ) -> Tuple[Tensor, Dict[str, Tensor]]:
    """D-step loss for one discriminator, esrgan.py:569-606 (relativistic, with
    gradient penalty).  ``epsilon`` is the (B,1,1,1) uniform draw of line 598."""
    pred_real = discriminator_forward(d_sd, real)                              # 569
    pred_fake = discriminator_forward(d_sd, fake.detach())                     # 570
    loss = ragan_loss_d(pred_real, pred_fake)                                  # 578-588
    parts = {"ragan": loss}
    if lambda_reg > 0:
        interp = (epsilon * real + (1 - epsilon) * fake.detach()).requires_grad_(True)   # 599-600
        pred_i = discriminator_forward(d_sd, interp)                           # 601
        grads = torch.autograd.grad(pred_i, interp, torch.ones_like(pred_i), create_graph=True)[0]  # 602
        grads = grads.view(grads.shape[0], -1)                                 # 604
        gp = ((grads.norm(2, dim=1) - 1) ** 2).mean() * lambda_reg / 2         # 605
        loss = loss + gp                                                       # 606
        parts["gp"] = gp
    return loss, parts


# ----------------------------------------------------------------------------
# Adam (torch.optim.Adam as used at esrgan.py:299,305)
# ----------------------------------------------------------------------------
def adam_step(p: Tensor, g: Tensor, m: Tensor, v: Tensor, step: int, *, lr=2e-4, b1=0.9, b2=0.999,
              eps=1e-8, weight_decay=0.0) -> None:
    """One in-place Adam update with coupled L2 decay, no amsgrad; ``step`` is
    the 1-based step count after increment."""
    if weight_decay != 0:
        g = g + weight_decay * p
    m.mul_(b1).add_(g, alpha=1 - b1)
    v.mul_(b2).addcmul_(g, g, value=1 - b2)
    bc1 = 1 - b1 ** step
    bc2 = 1 - b2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(m, denom, value=-lr / bc1)


# ----------------------------------------------------------------------------
# synthetic sparse jet images and deterministic weights
# ----------------------------------------------------------------------------
def synthetic_jets(batch: int, hr: int, factor: int, seed: int) -> Tuple[Tensor, Tensor]:
    """HR (B,1,hr,hr) with 20..60 non-zero pixels ~ Exp(mean 5), LR = SumPool(factor)(HR)
    as datasets.py:247 does."""
    g = torch.Generator().manual_seed(seed)
    img = torch.zeros(batch, 1, hr, hr)
    for b in range(batch):
        n = int(torch.randint(20, 61, (1,), generator=g))
        idx = torch.randperm(hr * hr, generator=g)[:n]
        e = -5.0 * torch.log(1 - torch.rand(n, generator=g))
        img[b, 0].view(-1)[idx] = e
    return img, sum_pool(img, factor)


def seeded_state_dict(shapes: Dict[str, Tuple[int, ...]], seed: int, gain: float = 1.0) -> SD:
    """Deterministic weights: each tensor ~ U(-a, a), a = gain/sqrt(fan_in) (the
    scale of nn.Conv2d's default init), from its own seeded generator, so the
    same dict can be rebuilt anywhere without the reference."""
    sd: SD = {}
    for i, (name, shape) in enumerate(shapes.items()):
        g = torch.Generator().manual_seed(seed * 100003 + i)
        if name in ("power", "multiplier"):
            continue
        fan_in = int(np.prod(shape[1:])) if len(shape) > 1 else None
        if fan_in is None:
            # bias: use the weight's fan_in (previous entry)
            prev = list(shapes.values())[i - 1]
            fan_in = int(np.prod(prev[1:]))
        a = gain / math.sqrt(fan_in)
        sd[name] = (torch.rand(shape, generator=g) * 2 - 1) * a
    return sd


def upsampling_kinds(num_upsample: int, use_transposed_conv=False, fully_tconv_upsample=False) -> List[str]:
    """'ps' (conv + LeakyReLU + PixelShuffle) or 'tconv' (ConvTranspose2d + LeakyReLU) per stage, models.py:70-91:
    use_transposed_conv is tested first and alternates ps / tconv."""
    if use_transposed_conv:
        return ["ps" if i % 2 == 0 else "tconv" for i in range(num_upsample)]
    if fully_tconv_upsample:
        return ["tconv"] * num_upsample
    return ["ps"] * num_upsample


def generator_shapes(channels=1, filters=64, num_res_blocks=10, num_upsample=2, use_transposed_conv=False,
                     fully_tconv_upsample=False, num_final_layer_res=0) -> Dict[str, Tuple[int, ...]]:
    """state_dict names/shapes of GeneratorRRDB in registration order (models.py:58-103;
    SURVEY.md section 5 'Checkpoint')."""
    s: Dict[str, Tuple[int, ...]] = {"power": (1,), "multiplier": (1,)}
    F_ = filters
    s["conv1.weight"] = (F_, channels, 3, 3); s["conv1.bias"] = (F_,)
    for i in range(num_res_blocks):
        for j in range(3):
            for k in range(1, 6):
                p = f"res_blocks.{i}.dense_blocks.{j}.b{k}.0."
                s[p + "weight"] = (F_, k * F_, 3, 3); s[p + "bias"] = (F_,)
    s["conv2.weight"] = (F_, F_, 3, 3); s["conv2.bias"] = (F_,)
    idx = 0
    for kind in upsampling_kinds(num_upsample, use_transposed_conv, fully_tconv_upsample):
        if kind == "ps":
            s[f"upsampling.{idx}.weight"] = (4 * F_, F_, 3, 3); s[f"upsampling.{idx}.bias"] = (4 * F_,)
            idx += 3
        else:
            s[f"upsampling.{idx}.weight"] = (F_, F_, 2, 2); s[f"upsampling.{idx}.bias"] = (F_,)
            idx += 2
    for i in range(num_final_layer_res):
        for j in range(3):
            for k in range(1, 6):
                p = f"res_blocks_final.{i}.dense_blocks.{j}.b{k}.0."
                s[p + "weight"] = (F_, k * F_, 3, 3); s[p + "bias"] = (F_,)
    s["conv3.0.weight"] = (F_, F_, 3, 3); s["conv3.0.bias"] = (F_,)
    s["conv3.2.weight"] = (channels, F_, 3, 3); s["conv3.2.bias"] = (channels,)
    return s


def discriminator_shapes(in_channels=1, channels=(16, 32, 32, 64)) -> Dict[str, Tuple[int, ...]]:
    """state_dict names/shapes of Markovian_Discriminator (models.py:149-174)."""
    s: Dict[str, Tuple[int, ...]] = {}
    cin, idx = in_channels, 0
    for c in channels:
        s[f"model.{idx}.weight"] = (c, cin, 3, 3); s[f"model.{idx}.bias"] = (c,)
        s[f"model.{idx + 2}.weight"] = (c, c, 3, 3); s[f"model.{idx + 2}.bias"] = (c,)
        cin, idx = c, idx + 4
    s[f"model.{idx}.weight"] = (1, cin, 3, 3); s[f"model.{idx}.bias"] = (1,)
    return s


def make_generator_sd(seed: int, *, channels=1, filters=64, num_res_blocks=10, num_upsample=2,
                      power=1.0, multiplier=1.0, gain=1.0, use_transposed_conv=False, fully_tconv_upsample=False,
                      num_final_layer_res=0) -> SD:
    shapes = generator_shapes(channels, filters, num_res_blocks, num_upsample, use_transposed_conv,
                              fully_tconv_upsample, num_final_layer_res)
    sd = {"power": torch.tensor([float(power)]), "multiplier": torch.tensor([float(multiplier)])}
    sd.update(seeded_state_dict(shapes, seed, gain))
    return {k: sd[k] for k in shapes}


def make_discriminator_sd(seed: int, *, in_channels=1, channels=(16, 32, 32, 64), gain=1.0) -> SD:
    return seeded_state_dict(discriminator_shapes(in_channels, channels), seed, gain)
