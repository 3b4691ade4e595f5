"""Device-side data path: constituent lists -> dense HR / LR batches.

Mirrors ``extract`` (datasets.py:136-145) followed by ``SumPool2d(factor)`` (datasets.py:247-249, models.py:297-305)
for a whole batch in one kernel launch instead of a Python loop per constituent and per item."""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib as L

Tensor = torch.Tensor


def extract_batch(pairs: Tensor, etaBins: int, phiBins: int, factor: Optional[int] = None,
                  check: bool = False) -> Tuple[Tensor, Optional[Tensor]]:
    """``pairs``: CUDA fp32 [B, N, 2] of (flat index, energy), zero padded.  Returns (hr [B,1,etaBins,phiBins],
    lr [B,1,etaBins/factor,phiBins/factor] or None).  ``check=True`` synchronises and raises IndexError if an index
    fell outside the image (what the reference's ``extract`` does); otherwise such entries are dropped."""
    if not pairs.is_cuda:
        raise RuntimeError("CUDA tensor required for extract_batch: super-resolution_b200 has no CPU path")
    if pairs.dim() != 3 or pairs.shape[-1] != 2:
        raise RuntimeError("CUDA srb200: extract_batch expects [B, N, 2] (index, energy) pairs")
    pairs = pairs.detach().float().contiguous()
    B, N, _ = pairs.shape
    hr = torch.empty(B, 1, etaBins, phiBins, device=pairs.device)
    lr = torch.empty(B, 1, etaBins // factor, phiBins // factor, device=pairs.device) if factor else None
    bad = torch.zeros(1, dtype=torch.int32, device=pairs.device) if check else None
    L.check(L.lib().srb_extract_batch(pairs.data_ptr(), B, N, etaBins, phiBins, factor or 1, hr.data_ptr(),
                                      lr.data_ptr() if lr is not None else None,
                                      bad.data_ptr() if bad is not None else None,
                                      torch.cuda.current_stream().cuda_stream), "extract_batch")
    if check and int(bad.item()):
        raise IndexError(f"extract_batch: {int(bad.item())} constituent indices outside the {etaBins}x{phiBins} image")
    return hr, lr


STAT_NAMES = ("max", "min", "mean", "nnz", "sum", "meannnz", "has_nan")


def image_stats(img: Tensor, thres: float = 0.0, nmax: int = 0):
    """Per-image statistics of a CUDA fp32 batch [B, ...] in one launch: the collector modes of
    evaluation/eval.py:323-335 plus ``get_nth_hardest`` (eval.py:98-105) for n = 1..nmax (<= 32: the default
    eval_modes reach E_30).  Returns a dict of [B] tensors keyed by STAT_NAMES and, if nmax > 0, "hardest" = [B, nmax]
    (column n-1 = n-th hardest constituent).  "has_nan" is 1 for images that contain a NaN: the reference skips a batch
    with any NaN (eval.py:319-321); max / min of such an image ignore the NaN here."""
    if not img.is_cuda:
        raise RuntimeError("CUDA tensor required for image_stats: super-resolution_b200 has no CPU path")
    x = img.detach().float().contiguous()
    B = x.shape[0]
    out = torch.empty(B, len(STAT_NAMES) + nmax, device=x.device)
    L.check(L.lib().srb_image_stats(x.data_ptr(), B, x[0].numel(), float(thres), int(nmax), out.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream), "image_stats")
    res = {n: out[:, i] for i, n in enumerate(STAT_NAMES)}
    if nmax:
        res["hardest"] = out[:, len(STAT_NAMES):]
    return res


def l1_metrics(gen_hr: Tensor, imgs_hr: Tensor, imgs_lr: Tensor, factor: int) -> Tuple[Tensor, Tensor]:
    """calculate_metrics, evaluation/eval.py:477-485, for one batch on the device: per-image mean |SumPool(gen) - lr|
    and mean |gen - hr| (the two lists the reference extends per batch after a .cpu() round trip)."""
    for t in (gen_hr, imgs_hr, imgs_lr):
        if not t.is_cuda:
            raise RuntimeError("CUDA tensor required for l1_metrics: super-resolution_b200 has no CPU path")
    g, h, l_ = (t.detach().float().contiguous() for t in (gen_hr, imgs_hr, imgs_lr))
    B, Cc, H, W = g.shape
    if Cc != 1 or h.shape != g.shape or tuple(l_.shape) != (B, 1, H // factor, W // factor):
        raise RuntimeError("CUDA srb200: l1_metrics expects [B,1,H,W] images and [B,1,H/f,W/f] low-resolution images")
    out = torch.empty(B, 2, device=g.device)
    L.check(L.lib().srb_l1_metrics(g.data_ptr(), h.data_ptr(), l_.data_ptr(), B, H, W, int(factor), out.data_ptr(),
                                   torch.cuda.current_stream().cuda_stream), "l1_metrics")
    return out[:, 0], out[:, 1]


def _rasters(img: Tensor, factor: int, thr_raster: float, thr_mean: float, energy: bool, raster: Optional[Tensor],
             mean_acc: Optional[Tensor]) -> None:
    if not img.is_cuda:
        raise RuntimeError("CUDA tensor required for the evaluation rasters: super-resolution_b200 has no CPU path")
    x = img.detach().float().contiguous()
    B, Cc, H, W = x.shape
    if Cc != 1:
        raise RuntimeError("CUDA srb200: evaluation rasters expect single-channel images")
    L.check(L.lib().srb_eval_rasters(x.data_ptr(), B, H, W, int(factor), float(thr_raster), float(thr_mean), int(energy),
                                     raster.data_ptr() if raster is not None else None,
                                     mean_acc.data_ptr() if mean_acc is not None else None,
                                     torch.cuda.current_stream().cuda_stream), "eval_rasters")


class SumRaster:
    """Device mirror of utils.SumRaster (utils.py:275-296; evaluation mode 'hitogram'): counts, per position inside the
    factor x factor cells, the pixels above ``threshold`` - accumulated on the device over add() calls, one launch per
    image stack instead of two np.split pyramids on the host."""

    def __init__(self, factor, height=None, width=None, threshold=.7):
        self.factor, self.threshold = factor, threshold
        self.height, self.width = height, width if width is not None else height
        self.sr = self.hr = None

    def add(self, SR: Tensor, HR: Tensor) -> None:
        if self.sr is None:
            self.sr, self.hr = (torch.zeros(self.factor, self.factor, dtype=torch.float64, device=SR.device)
                                for _ in range(2))
        _rasters(SR, self.factor, self.threshold, 0.0, False, self.sr, None)
        _rasters(HR, self.factor, self.threshold, 0.0, False, self.hr, None)

    def get_hist(self):
        return self.sr.cpu().numpy(), self.hr.cpu().numpy()


class MeanImage:
    """Device mirror of utils.MeanImage (utils.py:299-328; evaluation mode 'meanimg') without the optional
    preprocessing: per pixel, the count of images above ``threshold`` or (energy mode) the reference's
    sum(clip) * count product, accumulated over add() calls."""

    def __init__(self, factor, height=None, threshold=.7, preprocess=False, energy=False):
        if preprocess:
            raise RuntimeError("CUDA srb200: MeanImage(preprocess=True) is host-side image registration "
                               "(utils.preprocessing) and is not part of the device path")
        self.factor, self.threshold = factor, threshold
        self.energy = energy or threshold <= 0
        self.N = 0
        self.sr = self.hr = None

    def add(self, SR: Tensor, HR: Tensor) -> None:
        if self.sr is None:
            H, W = HR.shape[-2:]
            self.sr, self.hr = (torch.zeros(H, W, dtype=torch.float64, device=SR.device) for _ in range(2))
        self.N += len(SR)
        _rasters(SR, self.factor, 0.0, self.threshold, self.energy, None, self.sr)
        _rasters(HR, self.factor, 0.0, self.threshold, self.energy, None, self.hr)

    def get_hist(self):
        sr, hr = self.sr.cpu().numpy(), self.hr.cpu().numpy()
        return (sr / self.N, hr / self.N) if self.energy else (sr, hr)
