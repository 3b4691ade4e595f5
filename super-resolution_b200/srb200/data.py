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


STAT_NAMES = ("max", "min", "mean", "nnz", "sum", "meannnz")


def image_stats(img: Tensor, thres: float = 0.0, nmax: int = 0):
    """Per-image statistics of a CUDA fp32 batch [B, ...] in one launch: the collector modes of
    evaluation/eval.py:323-335 plus ``get_nth_hardest`` (eval.py:98-105) for n = 1..nmax.  Returns a dict of [B]
    tensors keyed by STAT_NAMES and, if nmax > 0, "hardest" = [B, nmax] (column n-1 = n-th hardest constituent)."""
    if not img.is_cuda:
        raise RuntimeError("CUDA tensor required for image_stats: super-resolution_b200 has no CPU path")
    x = img.detach().float().contiguous()
    B = x.shape[0]
    out = torch.empty(B, 6 + nmax, device=x.device)
    L.check(L.lib().srb_image_stats(x.data_ptr(), B, x[0].numel(), float(thres), int(nmax), out.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream), "image_stats")
    res = {n: out[:, i] for i, n in enumerate(STAT_NAMES)}
    if nmax:
        res["hardest"] = out[:, 6:]
    return res
