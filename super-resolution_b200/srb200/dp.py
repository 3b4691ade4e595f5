"""Data-parallel plumbing for the training step (absent from the reference, which is single-GPU).

Every rank holds full replicas and a shard of the batch.  The only exchange is the sum of each optimizer's flat fp32
gradient buffer (plus the 4-byte discriminator gate).  ``BucketedAllReduce`` cuts the flat buffer into buckets that
follow backward's completion order - the generator writes gradients from conv3 back to conv1, i.e. from the END of the
parameter-ordered buffer to its start - and launches one asynchronous all-reduce per bucket on a side stream as soon as
``mark_ready`` has covered it, so the transfer overlaps the rest of backward; ``finish`` joins the streams.  Works with
NCCL (CUDA tensors) and gloo (CPU tensors; used by the CPU tests)."""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Images [lo, hi) of a global batch of ``n`` that belong to ``rank`` (contiguous, sizes differ by at most one)."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def make_buckets(n: int, bucket_elems: int) -> List[Tuple[int, int]]:
    """[lo, hi) ranges covering [0, n), listed from the END of the buffer (the first gradients backward finishes)."""
    if n <= 0:
        return []
    bucket_elems = max(int(bucket_elems), 1)
    out, hi = [], n
    while hi > 0:
        lo = max(0, hi - bucket_elems)
        out.append((lo, hi))
        hi = lo
    return out


class BucketedAllReduce:
    def __init__(self, flat: torch.Tensor, group=None, bucket_bytes: Optional[int] = None):
        # 8 MB buckets: the LAST bucket (conv1 + the first RRDBs) can only start when backward is over, so its size is
        # what the step waits for (measured at N = 2 with 24 MB buckets: 0.37 ms exposed, profiles/r02_s7_phase_*);
        # SRB_DP_BUCKET_MB overrides
        if bucket_bytes is None:
            import os
            bucket_bytes = int(float(os.environ.get("SRB_DP_BUCKET_MB", "8")) * (1 << 20))
        self.flat, self.group = flat, group
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.buckets = make_buckets(flat.numel(), bucket_bytes // flat.element_size())
        self.cuda = flat.is_cuda
        self.stream = torch.cuda.Stream(device=flat.device) if (self.cuda and self.world > 1) else None
        self.reset()

    def reset(self) -> None:
        self._ready_lo = self.flat.numel()      # everything in [ready_lo, n) has its final local value
        self._next = 0                          # next bucket to launch
        self._works = []

    def mark_ready(self, lo: int) -> None:
        """Gradients at offsets >= ``lo`` are final on the current stream: launch every bucket now fully covered."""
        if self.world == 1:
            return
        self._ready_lo = min(self._ready_lo, int(lo))
        launch = []
        while self._next < len(self.buckets) and self.buckets[self._next][0] >= self._ready_lo:
            launch.append(self.buckets[self._next])
            self._next += 1
        if not launch:
            return
        if self.cuda:
            ev = torch.cuda.Event()
            ev.record()                                   # on the compute stream, after the producing kernels
            self.stream.wait_event(ev)
            with torch.cuda.stream(self.stream):
                for lo_, hi_ in launch:
                    dist.all_reduce(self.flat[lo_:hi_], group=self.group)
        else:
            for lo_, hi_ in launch:
                self._works.append(dist.all_reduce(self.flat[lo_:hi_], group=self.group, async_op=True))

    def finish(self) -> None:
        """Launch whatever is left and make the reduced buffer visible to the current stream."""
        if self.world == 1:
            return
        self.mark_ready(0)
        if self.cuda:
            torch.cuda.current_stream().wait_stream(self.stream)
        else:
            for w in self._works:
                w.wait()
        self.reset()


def sync_gate(gate: torch.Tensor, group=None) -> torch.Tensor:
    """Mean of the per-rank discriminator losses: the d_threshold decision (esrgan.py:623) must be the same on every
    rank, otherwise replicas diverge."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(gate, group=group)
        gate /= dist.get_world_size(group)
    return gate
