"""Data-parallel use of the rasterizer over camera views.

The hot path shards over VIEWS only: Gaussians are replicated on every rank, rank r renders its block of the
batch's views one after the other, the per-Gaussian gradients of those views are summed into ONE flat fp32 buffer
(3 + 3*M + 1 + 3 + 4 floats per Gaussian = 236 B at M = 16) and the buffer is all-reduced once per step
(`torch.distributed`, NCCL over NVLink/NVSwitch on GPUs, gloo in the CPU tests).  There is no collective inside a
view.  The reference has no multi-GPU path; its trainers sum image losses over a camera batch on one device
(train_stacked_transformer.py:206-222), which is what this reproduces across ranks.

Nothing here touches the rasterizer itself: `render_backward` is any callable returning the gradients of one view,
so the host logic is testable on CPU (tests/test_cpu_dp.py, world_size 2, gloo).
"""
from typing import Callable, Dict, List, Sequence

import torch
import torch.distributed as dist

LEAVES = ("means3D", "sh", "opacities", "scales", "rotations")


def shard_views(n_views: int, rank: int, world: int) -> List[int]:
    """Contiguous block of view indices of `rank` (rank r gets views r*B/G ... (r+1)*B/G - 1; when G does not divide
    B the first B % G ranks get one more)."""
    if n_views < 0 or world <= 0 or not (0 <= rank < world):
        raise ValueError("bad n_views / rank / world")
    q, r = divmod(n_views, world)
    lo = rank * q + min(rank, r)
    return list(range(lo, lo + q + (1 if rank < r else 0)))


class FlatGrads:
    """One contiguous fp32 buffer with a view per parameter, in LEAVES order."""

    def __init__(self, shapes: Dict[str, Sequence[int]], device, bucket_bytes: int = 0):
        self.shapes = {k: tuple(shapes[k]) for k in LEAVES if k in shapes}
        n = 0
        self.slices = {}
        for k, shp in self.shapes.items():
            m = 1
            for d in shp:
                m *= int(d)
            self.slices[k] = (n, n + m)
            n += m
        self.flat = torch.zeros(n, dtype=torch.float32, device=device)
        self.views = {k: self.flat[a:b].view(self.shapes[k]) for k, (a, b) in self.slices.items()}
        self.bucket_elems = bucket_bytes // 4 if bucket_bytes else 0
        self._first = True

    def zero_(self):
        self._first = True

    def add_(self, grads: Dict[str, torch.Tensor]):
        """Accumulate one view's gradients; the first view of a step overwrites (no memset pass)."""
        for k, v in self.views.items():
            g = grads[k].reshape(v.shape)
            if self._first:
                v.copy_(g)
            else:
                v.add_(g)
        self._first = False

    def all_reduce(self, group=None, async_op: bool = False):
        """SUM over ranks.  bucket_bytes > 0 splits the buffer so that early buckets start while later ones queue."""
        if self._first:
            self.flat.zero_()  # this rank rendered no view this step
            self._first = False
        if not dist.is_initialized() or dist.get_world_size(group) == 1:
            return []
        if not self.bucket_elems or self.bucket_elems >= self.flat.numel():
            w = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
            return [w] if async_op else []
        works = []
        for a in range(0, self.flat.numel(), self.bucket_elems):
            w = dist.all_reduce(self.flat[a:a + self.bucket_elems], op=dist.ReduceOp.SUM, group=group,
                                async_op=async_op)
            if async_op:
                works.append(w)
        return works


def dp_step(render_backward: Callable[[int], Dict[str, torch.Tensor]], n_views: int, flat: FlatGrads, rank: int,
            world: int, group=None) -> Dict[str, torch.Tensor]:
    """One data-parallel step over a batch of `n_views` views: render this rank's views (forward + backward each),
    sum their gradients locally, all-reduce once.  Returns the per-parameter views of the reduced buffer."""
    flat.zero_()
    for v in shard_views(n_views, rank, world):
        flat.add_(render_backward(v))
    for w in flat.all_reduce(group=group, async_op=True):
        w.wait()
    return flat.views
