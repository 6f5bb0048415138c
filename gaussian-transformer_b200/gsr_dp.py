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


# ---------------------------------------------------------------------------------------------------------------
# SH gradients by factor exchange.  For one view dL/dsh(i,k,c) = basis_k(dir_i) * f(i,c): a rank-1 product per
# Gaussian (csrc/sh_exchange.cuh).  Instead of all-reducing 12*M B per Gaussian (192 B of the 236 B at M = 16),
# ranks all-gather the 12 B factor of each of the batch's views and every rank rebuilds the sum locally; the
# remaining 44 B per Gaussian (means3D, opacity, scales, rotations) go through the flat all-reduce as before.
# Pays off while 12 B * n_views stays below the all-reduce's 2 * 12*M B per Gaussian, i.e. n_views <~ 2*M.
SMALL_LEAVES = ("means3D", "opacities", "scales", "rotations")


class capture_sh_factors:
    """Context manager: rasterizer backward passes inside it skip dL/dsh and record (campos, factors) per view."""

    def __enter__(self):
        import diff_gaussian_rasterization as dgr
        self._dgr, self._prev = dgr, dgr._sh_factor_sink
        dgr._sh_factor_sink = self.items = []
        return self.items

    def __exit__(self, *exc):
        self._dgr._sh_factor_sink = self._prev
        return False


def gather_sh_factors(items, per_rank: int, P: int, device, world: int, group=None):
    """items: this rank's [(campos (3,), factors (P,3))], at most `per_rank`; missing views are padded with zero
    factors (they add nothing).  Returns campos (world*per_rank, 3) and factors (world*per_rank, P, 3), identical
    on every rank."""
    if len(items) > per_rank:
        raise ValueError("more local views than per_rank")
    cam = torch.zeros(per_rank, 3, dtype=torch.float32, device=device)
    fac = torch.zeros(per_rank, P, 3, dtype=torch.float32, device=device) if len(items) < per_rank else \
        torch.empty(per_rank, P, 3, dtype=torch.float32, device=device)
    for j, (c, f) in enumerate(items):
        cam[j].copy_(c.reshape(3))
        fac[j].copy_(f.reshape(P, 3))
    if world == 1 or not dist.is_initialized():
        return cam, fac
    cam_all = torch.empty(world * per_rank, 3, dtype=torch.float32, device=device)
    fac_all = torch.empty(world * per_rank, P, 3, dtype=torch.float32, device=device)
    if dist.get_backend(group) == "nccl":
        w1 = dist.all_gather_into_tensor(fac_all, fac, group=group, async_op=True)
        w2 = dist.all_gather_into_tensor(cam_all, cam, group=group, async_op=True)
    else:
        w1 = dist.all_gather(list(fac_all.view(world, per_rank, P, 3).unbind(0)), fac, group=group, async_op=True)
        w2 = dist.all_gather(list(cam_all.view(world, per_rank, 3).unbind(0)), cam, group=group, async_op=True)
    w1.wait()
    w2.wait()
    return cam_all, fac_all


def _rebuild_cuda(means3D, campos, factors, degree, M):
    from diff_gaussian_rasterization import _C
    return _C.sh_grad_from_factors(means3D, campos, factors, degree, M)


def dp_step_factored(render_backward: Callable[[int], Dict[str, torch.Tensor]], n_views: int, flat_small: FlatGrads,
                     rank: int, world: int, means3D: torch.Tensor, degree: int, M: int, rebuild=_rebuild_cuda,
                     group=None) -> Dict[str, torch.Tensor]:
    """Like dp_step, but `render_backward(v)` returns the SMALL_LEAVES gradients plus "sh_factor": (campos,
    factors) of the view; the summed dL/dsh is rebuilt from the all-gathered factors.  Returns the reduced
    gradients of all five parameters."""
    per_rank = (n_views + world - 1) // world
    items = []
    flat_small.zero_()
    for v in shard_views(n_views, rank, world):
        g = render_backward(v)
        flat_small.add_(g)
        items.append(g["sh_factor"])
    works = flat_small.all_reduce(group=group, async_op=True)
    P = means3D.shape[0]
    cam_all, fac_all = gather_sh_factors(items, per_rank, P, means3D.device, world, group=group)
    for w in works:
        w.wait()
    out = dict(flat_small.views)
    out["sh"] = rebuild(means3D, cam_all, fac_all, degree, M)
    return out
