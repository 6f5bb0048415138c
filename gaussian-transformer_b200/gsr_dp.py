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

    @staticmethod
    def numel(shapes: Dict[str, Sequence[int]]) -> int:
        n = 0
        for k in LEAVES:
            if k in shapes:
                m = 1
                for d in shapes[k]:
                    m *= int(d)
                n += (m + 3) & ~3
        return n

    def __init__(self, shapes: Dict[str, Sequence[int]], device, bucket_bytes: int = 0, flat: torch.Tensor = None):
        """flat: caller-provided storage of FlatGrads.numel(shapes) floats (e.g. a symmetric-memory allocation)."""
        self.shapes = {k: tuple(shapes[k]) for k in LEAVES if k in shapes}
        n = 0
        self.slices = {}
        for k, shp in self.shapes.items():
            m = 1
            for d in shp:
                m *= int(d)
            self.slices[k] = (n, n + m)
            n += (m + 3) & ~3  # every view starts 16-byte aligned (the kernels' 128-bit stores); padding stays 0
        if flat is not None:
            if flat.numel() != n or flat.dtype != torch.float32 or not flat.is_contiguous():
                raise ValueError("flat must be a contiguous float32 tensor of %d elements" % n)
            flat.zero_()
        self.flat = flat if flat is not None else torch.zeros(n, dtype=torch.float32, device=device)
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
                if g.data_ptr() != v.data_ptr():  # grads_into(): the backward already wrote it in place
                    v.copy_(g)
            else:
                v.add_(g)
        self._first = False

    def all_reduce(self, group=None, async_op: bool = False, world=None):
        """SUM over ranks.  bucket_bytes > 0 splits the buffer so that early buckets start while later ones queue.
        world = 1: no collective even inside an initialised process group (a single-rank reference run)."""
        if self._first:
            self.flat.zero_()  # this rank rendered no view this step
            self._first = False
        if world == 1 or not dist.is_initialized() or dist.get_world_size(group) == 1:
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
    for w in flat.all_reduce(group=group, async_op=True, world=world):
        w.wait()
    return flat.views


_C_NAMES = {"means3D": "means3D", "sh": "sh", "opacities": "opacity", "scales": "scales", "rotations": "rot"}


class grads_into:
    """Context manager: the first rasterizer backward inside writes its parameter gradients straight into the
    views of `flat` (only while `flat` is still empty for this step), so that FlatGrads.add_ has nothing to copy."""

    def __init__(self, flat: FlatGrads):
        self.flat = flat

    def __enter__(self):
        import diff_gaussian_rasterization as dgr
        self._dgr, self._prev = dgr, dgr._grad_out
        dgr._grad_out = {_C_NAMES[k]: v for k, v in self.flat.views.items()} if self.flat._first else None
        return self.flat

    def __exit__(self, *exc):
        self._dgr._grad_out = self._prev
        return False


# ---------------------------------------------------------------------------------------------------------------
# SH gradients by factor exchange.  For one view dL/dsh(i,k,c) = basis_k(dir_i) * f(i,c): a rank-1 product per
# Gaussian (csrc/sh_exchange.cuh).  Instead of all-reducing 12*M B per Gaussian (192 B of the 236 B at M = 16),
# ranks all-gather the 12 B factor of each of the batch's views and every rank rebuilds the sum locally; the
# remaining 44 B per Gaussian (means3D, opacity, scales, rotations) go through the flat all-reduce as before.
# Pays off while 12 B * n_views stays below the all-reduce's 2 * 12*M B per Gaussian, i.e. n_views <~ 2*M.
SMALL_LEAVES = ("means3D", "opacities", "scales", "rotations")


class capture_sh_factors:
    """Context manager: rasterizer backward passes inside it skip dL/dsh and hand (campos, factors) of the view to
    `sink` -- a FactorExchange, or by default a list that collects them."""

    def __init__(self, sink=None):
        self.sink = [] if sink is None else sink

    def __enter__(self):
        import diff_gaussian_rasterization as dgr
        self._dgr, self._prev = dgr, dgr._sh_factor_sink
        dgr._sh_factor_sink = self.sink
        return self.sink

    def __exit__(self, *exc):
        self._dgr._sh_factor_sink = self._prev
        return False


class FactorExchange:
    """All-gather of the batch's SH factors, one collective per view SLOT (the j-th view of every rank), launched
    the moment this rank's j-th factors exist: the rasterizer calls emit() between its two backward stages, so
    the transfer runs under preprocess_bwd and under the following views.  Buffers persist across steps.

    Slot layout: fac_all[j, r] = factors of view j of rank r; ranks with fewer views contribute zeros (which add
    nothing to the rebuilt gradient)."""

    def __init__(self, per_rank: int, P: int, device, world: int, group=None):
        f32 = dict(dtype=torch.float32, device=device)
        self.per_rank, self.P, self.world, self.group = per_rank, P, world, group
        self.fac = torch.zeros(per_rank, P, 3, **f32)
        self.cam = torch.zeros(per_rank, 3, **f32)
        self.fac_all = torch.zeros(per_rank, world, P, 3, **f32)
        self.cam_all = torch.zeros(per_rank, world, 3, **f32)
        self.active = world > 1 and dist.is_initialized()
        self.nccl = self.active and dist.get_backend(group) == "nccl"
        self.begin()

    def begin(self):
        self.n, self.works = 0, []

    # -- the sink protocol of diff_gaussian_rasterization
    def buffer(self, P, device):
        if self.n >= self.per_rank:
            raise ValueError("more local views than per_rank")
        if P != self.P:
            raise ValueError("FactorExchange was sized for %d Gaussians, got %d" % (self.P, P))
        return self.fac[self.n]

    def emit(self, campos, factors):
        j = self.n
        if j >= self.per_rank:
            raise ValueError("more local views than per_rank")
        if factors.data_ptr() != self.fac[j].data_ptr():
            self.fac[j].copy_(factors.reshape(self.P, 3))
        self.cam[j].copy_(campos.reshape(3))
        self.n = j + 1
        self._gather(j)

    def _gather(self, j):
        if not self.active:
            self.fac_all[j, 0].copy_(self.fac[j])
            self.cam_all[j, 0].copy_(self.cam[j])
            return
        for dst, src in ((self.fac_all[j], self.fac[j]), (self.cam_all[j], self.cam[j])):
            if self.nccl:
                w = dist.all_gather_into_tensor(dst, src, group=self.group, async_op=True)
            else:
                w = dist.all_gather(list(dst.unbind(0)), src, group=self.group, async_op=True)
            self.works.append(w)

    def pad(self):
        """Launches the collectives of the slots this rank had no view for.  Every rank must issue its collectives
        in the same order, so call this before any other collective that follows the views (finish() does)."""
        while self.n < self.per_rank:
            j = self.n
            self.fac[j].zero_()
            self.cam[j].zero_()
            self.n = j + 1
            self._gather(j)

    def finish(self):
        """Joins the collectives.  Returns campos (V,3) and factors (V,P,3), V = per_rank * world, identical on
        every rank."""
        self.pad()
        for w in self.works:
            w.wait()
        self.works = []
        return self.cam_all.view(-1, 3), self.fac_all.view(-1, self.P, 3)


def gather_sh_factors(items, per_rank: int, P: int, device, world: int, group=None):
    """items: this rank's [(campos (3,), factors (P,3))], at most `per_rank`.  One-shot form of FactorExchange:
    returns campos (world*per_rank, 3) and factors (world*per_rank, P, 3), identical on every rank."""
    ex = FactorExchange(per_rank, P, device, world, group=group)
    for c, f in items:
        ex.emit(c, f)
    return ex.finish()


def _rebuild_cuda(means3D, campos, factors, degree, M):
    from diff_gaussian_rasterization import _C
    return _C.sh_grad_from_factors(means3D, campos, factors, degree, M)


def dp_step_factored(render_backward: Callable[..., Dict[str, torch.Tensor]], n_views: int, flat_small: FlatGrads,
                     rank: int, world: int, means3D: torch.Tensor, degree: int, M: int, rebuild=_rebuild_cuda,
                     group=None, exchange: "FactorExchange" = None) -> Dict[str, torch.Tensor]:
    """Like dp_step, but `render_backward(v, exchange)` returns only the SMALL_LEAVES gradients and hands the
    view's SH factors to `exchange` (run the rasterizer backward inside `capture_sh_factors(exchange)`, or call
    `exchange.emit(campos, factors)`); the summed dL/dsh is rebuilt from the all-gathered factors while the
    all-reduce of the small gradients is in flight.  Returns the reduced gradients of all five parameters."""
    per_rank = (n_views + world - 1) // world
    P = means3D.shape[0]
    ex = exchange if exchange is not None else FactorExchange(per_rank, P, means3D.device, world, group=group)
    if ex.per_rank != per_rank or ex.P != P or ex.world != world:
        raise ValueError("exchange was built for another batch shape")
    ex.begin()
    flat_small.zero_()
    for v in shard_views(n_views, rank, world):
        seen = ex.n
        flat_small.add_(render_backward(v, ex))
        if ex.n != seen + 1:
            raise RuntimeError("render_backward must hand exactly one view's SH factors to the exchange")
    ex.pad()  # same collective order on every rank: all slots first, then the all-reduce
    works = flat_small.all_reduce(group=group, async_op=True, world=world)
    cam_all, fac_all = ex.finish()
    out = dict(flat_small.views)
    out["sh"] = rebuild(means3D, cam_all, fac_all, degree, M)
    for w in works:
        w.wait()
    return out


# ---------------------------------------------------------------------------------------------------------------
# The same exchange with OUR kernels over NVLink peer memory instead of NCCL calls (csrc/peer_exchange.cuh).  The
# batch's SH factors and the small gradients live in ONE symmetric allocation per rank (torch's symmetric memory:
# plumbing -- allocation, peer mapping and the device-side barrier); after the local views,
#   barrier -> [ gsr_sh_grad_from_peers on a side stream  ||  gsr_peer_allreduce ] -> barrier.
# The gather of the factors never materialises (the rebuild kernel reads each view's factors from the rank that
# rendered it), and the all-reduce is one two-shot kernel (or NVSwitch multimem where the allocation has a multicast
# mapping).  CUDA only; the gloo tests cover the NCCL-shaped path above, bench.py's dp_check covers this one on GPUs.
class PeerExchange:
    """Factor sink (buffer / emit protocol of diff_gaussian_rasterization) + FlatGrads over a symmetric allocation."""

    def __init__(self, n_views: int, P: int, shapes_small: Dict[str, Sequence[int]], device, rank: int, world: int,
                 group=None, use_multimem: bool = False, gather: str = "push"):
        """gather: how the peers' view records reach this rank -- "push" (our kernel stores this rank's records into
        every peer's gather buffer, launched when the last view has emitted its factors, beside preprocess_bwd;
        fastest at 8 GPUs), "nccl" (one all-gather at the same point), "copy" (peer-to-peer copies by the copy
        engines) or "fused" (the rebuild kernel reads the peers' records in place over NVLink)."""
        import ctypes
        import torch.distributed._symmetric_memory as symm
        self.n_views, self.P, self.rank, self.world = n_views, P, rank, world
        self.per_rank = (n_views + world - 1) // world
        self.counts = [len(shard_views(n_views, r, world)) for r in range(world)]
        self.slot = (3 * P + 4 + 3) & ~3  # floats per view record: P*3 factors, 3 camera floats, padding
        n_flat = FlatGrads.numel(shapes_small)
        block = self.per_rank * self.slot
        n_gather = world * block if gather == "push" else 0  # peers store their record blocks here
        self.buf = symm.empty(block + n_flat + n_gather, dtype=torch.float32, device=device)
        self.hdl = symm.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
        self.flat_off = self.per_rank * self.slot
        self.flat = FlatGrads(shapes_small, device, flat=self.buf[self.flat_off:self.flat_off + n_flat])
        self.n_flat = n_flat
        ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.flat_ptrs = (ctypes.c_uint64 * world)(*[p + 4 * self.flat_off for p in ptrs])
        mc = int(self.hdl.multicast_ptr) if (use_multimem and getattr(self.hdl, "has_multicast_support", False)) else 0
        self.flat_mc = mc + 4 * self.flat_off if mc else 0
        # The batch's view records as the rebuild kernel sees them: this rank's own records in place, the peers'
        # records in a local staging copy that the copy engines fill over NVLink (no SM is spent on the gather, and
        # it runs under preprocess_bwd, which is HBM bound; the rebuild -- HBM bound too -- then runs under the
        # all-reduce, which is link bound).
        if gather not in ("nccl", "copy", "fused", "push"):
            raise ValueError("gather must be 'nccl', 'copy', 'fused' or 'push'")
        self.gather, self.group = gather, group
        pairs = [(j, r) for j in range(self.per_rank) for r in range(world) if j < self.counts[r]]
        remote = [(j, r) for (j, r) in pairs if r != rank]
        self.remote, self.gathered = [], None
        if gather == "copy":
            self.stage = torch.empty(max(len(remote), 1) * self.slot, dtype=torch.float32, device=device)
            self.remote = [(self.stage[k * self.slot:(k + 1) * self.slot],
                            self.hdl.get_buffer(r, (self.slot,), torch.float32, j * self.slot))
                           for k, (j, r) in enumerate(remote)]
            where = {jr: self.stage.data_ptr() + 4 * k * self.slot for k, jr in enumerate(remote)}
        elif gather == "nccl":
            self.gathered = torch.empty(world, self.per_rank * self.slot, dtype=torch.float32, device=device)
            where = {(j, r): self.gathered.data_ptr() + 4 * (r * self.per_rank + j) * self.slot for (j, r) in remote}
        elif gather == "push":
            g_off = self.flat_off + n_flat
            self.push_ptrs = (ctypes.c_uint64 * world)(*[p + 4 * g_off for p in ptrs])
            where = {(j, r): ptrs[rank] + 4 * (g_off + r * block + j * self.slot) for (j, r) in remote}
        else:
            where = {(j, r): ptrs[r] + 4 * j * self.slot for (j, r) in remote}
        views = [where.get((j, r), ptrs[rank] + 4 * j * self.slot) for (j, r) in pairs]
        self.view_ptrs = (ctypes.c_uint64 * max(len(views), 1))(*views)
        self.V = len(views)
        self.copy_streams = [torch.cuda.Stream(device=device, priority=-1) for _ in range(2)]
        self.push_ctas = 64
        self.side = torch.cuda.Stream(device=device, priority=-1)  # must get SMs while preprocess_bwd is dispatched
        self.n = 0
        self._sh = None
        self._early = None

    def begin(self, means3D: torch.Tensor = None, degree: int = 0, M: int = 16, out_sh: torch.Tensor = None):
        """Start of a step.  With means3D given, the SH rebuild is launched (side stream) the moment this rank's last
        view has emitted its factors -- i.e. between the two stages of that view's backward, so that the remote
        reads run under preprocess_bwd -- instead of inside exchange()."""
        self.n = 0
        self._sh = None
        self._early = None
        if means3D is not None:
            if M != 16:
                raise ValueError("PeerExchange is built for M = 16")
            self._early = (means3D, int(degree), int(M), out_sh)
        if self.counts[self.rank] == 0 and self._early is not None:
            self._launch_rebuild(*self._early)

    def _launch_rebuild(self, means3D, degree, M, out_sh):
        from diff_gaussian_rasterization import _lib
        dev = means3D.device
        sh = out_sh if out_sh is not None else torch.empty(self.P, M, 3, dtype=torch.float32, device=dev)
        cur = torch.cuda.current_stream(dev)
        if self.gather == "nccl":
            # one all-gather of the ranks' record blocks, launched here (between the two stages of the last view's
            # backward): it runs under preprocess_bwd on NCCL's stream
            work = dist.all_gather_into_tensor(self.gathered.view(-1), self.buf[:self.per_rank * self.slot],
                                               group=self.group, async_op=True)
            with torch.cuda.stream(self.side):
                work.wait()
                _lib.check(_lib.lib.gsr_sh_grad_from_peers(self.P, degree, M, self.V, means3D.data_ptr(),
                                                           self.view_ptrs, sh.data_ptr(), 0, self.side.cuda_stream))
        elif self.gather == "push":
            # our own all-gather: every rank stores its record block into every peer's gather buffer (posted NVLink
            # stores from a few CTAs, beside preprocess_bwd), then a barrier, then the rebuild from local memory
            self.side.wait_stream(cur)
            with torch.cuda.stream(self.side):
                block = self.per_rank * self.slot
                _lib.check(_lib.lib.gsr_peer_push(self.world, self.rank, self.push_ptrs, self.buf.data_ptr(), block,
                                                  block, self.push_ctas, self.side.cuda_stream))
                self.hdl.barrier(channel=0)  # every rank's pushes have landed
                _lib.check(_lib.lib.gsr_sh_grad_from_peers(self.P, degree, M, self.V, means3D.data_ptr(),
                                                           self.view_ptrs, sh.data_ptr(), 0, self.side.cuda_stream))
        else:
            self.side.wait_stream(cur)
            with torch.cuda.stream(self.side):
                self.hdl.barrier(channel=0)  # every rank's factors of this step are complete
            for cs in self.copy_streams:
                cs.wait_stream(self.side)
            for k, (dst, src) in enumerate(self.remote):  # peer -> local over NVLink, copy engines, two streams
                with torch.cuda.stream(self.copy_streams[k & 1]):
                    dst.copy_(src, non_blocking=True)
            with torch.cuda.stream(self.side):
                for cs in self.copy_streams:
                    self.side.wait_stream(cs)
                _lib.check(_lib.lib.gsr_sh_grad_from_peers(self.P, degree, M, self.V, means3D.data_ptr(),
                                                           self.view_ptrs, sh.data_ptr(), 0, self.side.cuda_stream))
        sh.record_stream(self.side)
        self._sh = sh

    # -- the sink protocol of diff_gaussian_rasterization
    def buffer(self, P, device):
        if self.n >= self.per_rank:
            raise ValueError("more local views than per_rank")
        if P != self.P:
            raise ValueError("PeerExchange was sized for %d Gaussians, got %d" % (self.P, P))
        j = self.n
        return self.buf[j * self.slot:j * self.slot + 3 * P].view(P, 3)

    def emit(self, campos, factors):
        j = self.n
        if j >= self.per_rank:
            raise ValueError("more local views than per_rank")
        rec = self.buf[j * self.slot:(j + 1) * self.slot]
        if factors.data_ptr() != rec.data_ptr():
            rec[:3 * self.P].copy_(factors.reshape(-1))
        rec[3 * self.P:3 * self.P + 3].copy_(campos.reshape(3))
        self.n = j + 1
        if self.n == self.counts[self.rank] and self._early is not None:
            self._launch_rebuild(*self._early)

    def exchange(self, means3D: torch.Tensor, degree: int, M: int, out_sh: torch.Tensor = None):
        """After this rank's views: all-reduce of the small gradients (in place in self.flat) and the batch's summed
        dL/dsh (P,M,3), rebuilt from every rank's factors.  Returns the SH gradient."""
        from diff_gaussian_rasterization import _lib
        if M != 16:
            raise ValueError("PeerExchange is built for M = 16")
        if self.n != self.counts[self.rank]:
            raise RuntimeError("this rank rendered %d views, its share of the batch is %d" % (self.n, self.counts[self.rank]))
        if self.flat._first:
            self.flat.flat.zero_()
            self.flat._first = False
        dev = means3D.device
        cur = torch.cuda.current_stream(dev)
        if self._sh is None:  # not launched early: factors barrier + rebuild now
            self._launch_rebuild(means3D, int(degree), int(M), out_sh)
        self.hdl.barrier(channel=1)  # every rank's small gradients are complete
        _lib.check(_lib.lib.gsr_peer_allreduce(self.world, self.rank, self.flat_ptrs, self.flat_mc, self.n_flat,
                                               cur.cuda_stream))
        cur.wait_stream(self.side)
        self.hdl.barrier(channel=2)  # all sums delivered; every peer is done reading this rank's factors
        sh, self._sh = self._sh, None
        return sh


def dp_step_peer(render_backward: Callable[..., Dict[str, torch.Tensor]], n_views: int, ex: "PeerExchange",
                 means3D: torch.Tensor, degree: int, M: int) -> Dict[str, torch.Tensor]:
    """dp_step_factored over peer memory: `render_backward(v, ex)` runs one view's forward + backward inside
    `capture_sh_factors(ex)` (and `grads_into(ex.flat)`), returning the SMALL_LEAVES gradients."""
    ex.begin(means3D, degree, M)
    ex.flat.zero_()
    for v in shard_views(n_views, ex.rank, ex.world):
        seen = ex.n
        ex.flat.add_(render_backward(v, ex))
        if ex.n != seen + 1:
            raise RuntimeError("render_backward must hand exactly one view's SH factors to the exchange")
    out = dict(ex.flat.views)
    out["sh"] = ex.exchange(means3D, degree, M)
    return out
