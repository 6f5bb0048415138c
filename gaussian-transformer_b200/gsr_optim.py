"""`gsr_optim.FusedAdam`: drop-in for the `torch.optim.Adam(l, lr=0.0, eps=1e-15)` of scene/gaussian_model.py:163
 on the library's one-launch multi-group Adam kernel (csrc/adam.cuh).

Same constructor arguments, `param_groups` (the trainer rewrites `group["lr"]` every iteration, gaussian_model.py:169-175)
and per-parameter `state` layout (`step`, `exp_avg`, `exp_avg_sq`) as torch's Adam, so the densification code that
cuts and concatenates optimizer state (gaussian_model.py:258-337) and `state_dict()` / `load_state_dict()`
(:73, :93) keep working.  CUDA float32 only; weight decay, amsgrad and maximize are not built (the reference
uses none of them) and raise.
"""
import ctypes

import torch

from diff_gaussian_rasterization import _lib


class _Group(ctypes.Structure):
    _fields_ = [("param", ctypes.c_void_p), ("grad", ctypes.c_void_p), ("exp_avg", ctypes.c_void_p),
                ("exp_avg_sq", ctypes.c_void_p), ("n", ctypes.c_int64), ("step", ctypes.c_int64),
                ("lr", ctypes.c_double)]


MAX_GROUPS = 8


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False, maximize=False):
        if weight_decay != 0 or amsgrad or maximize:
            raise ValueError("FusedAdam: weight_decay / amsgrad / maximize are not built")
        if not 0.0 <= lr or not 0.0 <= eps or not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0:
            raise ValueError("FusedAdam: invalid lr / eps / betas")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        # one launch per (device, betas, eps) bucket of at most MAX_GROUPS tensors
        buckets = {}
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda or p.dtype != torch.float32 or p.grad.dtype != torch.float32:
                    raise RuntimeError("FusedAdam: CUDA float32 parameters only (there is no CPU path)")
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                if not p.is_contiguous():
                    # e.g. _xyz built from a transposed numpy array (scene/dataset_readers.py:107-115 fetchPly): give
                    # the parameter dense storage once; the Parameter object (and every reference to it) stays
                    p.data = p.data.contiguous()
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)  # torch's layout: a CPU scalar tensor
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                st["step"] += 1
                for k in ("exp_avg", "exp_avg_sq"):  # densification may install non-contiguous cuts
                    if not st[k].is_contiguous():
                        st[k] = st[k].contiguous()
                g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                key = (p.device, tuple(group["betas"]), float(group["eps"]))
                buckets.setdefault(key, []).append((p, g, st, float(group["lr"])))
        for (dev, betas, eps), items in buckets.items():
            with torch.cuda.device(dev):
                stream = torch.cuda.current_stream(dev).cuda_stream
                for a in range(0, len(items), MAX_GROUPS):
                    part = items[a:a + MAX_GROUPS]
                    arr = (_Group * len(part))()
                    for i, (p, g, st, lr) in enumerate(part):
                        arr[i] = _Group(p.data_ptr(), g.data_ptr(), st["exp_avg"].data_ptr(),
                                        st["exp_avg_sq"].data_ptr(), p.numel(), int(st["step"]), lr)
                    _lib.check(_lib.lib.gsr_adam_step(len(part), arr, betas[0], betas[1], eps, stream))
        return loss


@torch.no_grad()
def densification_stats(radii, viewspace_grad, max_radii2D, xyz_gradient_accum, denom):
    """train.py:113-116 + GaussianModel.add_densification_stats (scene/gaussian_model.py:405-407) in one kernel and
    without the boolean-mask host synchronisation:

        visibility_filter = radii > 0
        max_radii2D[visibility_filter] = torch.max(max_radii2D[visibility_filter], radii[visibility_filter])
        xyz_gradient_accum[visibility_filter] += torch.norm(viewspace_grad[visibility_filter, :2], dim=-1, keepdim=True)
        denom[visibility_filter] += 1

    radii (P,) int32 and viewspace_grad (P,3) = screenspace_points.grad as the rasterizer returns them; the three
    statistics tensors ((P,), (P,1), (P,1) float32, contiguous) are updated in place."""
    P = radii.size(0)
    for t, n in ((viewspace_grad, 3 * P), (max_radii2D, P), (xyz_gradient_accum, P), (denom, P)):
        if not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != n:
            raise RuntimeError("densification_stats: dense float32 CUDA tensors of P rows expected")
    if not radii.is_cuda or radii.dtype != torch.int32 or not radii.is_contiguous():
        raise RuntimeError("densification_stats: radii must be the rasterizer's int32 CUDA tensor")
    if P:
        with torch.cuda.device(radii.device):
            _lib.check(_lib.lib.gsr_densification_stats(P, radii.data_ptr(), viewspace_grad.data_ptr(),
                                                        max_radii2D.data_ptr(), xyz_gradient_accum.data_ptr(),
                                                        denom.data_ptr(),
                                                        torch.cuda.current_stream(radii.device).cuda_stream))
