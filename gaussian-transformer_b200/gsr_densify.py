"""Densify / clone / split / prune of a `GaussianModel` on the library's compaction kernels.

Mirrors scene/gaussian_model.py:273-403 of the reference -- `prune_points`, `densify_and_clone`,
`densify_and_split`, `densify_and_prune`, with `_prune_optimizer` / `cat_tensors_to_optimizer` /
`densification_postfix` folded in -- for any object with the reference's attribute names (`_xyz`, `_features_dc`,
`_features_rest`, `_opacity`, `_scaling`, `_rotation`, `xyz_gradient_accum`, `denom`, `max_radii2D`, `optimizer`
with one parameter per group named xyz / f_dc / f_rest / opacity / scaling / rotation, `percent_dense`).

The reference rebuilds every parameter, both Adam moments and the statistics with `tensor[mask]` and `torch.cat`
(~40 nonzero / index / cat launches per call).  Here a mask becomes ONE index list (csrc/compact.cuh) and ONE launch
moves the rows of all 18 tensors; clone is a single gather ("the old rows, then the selected rows again"), and
split's append + removal of the split Gaussians is a single gather too.  Rows, their order, the fresh zero moments
and the random samples (`torch.normal` is called with the reference's shapes) are bit-identical to the reference's
statements (tests/test_gpu_densify.py).  The trainer-side change is three lines.
"""
import ctypes

import torch
from torch import nn

from diff_gaussian_rasterization import _lib

_NAMES = (("xyz", "_xyz"), ("f_dc", "_features_dc"), ("f_rest", "_features_rest"), ("opacity", "_opacity"),
          ("scaling", "_scaling"), ("rotation", "_rotation"))


class _Table(ctypes.Structure):
    _fields_ = [("src", ctypes.c_void_p), ("dst", ctypes.c_void_p), ("idx", ctypes.c_void_p),
                ("row_floats", ctypes.c_int32), ("reserved", ctypes.c_int32)]


def _stream(dev):
    return torch.cuda.current_stream(dev).cuda_stream


def mask_indices(mask, out=None, add=0):
    """Ascending indices of the set entries of a bool (P,) CUDA mask -> (int32 tensor, count); the order of
    tensor[mask].  `out`: write into the head of an existing int32 tensor (must be long enough).  One host
    synchronisation (the count), as `tensor[mask]` has."""
    dev = mask.device
    P = mask.numel()
    m = mask.contiguous().view(torch.uint8)
    nb = ctypes.c_size_t()
    _lib.check(_lib.lib.gsr_compact_scratch_bytes(P, ctypes.byref(nb)))
    scratch = torch.empty(nb.value, dtype=torch.uint8, device=dev)
    idx = out if out is not None else torch.empty(P, dtype=torch.int32, device=dev)
    count = torch.zeros(1, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib.gsr_compact_indices(P, m.data_ptr(), idx.data_ptr(), int(add), count.data_ptr(),
                                                scratch.data_ptr(), _stream(dev)))
    n = int(count.item())
    return idx, n


def gather_rows(pairs, n_rows, idx):
    """pairs: [(src, row_floats, idx_override or None)] -> list of new (n_rows * row_floats) float32 tensors, all
    moved by ONE launch; idx entries < 0 give zero rows."""
    dev = idx.device
    outs = [torch.empty(n_rows * w, dtype=torch.float32, device=dev) for (_, w, _) in pairs]
    if n_rows and pairs:
        arr = (_Table * len(pairs))()
        for k, ((src, w, ov), dst) in enumerate(zip(pairs, outs)):
            arr[k] = _Table(src.data_ptr(), dst.data_ptr(), ov.data_ptr() if ov is not None else None, w, 0)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib.gsr_gather_rows(len(pairs), arr, n_rows, idx.data_ptr(), _stream(dev)))
    return outs


def _rows(t):
    w = 1
    for d in t.shape[1:]:
        w *= int(d)
    return w


def _regather(model, n_rows, idx_param, idx_state, stats="gather"):
    """Rebuild the six parameters (rows idx_param), their Adam moments (rows idx_state, < 0 = fresh zeros) and the
    three statistics tensors in one launch, and install them like _prune_optimizer / cat_tensors_to_optimizer."""
    pairs, meta = [], []
    for group in model.optimizer.param_groups:
        assert len(group["params"]) == 1
        p = group["params"][0]
        w = _rows(p)
        st = model.optimizer.state.get(p, None)
        pairs.append((p.data.contiguous(), w, None))
        if st is not None:
            pairs.append((st["exp_avg"].contiguous(), w, idx_state))
            pairs.append((st["exp_avg_sq"].contiguous(), w, idx_state))
        meta.append((group, p, st))
    if stats == "gather":
        pairs += [(model.xyz_gradient_accum.contiguous(), 1, None), (model.denom.contiguous(), 1, None),
                  (model.max_radii2D.contiguous(), 1, None)]
    outs = gather_rows(pairs, n_rows, idx_param)
    k = 0
    new = {}
    for group, p, st in meta:
        shape = (n_rows,) + tuple(p.shape[1:])
        q = nn.Parameter(outs[k].view(shape).requires_grad_(True))
        k += 1
        if st is not None:
            st["exp_avg"] = outs[k].view(shape)
            st["exp_avg_sq"] = outs[k + 1].view(shape)
            k += 2
            del model.optimizer.state[p]
            model.optimizer.state[q] = st
        group["params"][0] = q
        new[group["name"]] = q
    for name, attr in _NAMES:
        setattr(model, attr, new[name])
    dev = idx_param.device
    if stats == "gather":
        model.xyz_gradient_accum = outs[k].view(n_rows, 1)
        model.denom = outs[k + 1].view(n_rows, 1)
        model.max_radii2D = outs[k + 2].view(n_rows)
    else:  # densification_postfix (gaussian_model.py:345-347): the statistics restart from zero
        model.xyz_gradient_accum = torch.zeros((n_rows, 1), device=dev)
        model.denom = torch.zeros((n_rows, 1), device=dev)
        model.max_radii2D = torch.zeros((n_rows,), device=dev)


def prune_points(model, mask):
    """scene/gaussian_model.py:291-305: drop the rows where `mask` is set (parameters, moments, statistics)."""
    keep = ~mask
    idx, n = mask_indices(keep)
    _regather(model, n, idx, idx, stats="gather")


def _build_rotation(r):
    """utils/general_utils.py:78-100 build_rotation (normalised quaternion (w,x,y,z) -> 3x3), same expressions."""
    norm = torch.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])
    q = r / norm[:, None]
    R = torch.zeros((q.size(0), 3, 3), device=r.device)
    r_, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R[:, 0, 0] = 1 - 2 * (y * y + z * z)
    R[:, 0, 1] = 2 * (x * y - r_ * z)
    R[:, 0, 2] = 2 * (x * z + r_ * y)
    R[:, 1, 0] = 2 * (x * y + r_ * z)
    R[:, 1, 1] = 1 - 2 * (x * x + z * z)
    R[:, 1, 2] = 2 * (y * z - r_ * x)
    R[:, 2, 0] = 2 * (x * z - r_ * y)
    R[:, 2, 1] = 2 * (y * z + r_ * x)
    R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def densify_and_clone(model, grads, grad_threshold, scene_extent):
    """scene/gaussian_model.py:374-387: small Gaussians with a large view-space gradient are duplicated.  One
    gather: rows [0..P) then the selected rows again; their moments are zero rows."""
    P = model._xyz.shape[0]
    dev = model._xyz.device
    sel = torch.where(torch.norm(grads, dim=-1) >= grad_threshold, True, False)
    sel = torch.logical_and(sel, torch.max(torch.exp(model._scaling), dim=1).values <= model.percent_dense * scene_extent)
    idx = torch.empty(2 * P, dtype=torch.int32, device=dev)
    torch.arange(P, dtype=torch.int32, device=dev, out=idx[:P])
    _, C = mask_indices(sel, out=idx[P:])
    idx_state = idx[:P + C].clone()
    idx_state[P:] = -1
    _regather(model, P + C, idx[:P + C], idx_state, stats="zero")
    return C


def densify_and_split(model, grads, grad_threshold, scene_extent, N=2):
    """scene/gaussian_model.py:349-372: large Gaussians with a large gradient are replaced by N samples of
    themselves, scaled down by 0.8 N.  The reference appends the N * S new rows and then prunes the S originals
    (two rebuilds of every tensor); here one gather produces [the kept old rows, the new rows]."""
    n0 = model._xyz.shape[0]
    dev = model._xyz.device
    padded_grad = torch.zeros((n0,), device=dev)
    padded_grad[:grads.shape[0]] = grads.squeeze()
    sel = torch.where(padded_grad >= grad_threshold, True, False)
    scaling = torch.exp(model._scaling)
    sel = torch.logical_and(sel, torch.max(scaling, dim=1).values > model.percent_dense * scene_extent)
    idx = torch.empty(n0 * (N + 1), dtype=torch.int32, device=dev)
    _, kept = mask_indices(~sel, out=idx)
    S = n0 - kept
    sel_idx = idx[kept:kept + S]
    mask_indices(sel, out=sel_idx)
    for k in range(1, N):
        idx[kept + k * S:kept + (k + 1) * S] = sel_idx
    n_new = kept + N * S
    sel_long = sel_idx.long()
    # the new rows' values that are not copies (same statements, same shapes, same RNG consumption as :357-362)
    stds = scaling.index_select(0, sel_long).repeat(N, 1)
    means = torch.zeros((stds.size(0), 3), device=dev)
    samples = torch.normal(mean=means, std=stds)
    rots = _build_rotation(model._rotation.index_select(0, sel_long)).repeat(N, 1, 1)
    new_xyz = torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + model._xyz.index_select(0, sel_long).repeat(N, 1)
    new_scaling = torch.log(scaling.index_select(0, sel_long).repeat(N, 1) / (0.8 * N))
    idx_state = idx[:n_new].clone()
    idx_state[kept:] = -1
    _regather(model, n_new, idx[:n_new], idx_state, stats="zero")
    with torch.no_grad():
        model._xyz.data[kept:] = new_xyz
        model._scaling.data[kept:] = new_scaling
    return S


def densify_and_prune(model, max_grad, min_opacity, extent, max_screen_size):
    """scene/gaussian_model.py:389-403."""
    grads = model.xyz_gradient_accum / model.denom
    grads[grads.isnan()] = 0.0
    densify_and_clone(model, grads, max_grad, extent)
    densify_and_split(model, grads, max_grad, extent)
    prune_mask = (torch.sigmoid(model._opacity) < min_opacity).squeeze()
    if max_screen_size:
        big_points_vs = model.max_radii2D > max_screen_size
        big_points_ws = torch.exp(model._scaling).max(dim=1).values > 0.1 * extent
        prune_mask = torch.logical_or(torch.logical_or(prune_mask, big_points_vs), big_points_ws)
    prune_points(model, prune_mask)
    torch.cuda.empty_cache()
