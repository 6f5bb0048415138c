"""Drop-in `diff_gaussian_rasterization` for stu214634/gaussian-transformer on B200.

Public surface (identical names, argument order and semantics to
DGR/diff_gaussian_rasterization/__init__.py:21-220):

    GaussianRasterizationSettings   NamedTuple, 12 fields, positional order kept
    GaussianRasterizer              nn.Module: forward(...) -> (color, radii); markVisible(positions)
    rasterize_gaussians             autograd entry returning (color, radii)

`gaussian_renderer.render()` imports exactly the first two
(gaussian_renderer/__init__.py:14) and runs unchanged.  Underneath is the
C-ABI library libgsrast.so (hand-written sm_100a kernels); there is no CPU or
PyTorch fallback.
"""
from typing import NamedTuple

import torch
from torch import nn

from . import _C

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


# Extension used by gsr_dp (data-parallel over views): while a sink is installed, the backward of a call that
# used SHs does not compute dL/dsh (the gradient returned for `shs` is None); it hands (campos, factors (P,3)) of
# the view to the sink instead -- BETWEEN the blend and the geometry stage of the backward, so that the exchange
# the sink launches overlaps the latter -- and gsr_dp rebuilds the summed SH gradient from the exchanged factors.
# A sink is a list (gets (campos, factors) appended) or an object with buffer(P, device) and emit(campos, factors).
# None (the default) = reference behaviour.
_sh_factor_sink = None


# Extension used by gsr_dp.grads_into: a dict {"means3D" | "sh" | "opacity" | "scales" | "rot": tensor}; the NEXT
# backward writes those gradients into the given tensors instead of allocating (consumed by that one call).
_grad_out = None


class _ListSink:
    def __init__(self, items):
        self.items = items

    def buffer(self, P, device):
        return torch.empty(P, 3, dtype=torch.float32, device=device)

    def emit(self, campos, factors):
        self.items.append((campos, factors))


def _snapshot(args, path, note):
    """debug=True contract (DGR __init__.py:83-90,132-139): dump the argument tuple on failure."""
    cpu = tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in args)

    def write():
        torch.save(cpu, path)
        print(note)
    return write


def _take_out():
    global _grad_out
    out, _grad_out = _grad_out, None
    return out


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, settings,
                sh_rest=None, raw_params=False):
        s = settings
        call = (s.bg, means3D, colors_precomp, opacities, scales, rotations, s.scale_modifier, cov3Ds_precomp,
                s.viewmatrix, s.projmatrix, s.tanfovx, s.tanfovy, s.image_height, s.image_width, sh, s.sh_degree,
                s.campos, s.prefiltered, s.debug)
        dump = _snapshot(call, "snapshot_fw.dump",
                         "\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.") \
            if s.debug else None
        try:
            num_rendered, color, radii, geom, binning, img = _C.rasterize_gaussians(*call, sh_rest=sh_rest,
                                                                                    raw_params=raw_params)
        except Exception:
            if dump:
                dump()
            raise
        ctx.raster_settings = s
        ctx.raw_params = bool(raw_params)
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning,
                              img, sh_rest)
        ctx.mark_non_differentiable(radii)
        return color, radii

    @staticmethod
    def backward(ctx, grad_color, _grad_radii):
        s = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img,
         sh_rest) = ctx.saved_tensors
        call = (s.bg, means3D, radii, colors_precomp, scales, rotations, s.scale_modifier, cov3Ds_precomp,
                s.viewmatrix, s.projmatrix, s.tanfovx, s.tanfovy, grad_color, sh, s.sh_degree, s.campos, geom,
                ctx.num_rendered, binning, img, s.debug)
        dump = _snapshot(call, "snapshot_bw.dump",
                         "\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n") \
            if s.debug else None
        sink = _sh_factor_sink if (sh is not None and sh.numel() != 0) else None
        if isinstance(sink, list):
            sink = _ListSink(sink)
        try:
            (g_means2D, g_colors, g_opac, g_means3D, g_cov3D, g_sh, g_scales,
             g_rot) = _C.rasterize_gaussians_backward(*call, sh_grad=sink is None, factor_sink=sink, out=_take_out(),
                                                      sh_rest=sh_rest, raw_params=ctx.raw_params)
        except Exception:
            if dump:
                dump()
            raise
        g_rest = None
        if sh_rest is not None:
            g_sh, g_rest = g_sh if g_sh is not None else (None, None)
        # order of forward's inputs; the settings tuple gets None (DGR __init__.py:143-153)
        return g_means3D, g_means2D, g_sh, g_colors, g_opac, g_scales, g_rot, g_cov3D, None, g_rest, None


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings, raw_params=False):
    """`sh` may also be the pair (features_dc (P,1,3), features_rest (P,M-1,3)) of scene/gaussian_model.py:108-111:
    the kernels then read (and write the gradients of) the two arrays in place instead of their torch.cat.
    raw_params=True (extension): `opacities`, `scales`, `rotations` are the model's raw `_opacity`, `_scaling`,
    `_rotation`; sigmoid / exp / normalize (gaussian_model.py:95-106) run inside preprocess_fwd and their chain rule
    inside preprocess_bwd, so the returned gradients are those of the raw parameters."""
    if raw_params and (scales is None or rotations is None or scales.numel() == 0 or rotations.numel() == 0):
        raise Exception('raw_params needs the raw scale/rotation pair (not a precomputed 3D covariance)')
    if isinstance(sh, (tuple, list)):
        dc, rest = sh
        if _C.split_sh_ok(dc, rest):
            return _RasterizeGaussians.apply(means3D, means2D, dc, colors_precomp, opacities, scales, rotations,
                                             cov3Ds_precomp, raster_settings, rest, raw_params)
        sh = torch.cat((dc, rest), dim=1)
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings, None, raw_params)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """Boolean mask of points in front of the near plane (DGR __init__.py:176-185)."""
        s = self.raster_settings
        with torch.no_grad():
            return _C.mark_visible(positions, s.viewmatrix, s.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, raw_params=False):
        # same validation and messages as DGR __init__.py:191-195
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        has_sr = scales is not None or rotations is not None
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (has_sr and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        empty = torch.Tensor([])  # "absent" marker, as in the reference (:197-207)
        pick = lambda t: empty if t is None else t
        return rasterize_gaussians(means3D, means2D, pick(shs), pick(colors_precomp), opacities, pick(scales),
                                   pick(rotations), pick(cov3D_precomp), self.raster_settings, raw_params)
