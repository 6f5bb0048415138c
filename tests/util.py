"""Helpers shared by the parity tests (test infrastructure)."""
import ctypes

import numpy as np
import torch

TENSOR_KEYS = ("means3D", "scales", "rotations", "opacities", "sh", "bg", "viewmatrix", "projmatrix", "campos")


def to_device(sc, dev="cuda"):
    d = dict(sc)
    for k in TENSOR_KEYS:
        d[k] = torch.from_numpy(np.ascontiguousarray(sc[k])).to(dev)
    return d


def settings_from(d, debug=False, prefiltered=False):
    from diff_gaussian_rasterization import GaussianRasterizationSettings
    return GaussianRasterizationSettings(
        image_height=int(d["H"]), image_width=int(d["W"]), tanfovx=d["tanfovx"], tanfovy=d["tanfovy"], bg=d["bg"],
        scale_modifier=d["scale_modifier"], viewmatrix=d["viewmatrix"], projmatrix=d["projmatrix"],
        sh_degree=int(d["sh_degree"]), campos=d["campos"], prefiltered=prefiltered, debug=debug)


def _slice(buf, ptr, nbytes, dtype):
    off = ptr - buf.data_ptr()
    assert 0 <= off and off + nbytes <= buf.numel(), (off, nbytes, buf.numel())
    return buf[off:off + nbytes].view(dtype)


def ours_forward(d, colors_precomp=None, cov3D_precomp=None, debug=False):
    """Run the product path through the _C-shaped shim and decode its scratch buffers (CPU tensors)."""
    from diff_gaussian_rasterization import _C, _lib
    empty = torch.Tensor([])
    use_sh = colors_precomp is None
    R, color, radii, geom, binning, img = _C.rasterize_gaussians(
        d["bg"], d["means3D"], empty if use_sh else colors_precomp, d["opacities"],
        d["scales"] if cov3D_precomp is None else empty, d["rotations"] if cov3D_precomp is None else empty,
        d["scale_modifier"], empty if cov3D_precomp is None else cov3D_precomp, d["viewmatrix"], d["projmatrix"],
        d["tanfovx"], d["tanfovy"], d["H"], d["W"], d["sh"] if use_sh else empty, d["sh_degree"], d["campos"], False,
        debug)
    torch.cuda.synchronize()
    P, W, H = int(d["P"]), int(d["W"]), int(d["H"])
    out = dict(num_rendered=R, color=color, radii=radii, buffers=(geom, binning, img))
    if P == 0:
        return out
    v = _lib.gsr_views()
    _lib.check(_lib.lib.gsr_get_views(P, R, W, H, geom.data_ptr(), binning.data_ptr() if R > 0 else 0,
                                      img.data_ptr(), ctypes.byref(v)))
    N, T = W * H, ((W + 15) // 16) * ((H + 15) // 16)
    g0 = _slice(geom, v.xy_depth_radius, 16 * P, torch.float32).view(P, 4).cpu()
    g1 = _slice(geom, v.conic_opacity, 16 * P, torch.float32).view(P, 4).cpu()
    g2 = _slice(geom, v.rgb_flags, 16 * P, torch.float32).view(P, 4).cpu()
    flags = g2[:, 3].contiguous().view(torch.int32)
    out.update(
        means2D=g0[:, :2].contiguous(), depths=g0[:, 2].contiguous(), conic_opacity=g1, rgb=g2[:, :3].contiguous(),
        clamped=torch.stack([(flags >> k) & 1 for k in range(3)], 1).to(torch.uint8), tiles_touched=flags >> 3,
        final_T=_slice(img, v.final_T, 4 * N, torch.float32).view(H, W).cpu(),
        n_contrib=_slice(img, v.n_contrib, 4 * N, torch.int32).view(H, W).cpu(),
        ranges=_slice(img, v.ranges, 8 * T, torch.int32).view(T, 2).cpu(),
        tile_count=_slice(img, v.tile_count, 4 * T, torch.int32).cpu(),
        point_list=_slice(binning, v.point_list, 4 * R, torch.int32).cpu() if R > 0
        else torch.zeros(0, dtype=torch.int32))
    return out


def ours_backward(d, fwd, dL_dpix, colors_precomp=None, cov3D_precomp=None, debug=False, dbg=None):
    from diff_gaussian_rasterization import _C
    empty = torch.Tensor([])
    use_sh = colors_precomp is None
    geom, binning, img = fwd["buffers"]
    r = _C.rasterize_gaussians_backward(
        d["bg"], d["means3D"], fwd["radii"], empty if use_sh else colors_precomp,
        d["scales"] if cov3D_precomp is None else empty, d["rotations"] if cov3D_precomp is None else empty,
        d["scale_modifier"], empty if cov3D_precomp is None else cov3D_precomp, d["viewmatrix"], d["projmatrix"],
        d["tanfovx"], d["tanfovy"], dL_dpix, d["sh"] if use_sh else empty, d["sh_degree"], d["campos"], geom,
        fwd["num_rendered"], binning, img, debug, debug_out=(dbg := {} if dbg is None else dbg))
    torch.cuda.synchronize()
    names = ("means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rot")
    g = dict(zip(names, r))
    g["conic"] = dbg["conic"].view(-1, 4)
    return g


def rel_err(a, b):
    """max |a-b| / max(|b|_inf, tiny): error relative to the largest entry of the checker's array."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    if a.size == 0:
        return 0.0
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def np_(t):
    return t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)
