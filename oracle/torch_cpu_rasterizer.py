"""TEST / BASELINE INFRASTRUCTURE -- a pure-PyTorch (CPU, fp32) implementation of `rasterize_gaussians`.

BASELINE.json's second baseline ("a pure-PyTorch CPU implementation of the same math timed on the box's host cores",
configs[0]): the reference has no CPU rasterizer, so this restates the math of its CUDA path with torch tensor ops and
lets autograd produce the backward:
  preprocess   DGR/cuda_rasterizer/forward.cu:155-256 (in_frustum auxiliary.h:139-164, computeCov3D :118-152,
               computeCov2D :74-113, SH -> RGB :20-71 as in utils/sh_utils.py:57-112, getRect auxiliary.h:46-56)
  binning      rasterizer_impl.cu:70-138 semantics (per tile: the Gaussians whose rectangle covers it, by depth, ties
               by index) -- evaluated per tile with a stable argsort instead of a global key sort
  blending     forward.cu:261-374: alpha = min(0.99, o exp(power)), skipped if power > 0 or alpha < 1/255, the pixel
               stops when T (1 - alpha) < 1e-4; vectorised over the tile's 256 pixels x list length with a cumulative
               product for T.
Only tests/, bench.py's cpu_baseline leg and tools/ import this; the product never does.  It is checked against the C
oracle (tests/test_cpu_torch_rasterizer.py); it is a float32 torch program, not bit-exact with the CUDA rounding.
"""
import math

import torch

C0 = 0.28209479177387814
C1 = 0.4886025119029199
C2 = (1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396)
C3 = (-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
      1.445305721320277, -0.5900435899266435)


def _sh_to_rgb(deg, sh, dirs):
    """sh (V,M,3), dirs (V,3) unit -> (V,3) before the +0.5 / clamp."""
    x, y, z = dirs[:, 0:1], dirs[:, 1:2], dirs[:, 2:3]
    res = C0 * sh[:, 0]
    if deg > 0:
        res = res - C1 * y * sh[:, 1] + C1 * z * sh[:, 2] - C1 * x * sh[:, 3]
    if deg > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        res = (res + C2[0] * xy * sh[:, 4] + C2[1] * yz * sh[:, 5] + C2[2] * (2.0 * zz - xx - yy) * sh[:, 6]
               + C2[3] * xz * sh[:, 7] + C2[4] * (xx - yy) * sh[:, 8])
        if deg > 2:
            res = (res + C3[0] * y * (3.0 * xx - yy) * sh[:, 9] + C3[1] * xy * z * sh[:, 10]
                   + C3[2] * y * (4.0 * zz - xx - yy) * sh[:, 11] + C3[3] * z * (2.0 * zz - 3.0 * xx - 3.0 * yy) * sh[:, 12]
                   + C3[4] * x * (4.0 * zz - xx - yy) * sh[:, 13] + C3[5] * z * (xx - yy) * sh[:, 14]
                   + C3[6] * x * (xx - 3.0 * yy) * sh[:, 15])
    return res


def _rotation_matrices(q):
    """Rot (V,3,3) of unnormalised quaternions (r,x,y,z) exactly as computeCov3D uses them (forward.cu:127-138)."""
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    rows = [1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
            2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
            2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)]
    return torch.stack(rows, 1).view(-1, 3, 3)


def preprocess(means3D, sh, opacities, scales, rotations, viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy,
               sh_degree, scale_modifier=1.0):
    """-> dict of per-Gaussian tensors for the visible ones (index `vis` into the inputs) + radii (P,) int32."""
    P = means3D.shape[0]
    ones = torch.ones(P, 1, dtype=means3D.dtype)
    hom = torch.cat([means3D, ones], 1)
    p_view = hom @ viewmatrix          # row-vector convention (scene/cameras.py:54-56)
    p_hom = hom @ projmatrix
    in_front = p_view[:, 2] > 0.2      # auxiliary.h:154
    p_w = 1.0 / (p_hom[:, 3] + 1e-7)
    ndc = p_hom[:, :2] * p_w[:, None]
    # cov3D = Rot S^2 Rot^T
    Rm = _rotation_matrices(rotations)
    # value mod * s; gradient as the reference returns it: computeCov3D's backward (backward.cu:278-341) gives
    # dL/d(mod * s) for dL_dscales, without the factor mod
    s_eff = scales + (scale_modifier - 1.0) * scales.detach()
    L = Rm * s_eff[:, None, :]
    Sigma = L @ L.transpose(1, 2)
    # EWA (forward.cu:74-113)
    fx, fy = W / (2.0 * tanfovx), H / (2.0 * tanfovy)
    tz = p_view[:, 2]
    tzs = torch.where(in_front, tz, torch.ones_like(tz))  # keep culled rows finite; they are masked out below
    tx = torch.clamp(p_view[:, 0] / tzs, -1.3 * tanfovx, 1.3 * tanfovx) * tzs
    ty = torch.clamp(p_view[:, 1] / tzs, -1.3 * tanfovy, 1.3 * tanfovy) * tzs
    zero = torch.zeros_like(tz)
    J = torch.stack([fx / tzs, zero, -fx * tx / (tzs * tzs), zero, fy / tzs, -fy * ty / (tzs * tzs)], 1).view(-1, 2, 3)
    Rwc = viewmatrix[:3, :3].t()       # world -> camera rotation
    A = J @ Rwc
    cov = A @ Sigma @ A.transpose(1, 2)
    a, b, c = cov[:, 0, 0] + 0.3, cov[:, 0, 1], cov[:, 1, 1] + 0.3
    det = a * c - b * b
    ok = in_front & (det != 0)
    dets = torch.where(ok, det, torch.ones_like(det))
    conic = torch.stack([c / dets, -b / dets, a / dets], 1)
    mid = 0.5 * (a + c)
    lam = mid + torch.sqrt(torch.clamp(mid * mid - det, min=0.1))
    radius = torch.ceil(3.0 * torch.sqrt(lam.detach()))
    px = ((ndc[:, 0].double() + 1.0) * W - 1.0) * 0.5   # ndc2Pix in double, as the reference evaluates it
    py = ((ndc[:, 1].double() + 1.0) * H - 1.0) * 0.5
    px, py = px.to(means3D.dtype), py.to(means3D.dtype)
    gx, gy = (W + 15) // 16, (H + 15) // 16
    pxd, pyd = px.detach(), py.detach()
    rect = torch.stack([torch.clamp(((pxd - radius) / 16).trunc(), 0, gx), torch.clamp(((pyd - radius) / 16).trunc(), 0, gy),
                        torch.clamp(((pxd + radius + 15) / 16).trunc(), 0, gx),
                        torch.clamp(((pyd + radius + 15) / 16).trunc(), 0, gy)], 1).to(torch.int64)
    tiles = (rect[:, 2] - rect[:, 0]) * (rect[:, 3] - rect[:, 1])
    ok = ok & (radius > 0) & (tiles > 0) & torch.isfinite(radius)
    vis = ok.nonzero().squeeze(1)
    dirs = means3D[vis] - campos
    dirs = dirs / dirs.norm(dim=1, keepdim=True)
    rgb = torch.clamp_min(_sh_to_rgb(sh_degree, sh[vis], dirs) + 0.5, 0.0)
    radii = torch.zeros(P, dtype=torch.int32)
    radii[vis] = radius[vis].to(torch.int32)
    return dict(vis=vis, xy=torch.stack([px, py], 1)[vis], depth=tz[vis], conic=conic[vis], opacity=opacities[vis, 0],
                rgb=rgb, rect=rect[vis], radii=radii, num_rendered=int(tiles[vis].sum()))


def render(pre, bg, W, H):
    """Tile by tile front-to-back blending -> color (3,H,W), final_T (H,W), n_contrib (H,W)."""
    gx, gy = (W + 15) // 16, (H + 15) // 16
    rect, depth = pre["rect"], pre["depth"].detach()
    rows = []
    T_img = torch.ones(H, W)
    n_img = torch.zeros(H, W, dtype=torch.int32)
    ar = torch.arange(16, dtype=torch.float32)
    for ty in range(gy):
        row_tiles = []
        in_row = ((rect[:, 1] <= ty) & (rect[:, 3] > ty)).nonzero().squeeze(1)
        h = min(16, H - 16 * ty)
        for tx in range(gx):
            w = min(16, W - 16 * tx)
            sel = in_row[(rect[in_row, 0] <= tx) & (rect[in_row, 2] > tx)]
            if sel.numel() == 0:
                row_tiles.append(bg.view(3, 1, 1).expand(3, h, w))
                continue
            sel = sel[torch.argsort(depth[sel], stable=True)]           # depth, ties by index
            xs = (16 * tx + ar[:w]).view(1, 1, w)
            ys = (16 * ty + ar[:h]).view(1, h, 1)
            dx = pre["xy"][sel, 0].view(-1, 1, 1) - xs
            dy = pre["xy"][sel, 1].view(-1, 1, 1) - ys
            con = pre["conic"][sel]
            power = -0.5 * (con[:, 0].view(-1, 1, 1) * dx * dx + con[:, 2].view(-1, 1, 1) * dy * dy) \
                - con[:, 1].view(-1, 1, 1) * dx * dy
            a_raw = pre["opacity"][sel].view(-1, 1, 1) * torch.exp(power)
            # value min(0.99, a); gradient passed through unclamped, as backward.cu:499-554 does (no clamp term)
            alpha = a_raw - (a_raw - torch.clamp_max(a_raw, 0.99)).detach()
            alpha = torch.where((power > 0) | (alpha < 1.0 / 255.0), torch.zeros_like(alpha), alpha)
            T_incl = torch.cumprod(1.0 - alpha, 0)
            live = T_incl >= 0.0001                                      # the pixel is done once this fails
            T_excl = torch.cat([torch.ones(1, h, w), T_incl[:-1]], 0)
            wgt = alpha * T_excl * live
            col = (wgt.unsqueeze(1) * pre["rgb"][sel].view(-1, 3, 1, 1)).sum(0)
            n_live = live.sum(0)
            T_fin = torch.where(n_live > 0, T_incl.gather(0, (n_live - 1).clamp_min(0).unsqueeze(0)).squeeze(0),
                                torch.ones(h, w))
            row_tiles.append(col + T_fin.unsqueeze(0) * bg.view(3, 1, 1))
            with torch.no_grad():
                T_img[16 * ty:16 * ty + h, 16 * tx:16 * tx + w] = T_fin
                contrib = (wgt > 0)
                idx = torch.arange(1, sel.numel() + 1).view(-1, 1, 1)
                n_img[16 * ty:16 * ty + h, 16 * tx:16 * tx + w] = (contrib * idx).amax(0).to(torch.int32)
        rows.append(torch.cat(row_tiles, 2))
    return torch.cat(rows, 1), T_img, n_img


def rasterize_gaussians_torch(means3D, sh, opacities, scales, rotations, viewmatrix, projmatrix, campos, bg, W, H,
                              tanfovx, tanfovy, sh_degree, scale_modifier=1.0):
    """-> (color (3,H,W), radii (P,) int32, info dict); differentiable w.r.t. the five parameter tensors."""
    pre = preprocess(means3D, sh, opacities, scales, rotations, viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy,
                     sh_degree, scale_modifier)
    color, final_T, n_contrib = render(pre, bg, W, H)
    return color, pre["radii"], dict(final_T=final_T, n_contrib=n_contrib, num_rendered=pre["num_rendered"])


def scene_tensors(sc, requires_grad=False):
    """numpy scene dict (gsr_scenes) -> the argument tuple of rasterize_gaussians_torch (CPU float32 tensors)."""
    import numpy as np
    t = lambda k: torch.from_numpy(np.ascontiguousarray(sc[k])).float()
    leaves = [t(k).requires_grad_(requires_grad) for k in ("means3D", "sh", "opacities", "scales", "rotations")]
    return leaves, dict(viewmatrix=t("viewmatrix"), projmatrix=t("projmatrix"), campos=t("campos"), bg=t("bg"),
                        W=int(sc["W"]), H=int(sc["H"]), tanfovx=float(sc["tanfovx"]), tanfovy=float(sc["tanfovy"]),
                        sh_degree=int(sc["sh_degree"]), scale_modifier=float(sc["scale_modifier"]))


def fwd_bwd(sc, dL_dpix):
    """One forward + backward on the CPU; returns (color, radii, grads list in LEAVES order, info)."""
    leaves, kw = scene_tensors(sc, requires_grad=True)
    color, radii, info = rasterize_gaussians_torch(*leaves, **kw)
    color.backward(torch.as_tensor(dL_dpix))
    return color.detach(), radii, [x.grad for x in leaves], info


__all__ = ["rasterize_gaussians_torch", "preprocess", "render", "scene_tensors", "fwd_bwd"]
_ = math  # (kept for interactive use)
