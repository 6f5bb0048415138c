"""`_C`-shaped shim over the C-ABI library.

Same three entry points, argument order and return tuples as the reference's
pybind module (DGR/ext.cpp:15-19, DGR/rasterize_points.h:18-67), so that
snapshot_fw.dump / snapshot_bw.dump argument tuples replay unchanged.  All
tensors are torch-owned; the library only sees raw device pointers
(torch is plumbing here: memory, streams).
"""
import ctypes
import threading

import torch

from . import _lib

_tls = threading.local()


def _host_result():
    """Pinned int64[2] per thread: [num_rendered, flags] written by gsr_forward_pre."""
    buf = getattr(_tls, "host_result", None)
    if buf is None:
        buf = torch.zeros(2, dtype=torch.int64).pin_memory()
        _tls.host_result = buf
    return buf


def _f32(t, device):
    """Dense fp32 CUDA view of `t` (DGR/rasterize_points.cu:94-111 .contiguous()), or None when absent."""
    if t is None or t.numel() == 0:
        return None
    if t.device != device:
        raise RuntimeError(f"gsrast: tensor on {t.device}, expected {device} (there is no CPU path)")
    if t.dtype != torch.float32:
        t = t.float()
    t = t.contiguous()
    if t.data_ptr() % 16:
        t = t.clone()
    return t


def _ptr(t):
    return 0 if t is None else t.data_ptr()


def _inputs(P, D, M, W, H, tanfovx, tanfovy, scale_modifier, prefiltered, debug, bg, means3D, sh, colors,
            opacity, scales, rotations, cov3D, view, proj, campos, sh_rest=None, raw_params=False):
    a = _lib.gsr_inputs()
    a.P, a.D, a.M, a.W, a.H = P, D, M, W, H
    a.tanfovx, a.tanfovy, a.scale_modifier = tanfovx, tanfovy, scale_modifier
    a.prefiltered, a.debug = int(bool(prefiltered)), int(bool(debug))
    a.bg, a.means3D, a.sh, a.colors_precomp = _ptr(bg), _ptr(means3D), _ptr(sh), _ptr(colors)
    a.opacities, a.scales, a.rotations, a.cov3D_precomp = _ptr(opacity), _ptr(scales), _ptr(rotations), _ptr(cov3D)
    a.viewmatrix, a.projmatrix, a.campos = _ptr(view), _ptr(proj), _ptr(campos)
    a.sh_rest = _ptr(sh_rest)
    a.raw_params = int(bool(raw_params))
    return a


def split_sh_ok(dc, rest):
    """Can (features_dc (P,1,3), features_rest (P,M-1,3)) be passed without concatenation?  (M = 16 or 4, dense
    float32 CUDA tensors on 16-byte boundaries -- anything else goes through torch.cat.)"""
    return (dc.is_cuda and rest.is_cuda and dc.dtype == rest.dtype == torch.float32 and dc.dim() == rest.dim() == 3
            and dc.size(1) == 1 and dc.size(2) == rest.size(2) == 3 and dc.size(0) == rest.size(0)
            and rest.size(1) + 1 in (4, 16) and dc.is_contiguous() and rest.is_contiguous()
            and dc.data_ptr() % 16 == 0 and rest.data_ptr() % 16 == 0)


def _check_means(means3D):
    if means3D.ndim != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")  # rasterize_points.cu:57-59
    if not means3D.is_cuda:
        raise RuntimeError("gsrast: means3D must be a CUDA tensor (there is no CPU path)")


_r_hint = {}  # (device index, P, W, H) -> recent num_rendered (sizes the speculative binning buffer)


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                        prefiltered, debug, sh_rest=None, raw_params=False):
    """-> (num_rendered, out_color (3,H,W), radii (P,) int32, geomBuffer, binningBuffer, imgBuffer).

    sh_rest (extension): split SH storage -- `sh` is features_dc (P,1,3) and sh_rest features_rest (P,M-1,3), read
    in place instead of through the torch.cat of scene/gaussian_model.py:108-111 (see split_sh_ok).
    raw_params (extension): opacity / scales / rotations are the raw _opacity / _scaling / _rotation; the kernels
    apply sigmoid / exp / normalize (gaussian_model.py:95-106) themselves."""
    _check_means(means3D)
    dev = means3D.device
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    u8 = dict(dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        if P == 0:  # rasterize_points.cu:66-67,80-114: image stays 0, buffers empty
            return (0, torch.zeros(3, H, W, dtype=torch.float32, device=dev),
                    torch.zeros(0, dtype=torch.int32, device=dev), torch.empty(0, **u8), torch.empty(0, **u8),
                    torch.empty(0, **u8))
        M = sh.size(1) if sh is not None and sh.numel() != 0 else 0
        if sh_rest is not None:
            M += sh_rest.size(1)
        t = [_f32(x, dev) for x in (background, means3D, sh, colors, opacity, scales, rotations, cov3D_precomp,
                                    viewmatrix, projmatrix, campos)]
        args = _inputs(P, int(degree), M, W, H, float(tan_fovx), float(tan_fovy), float(scale_modifier), prefiltered,
                       debug, *t, sh_rest=sh_rest, raw_params=raw_params)
        gb, _, ib, _ = _lib.scratch_bytes(P, 0, W, H)
        geom, image = torch.empty(gb, **u8), torch.empty(ib, **u8)
        radii = torch.empty(P, dtype=torch.int32, device=dev)
        out_color = torch.empty(3, H, W, dtype=torch.float32, device=dev)
        host = _host_result()
        stream = torch.cuda.current_stream(dev)
        # Phase 2 is launched speculatively against a binning buffer sized from the num_rendered this scene size
        # produced recently on this device (+25 %), so the GPU does not idle during the one host round trip (rasterizer_impl.cu:281);
        # a miss (first call, or a much denser view) falls back to allocate-then-launch.
        key = (dev.index, P, W, H)  # a scene of another size starts without a hint (one allocate-then-launch call)
        hint = _r_hint.get(key, 0)
        cap = min(hint + hint // 4 + 65536, 0x7FFFFFFF) if hint > 0 else 0
        binning = torch.empty(_lib.scratch_bytes(P, cap, W, H)[1] if cap > 0 else 0, **u8)
        _lib.check(_lib.lib.gsr_forward(ctypes.byref(args), geom.data_ptr(), _ptr(binning) if cap > 0 else 0, cap,
                                        image.data_ptr(), radii.data_ptr(), host.data_ptr(), out_color.data_ptr(),
                                        stream.cuda_stream))
        R, flags = int(host[0]), int(host[1])
        if flags & 1:
            raise RuntimeError("gsrast: point is filtered although prefiltered is set. This shouldn't happen!")
        if R > 0x7FFFFFFF:
            raise RuntimeError(f"gsrast: num_rendered = {R} does not fit a 32-bit int")
        if len(_r_hint) > 64:
            _r_hint.clear()
        _r_hint[key] = max(R, hint - hint // 16)  # follows shrinking scenes too (decays ~6 % per call)
        if R > cap or cap == 0:
            _, bb, _, _ = _lib.scratch_bytes(P, R, W, H)
            binning = torch.empty(bb if R > 0 else 0, **u8)
            _lib.check(_lib.lib.gsr_forward_post(ctypes.byref(args), geom.data_ptr(), _ptr(binning) if R > 0 else 0,
                                                 image.data_ptr(), R, out_color.data_ptr(), stream.cuda_stream))
    return R, out_color, radii, geom, binning, image


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                 cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh,
                                 degree, campos, geomBuffer, R, binningBuffer, imageBuffer, debug, sh_grad=True,
                                 factor_sink=None, out=None, sh_rest=None, debug_out=None, raw_params=False):
    """-> (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations).

    sh_grad=False (extension, used by gsr_dp): dL_dsh is not computed and returned as None; the caller rebuilds it
    from the view's SH factors and sh_grad_from_factors().  factor_sink (with sh_grad=False): an object with
    `buffer(P, device) -> (P,3) float32 tensor` and `emit(campos, factors)`; the factors are written into the buffer
    BETWEEN the two backward stages and emit() is called there, so that an exchange it launches overlaps the
    geometry stage.  out (extension): {"means3D" | "sh" | "opacity" | "scales" | "rot": tensor} -- gradients written
    into the caller's tensors (same shape, float32, contiguous) instead of freshly allocated ones.  sh_rest
    (extension): split SH storage as in rasterize_gaussians; dL_dsh is then the pair (dL_dfeatures_dc (P,1,3),
    dL_dfeatures_rest (P,M-1,3)).  debug_out (test hook): a dict that receives "conic" = dL_dconic (P,2,2) and
    "work" = the backward scratch buffer; nothing is retained otherwise.  raw_params (extension): scales /
    rotations are the raw parameters and dL_dopacity / dL_dscales / dL_drotations are the gradients of the RAW
    parameters (the activations' chain rule is folded into preprocess_bwd)."""
    _check_means(means3D)
    dev = means3D.device
    P = means3D.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if sh is not None and sh.numel() != 0 else 0
    split = sh_rest is not None
    if split:
        M += sh_rest.size(1)
    f32 = dict(dtype=torch.float32, device=dev)
    g_rest = torch.empty(P, M - 1, 3, **f32) if (split and sh_grad) else None
    g = dict(means3D=torch.empty(P, 3, **f32), means2D=torch.empty(P, 3, **f32),
             sh=torch.empty(P, 1 if split else M, 3, **f32) if sh_grad else None,
             colors=torch.empty(P, 3, **f32), opacity=torch.empty(P, 1, **f32), scales=torch.empty(P, 3, **f32),
             rot=torch.empty(P, 4, **f32), cov3D=torch.empty(P, 6, **f32), conic=torch.empty(P, 2, 2, **f32))
    for k, t in (out or {}).items():
        if k not in ("means3D", "sh", "opacity", "scales", "rot") or g[k] is None or (k == "sh" and split):
            continue
        if t.shape != g[k].shape or t.dtype != torch.float32 or t.device != dev or not t.is_contiguous():
            raise RuntimeError("out[%r] must be a contiguous float32 tensor of shape %s on %s" % (k, tuple(g[k].shape), dev))
        g[k] = t.view(t.shape)  # a fresh alias: autograd may adopt it as the leaf's .grad without a copy
    if P != 0:
        with torch.cuda.device(dev):
            t = [_f32(x, dev) for x in (background, means3D, sh, colors, None, scales, rotations, cov3D_precomp,
                                        viewmatrix, projmatrix, campos)]
            # opacities are not needed by backward (they live in geomBuffer); pass a valid pointer for the checker
            t[4] = t[1]
            args = _inputs(P, int(degree), M, W, H, float(tan_fovx), float(tan_fovy), float(scale_modifier), False,
                           debug, *t, sh_rest=sh_rest, raw_params=raw_params)
            dpix = _f32(dL_dout_color, dev)
            _, _, _, wb = _lib.scratch_bytes(P, int(R), W, H)
            work = torch.empty(wb, dtype=torch.uint8, device=dev)
            stream = torch.cuda.current_stream(dev).cuda_stream
            outs = (g["means3D"].data_ptr(), g["means2D"].data_ptr(), _ptr(g["sh"]) if (M and sh_grad) else 0,
                    g["colors"].data_ptr(), g["opacity"].data_ptr(), g["scales"].data_ptr(), g["rot"].data_ptr(),
                    g["cov3D"].data_ptr(), g["conic"].data_ptr())
            bin_ptr = binningBuffer.data_ptr() if binningBuffer.numel() else 0
            if split:
                outs = outs[:3] + (_ptr(g_rest) if sh_grad else 0,) + outs[3:]
                geometry = _lib.lib.gsr_backward_geometry_split
            else:
                geometry = _lib.lib.gsr_backward_geometry
            if factor_sink is None and not split:
                _lib.check(_lib.lib.gsr_backward(ctypes.byref(args), radii.data_ptr(), geomBuffer.data_ptr(), bin_ptr,
                                                 imageBuffer.data_ptr(), int(R), dpix.data_ptr(), *outs,
                                                 work.data_ptr(), stream))
            elif factor_sink is None:
                _lib.check(_lib.lib.gsr_backward_blend(ctypes.byref(args), geomBuffer.data_ptr(), bin_ptr,
                                                       imageBuffer.data_ptr(), int(R), dpix.data_ptr(),
                                                       work.data_ptr(), stream))
                _lib.check(geometry(ctypes.byref(args), radii.data_ptr(), geomBuffer.data_ptr(), *outs,
                                    work.data_ptr(), stream))
            else:
                _lib.check(_lib.lib.gsr_backward_blend(ctypes.byref(args), geomBuffer.data_ptr(), bin_ptr,
                                                       imageBuffer.data_ptr(), int(R), dpix.data_ptr(),
                                                       work.data_ptr(), stream))
                fac = factor_sink.buffer(P, dev)
                if fac.shape != (P, 3) or fac.dtype != torch.float32 or not fac.is_contiguous():
                    raise RuntimeError("factor_sink.buffer() must return a contiguous (P,3) float32 tensor")
                _lib.check(_lib.lib.gsr_backward_sh_factors(P, geomBuffer.data_ptr(), work.data_ptr(),
                                                            fac.data_ptr(), stream))
                factor_sink.emit(campos.detach().reshape(3), fac)
                _lib.check(geometry(ctypes.byref(args), radii.data_ptr(), geomBuffer.data_ptr(), *outs,
                                    work.data_ptr(), stream))
    elif factor_sink is not None:
        factor_sink.emit(campos.detach().reshape(3), factor_sink.buffer(0, dev))
    if P == 0 and out:
        for k in out:
            if g.get(k) is not None:
                g[k].zero_()
    if debug_out is not None:
        debug_out["conic"] = g["conic"]
        debug_out["work"] = work if P != 0 else None
    g_sh = (g["sh"], g_rest) if split else g["sh"]
    return (g["means2D"], g["colors"], g["opacity"], g["means3D"], g["cov3D"], g_sh, g["scales"], g["rot"])


def sh_color_factors(geomBuffer, dL_dcolors):
    """(P,3) dL/dcolor of one view with the channels the forward clamped zeroed: the per-Gaussian factor of that
    view's rank-1 SH gradient (include/gsrast.h gsr_sh_color_factors)."""
    P = dL_dcolors.size(0)
    dev = dL_dcolors.device
    out = torch.empty(P, 3, dtype=torch.float32, device=dev)
    if P != 0:
        with torch.cuda.device(dev):
            c = _f32(dL_dcolors, dev)
            _lib.check(_lib.lib.gsr_sh_color_factors(P, geomBuffer.data_ptr(), c.data_ptr(), out.data_ptr(),
                                                     torch.cuda.current_stream(dev).cuda_stream))
    return out


def sh_grad_from_factors(means3D, campos, factors, degree, M, out=None):
    """(P,M,3) sum over views v of basis(normalize(means3D - campos[v])) (x) factors[v]; campos (V,3), factors
    (V,P,3).  out: accumulate into an existing (P,M,3) tensor instead of allocating."""
    _check_means(means3D)
    dev = means3D.device
    P, V = means3D.size(0), campos.size(0)
    if factors.shape != (V, P, 3):
        raise RuntimeError("factors must have dimensions (num_views, num_points, 3)")
    acc = out is not None
    if not acc:
        out = torch.empty(P, M, 3, dtype=torch.float32, device=dev)
    if P != 0:
        with torch.cuda.device(dev):
            m, c, f = _f32(means3D, dev), _f32(campos, dev), _f32(factors, dev)
            _lib.check(_lib.lib.gsr_sh_grad_from_factors(P, int(degree), int(M), V, m.data_ptr(), _ptr(c), _ptr(f),
                                                         out.data_ptr(), int(acc),
                                                         torch.cuda.current_stream(dev).cuda_stream))
    return out


def mark_visible(means3D, viewmatrix, projmatrix):
    """bool (P,): view-space z > 0.2 (DGR/rasterize_points.cu:198-217)."""
    _check_means(means3D)
    dev = means3D.device
    P = means3D.size(0)
    present = torch.zeros(P, dtype=torch.bool, device=dev)
    if P != 0:
        with torch.cuda.device(dev):
            m, v, p = _f32(means3D, dev), _f32(viewmatrix, dev), _f32(projmatrix, dev)
            _lib.check(_lib.lib.gsr_mark_visible(P, m.data_ptr(), v.data_ptr(), p.data_ptr(), present.data_ptr(),
                                                 torch.cuda.current_stream(dev).cuda_stream))
    return present
