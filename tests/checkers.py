"""TEST INFRASTRUCTURE: ctypes bindings of the two checkers.

  * Oracle     -- oracle/_build/liboracle.so, the plain-C CPU restatement (oracle/oracle.c)
  * Reference  -- oracle/_ref/libref_rasterizer.so, the UNMODIFIED reference CUDA rasterizer
                  behind our extern "C" shim (oracle/ref_shim.cu); needs a GPU

Only tests/, __graft_entry__.smoke() and bench.py's baseline legs import this.
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "_build", "liboracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libref_rasterizer.so")
REF_KNN_SO = os.path.join(ROOT, "oracle", "_ref", "libref_knn.so")

_c = ctypes
_fp = _c.POINTER(_c.c_float)


def build_oracle():
    if not os.path.exists(ORACLE_SO):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], stdout=subprocess.DEVNULL)
    return ORACLE_SO


def _np_f(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return None if a is None else a.ctypes.data_as(_c.c_void_p)


class Oracle:
    """CPU oracle.  forward()/backward() take and return numpy arrays."""

    def __init__(self):
        self.lib = _c.CDLL(build_oracle())
        L = self.lib
        L.orc_create.restype = _c.c_void_p
        L.orc_destroy.argtypes = [_c.c_void_p]
        vp = _c.c_void_p
        L.orc_forward.restype = _c.c_long
        L.orc_forward.argtypes = [vp, _c.c_int, _c.c_int, _c.c_int, vp, _c.c_int, _c.c_int, vp, vp, vp, vp, vp,
                                  _c.c_float, vp, vp, vp, vp, vp, _c.c_float, _c.c_float, vp, vp]
        L.orc_backward.restype = None
        L.orc_backward.argtypes = [vp, _c.c_int, _c.c_int, _c.c_int, vp, _c.c_int, _c.c_int, vp, vp, vp, vp,
                                   _c.c_float, vp, vp, vp, vp, vp, _c.c_float, _c.c_float, vp, vp] + [vp] * 9
        L.orc_mark_visible.argtypes = [_c.c_int, vp, vp, vp, vp]
        L.orc_higher_msb.restype = _c.c_uint32
        L.orc_higher_msb.argtypes = [_c.c_uint32]
        L.orc_num_threads.restype = _c.c_int
        for name, ty in (("depths", _c.c_float), ("means2D", _c.c_float), ("cov3D", _c.c_float),
                         ("conic_opacity", _c.c_float), ("rgb", _c.c_float), ("clamped", _c.c_ubyte),
                         ("tiles_touched", _c.c_uint32), ("point_offsets", _c.c_uint32), ("keys", _c.c_uint64),
                         ("point_list", _c.c_uint32), ("final_T", _c.c_float), ("n_contrib", _c.c_uint32),
                         ("ranges", _c.c_uint32)):
            f = getattr(L, "orc_" + name)
            f.restype = _c.POINTER(ty)
            f.argtypes = [vp]
        self.h = L.orc_create()
        self.sc = None

    def __del__(self):
        try:
            self.lib.orc_destroy(self.h)
        except Exception:
            pass

    def num_threads(self):
        return self.lib.orc_num_threads()

    def _view(self, name, n, dtype):
        ptr = getattr(self.lib, "orc_" + name)(self.h)
        if n == 0 or not ptr:
            return np.zeros(0, dtype=dtype)
        return np.ctypeslib.as_array(ptr, shape=(n,)).copy()

    def forward(self, sc, colors_precomp=None, cov3D_precomp=None, internals=True):
        P, W, H = int(sc["P"]), int(sc["W"]), int(sc["H"])
        M, D = int(sc["M"]), int(sc["sh_degree"])
        use_sh = colors_precomp is None
        a = dict(bg=_np_f(sc["bg"]), means3D=_np_f(sc["means3D"]), sh=_np_f(sc["sh"]) if use_sh else None,
                 colors=_np_f(colors_precomp), opac=_np_f(sc["opacities"]),
                 scales=_np_f(sc["scales"]) if cov3D_precomp is None else None,
                 rot=_np_f(sc["rotations"]) if cov3D_precomp is None else None, cov=_np_f(cov3D_precomp),
                 view=_np_f(sc["viewmatrix"]), proj=_np_f(sc["projmatrix"]), campos=_np_f(sc["campos"]))
        self.args = a
        self.sc = sc
        color = np.zeros((3, H, W), dtype=np.float32)
        radii = np.zeros(P, dtype=np.int32)
        R = self.lib.orc_forward(self.h, P, D, M if use_sh else 0, _p(a["bg"]), W, H, _p(a["means3D"]), _p(a["sh"]),
                                 _p(a["colors"]), _p(a["opac"]), _p(a["scales"]), float(sc["scale_modifier"]),
                                 _p(a["rot"]), _p(a["cov"]), _p(a["view"]), _p(a["proj"]), _p(a["campos"]),
                                 float(sc["tanfovx"]), float(sc["tanfovy"]), _p(color), _p(radii))
        out = dict(num_rendered=int(R), color=color, radii=radii)
        self.radii = radii
        if internals and P > 0 and R >= 0:
            N, T = W * H, ((W + 15) // 16) * ((H + 15) // 16)
            out.update(
                depths=self._view("depths", P, np.float32), means2D=self._view("means2D", 2 * P, np.float32).reshape(P, 2),
                conic_opacity=self._view("conic_opacity", 4 * P, np.float32).reshape(P, 4),
                rgb=self._view("rgb", 3 * P, np.float32).reshape(P, 3),
                clamped=self._view("clamped", 3 * P, np.uint8).reshape(P, 3),
                tiles_touched=self._view("tiles_touched", P, np.uint32),
                point_list=self._view("point_list", R, np.uint32), keys=self._view("keys", R, np.uint64),
                final_T=self._view("final_T", N, np.float32).reshape(H, W),
                n_contrib=self._view("n_contrib", N, np.uint32).reshape(H, W),
                ranges=self._view("ranges", 2 * T, np.uint32).reshape(T, 2))
        return out

    def backward(self, dL_dpix):
        sc, a = self.sc, self.args
        P, W, H = int(sc["P"]), int(sc["W"]), int(sc["H"])
        use_sh = a["colors"] is None
        M, D = (int(sc["M"]) if use_sh else 0), int(sc["sh_degree"])
        dpix = _np_f(dL_dpix)
        g = dict(means2D=np.zeros((P, 3), np.float32), conic=np.zeros((P, 4), np.float32),
                 opacity=np.zeros((P, 1), np.float32), colors=np.zeros((P, 3), np.float32),
                 means3D=np.zeros((P, 3), np.float32), cov3D=np.zeros((P, 6), np.float32),
                 sh=np.zeros((P, max(M, 0), 3), np.float32), scales=np.zeros((P, 3), np.float32),
                 rot=np.zeros((P, 4), np.float32))
        self.lib.orc_backward(self.h, P, D, M, _p(a["bg"]), W, H, _p(a["means3D"]), _p(a["sh"]), _p(a["colors"]),
                              _p(a["scales"]), float(sc["scale_modifier"]), _p(a["rot"]), _p(a["cov"]),
                              _p(a["view"]), _p(a["proj"]), _p(a["campos"]), float(sc["tanfovx"]),
                              float(sc["tanfovy"]), _p(self.radii), _p(dpix), _p(g["means2D"]), _p(g["conic"]),
                              _p(g["opacity"]), _p(g["colors"]), _p(g["means3D"]), _p(g["cov3D"]),
                              _p(g["sh"]) if M > 0 else None, _p(g["scales"]), _p(g["rot"]))
        return g

    def mark_visible(self, means3D, view, proj):
        m, v, p = _np_f(means3D), _np_f(view), _np_f(proj)
        out = np.zeros(len(m), dtype=np.uint8)
        self.lib.orc_mark_visible(len(m), _p(m), _p(v), _p(p), _p(out))
        return out.astype(bool)

    def higher_msb(self, n):
        return int(self.lib.orc_higher_msb(n))

    def l1_ssim(self, img, gt, lambda_dssim, grad=True):
        """(L1, SSIM, loss) as float64 and d loss / d img (oracle.c orc_l1_ssim).  img, gt: (C,H,W) float32."""
        img, gt = _np_f(img), _np_f(gt)
        C, H, W = img.shape
        out3 = np.zeros(3, dtype=np.float64)
        g = np.zeros_like(img) if grad else None
        self.lib.orc_l1_ssim.argtypes = [_c.c_int, _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p, _c.c_double,
                                         _c.c_void_p, _c.c_void_p]
        self.lib.orc_l1_ssim.restype = None
        self.lib.orc_l1_ssim(C, H, W, _p(img), _p(gt), float(lambda_dssim), _p(out3), _p(g))
        return out3, g

    def adam_step(self, p, g, m, v, step, lr, beta1=0.9, beta2=0.999, eps=1e-15):
        """In-place Adam update of float32 numpy arrays p, m, v (oracle.c orc_adam_step)."""
        for a in (p, g, m, v):
            assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
        self.lib.orc_adam_step.argtypes = [_c.c_long, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_long,
                                           _c.c_double, _c.c_double, _c.c_double, _c.c_double]
        self.lib.orc_adam_step.restype = None
        self.lib.orc_adam_step(p.size, _p(p), _p(g), _p(m), _p(v), int(step), lr, beta1, beta2, eps)

    def knn3(self, points):
        """(P,) mean squared distance to the 3 nearest neighbours, brute force (oracle.c orc_knn3)."""
        pts = _np_f(points)
        out = np.zeros(len(pts), dtype=np.float32)
        self.lib.orc_knn3.argtypes = [_c.c_int, _c.c_void_p, _c.c_void_p]
        self.lib.orc_knn3(len(pts), _p(pts), _p(out))
        return out


def ref_knn(points):
    """The UNMODIFIED reference simple_knn (oracle/_ref/libref_knn.so) on a CUDA tensor (P,3) -> (P,) tensor."""
    import torch
    lib = _c.CDLL(REF_KNN_SO)
    lib.ref_knn.argtypes = [_c.c_int, _c.c_void_p, _c.c_void_p]
    pts = points.to(torch.float32).contiguous()
    out = torch.zeros(pts.shape[0], dtype=torch.float32, device=pts.device)
    torch.cuda.synchronize()
    rc = lib.ref_knn(pts.shape[0], pts.data_ptr(), out.data_ptr())
    assert rc == 0, f"reference simple_knn failed: cuda error {rc}"
    return out


class _RefViews(_c.Structure):
    _fields_ = [(n, _c.c_void_p) for n in (
        "depths", "clamped", "internal_radii", "means2D", "cov3D", "conic_opacity", "rgb", "tiles_touched",
        "point_offsets", "accum_alpha", "n_contrib", "ranges", "point_list", "point_list_keys",
        "point_list_unsorted", "point_list_keys_unsorted")]


class Reference:
    """The reference CUDA rasterizer (oracle/_ref).  Works on torch CUDA tensors (legacy default stream)."""

    def __init__(self):
        import torch
        self.torch = torch
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO + " (built by `make -C oracle ref` where /root/reference exists)")
        self.lib = _c.CDLL(REF_SO)
        L = self.lib
        vp, ci, cf = _c.c_void_p, _c.c_int, _c.c_float
        L.ref_create.restype = vp
        L.ref_destroy.argtypes = [vp]
        L.ref_last_error.restype = _c.c_char_p
        L.ref_forward.restype = ci
        L.ref_forward.argtypes = [vp, ci, ci, ci, vp, ci, ci, vp, vp, vp, vp, vp, cf, vp, vp, vp, vp, vp, cf, cf, ci,
                                  vp, vp, ci, ci]
        L.ref_backward.restype = ci
        L.ref_backward.argtypes = [vp, ci, ci, ci, ci, vp, ci, ci, vp, vp, vp, vp, cf, vp, vp, vp, vp, vp, cf, cf,
                                   vp, vp] + [vp] * 9 + [ci, ci]
        L.ref_mark_visible.argtypes = [ci, vp, vp, vp, vp]
        L.ref_views.argtypes = [vp, _c.POINTER(_RefViews)]
        L.ref_copy.argtypes = [vp, vp, _c.c_size_t]
        self.h = L.ref_create()

    def __del__(self):
        try:
            self.torch.cuda.synchronize()
            self.lib.ref_destroy(self.h)
        except Exception:
            pass

    @staticmethod
    def _dp(t):
        return 0 if t is None else t.data_ptr()

    def forward(self, d, colors_precomp=None, cov3D_precomp=None, internals=True, sync=True):
        """d: dict of torch CUDA tensors + scalars (see tests/util.to_device)."""
        torch = self.torch
        P, W, H = int(d["P"]), int(d["W"]), int(d["H"])
        use_sh = colors_precomp is None
        M = int(d["M"]) if use_sh else 0
        dev = d["means3D"].device
        color = torch.empty(3, H, W, dtype=torch.float32, device=dev)
        radii = torch.empty(P, dtype=torch.int32, device=dev)
        torch.cuda.current_stream().synchronize()  # the reference uses the legacy default stream
        self.a = dict(sh=d["sh"] if use_sh else None, colors=colors_precomp,
                      scales=d["scales"] if cov3D_precomp is None else None,
                      rot=d["rotations"] if cov3D_precomp is None else None, cov=cov3D_precomp)
        self.d, self.M = d, M
        a = self.a
        R = self.lib.ref_forward(self.h, P, int(d["sh_degree"]), M, self._dp(d["bg"]), W, H, self._dp(d["means3D"]),
                                 self._dp(a["sh"]), self._dp(a["colors"]), self._dp(d["opacities"]),
                                 self._dp(a["scales"]), float(d["scale_modifier"]), self._dp(a["rot"]),
                                 self._dp(a["cov"]), self._dp(d["viewmatrix"]), self._dp(d["projmatrix"]),
                                 self._dp(d["campos"]), float(d["tanfovx"]), float(d["tanfovy"]), 0,
                                 color.data_ptr(), radii.data_ptr(), 0, 1)
        if R < 0:
            raise RuntimeError("reference forward failed: " + self.lib.ref_last_error().decode())
        if sync:
            torch.cuda.synchronize()
        self.R, self.radii = R, radii
        out = dict(num_rendered=R, color=color, radii=radii)
        if internals and P > 0:
            torch.cuda.synchronize()
            v = _RefViews()
            self.lib.ref_views(self.h, _c.byref(v))
            N, T = W * H, ((W + 15) // 16) * ((H + 15) // 16)

            def grab(ptr, n, dtype):
                if not ptr or n == 0:
                    return torch.zeros(0, dtype=dtype)
                t = torch.empty(n, dtype=dtype, device=dev)
                nbytes = n * t.element_size()
                rc = self.lib.ref_copy(t.data_ptr(), ptr, nbytes)
                assert rc == 0, rc
                return t.cpu()
            out.update(depths=grab(v.depths, P, torch.float32), means2D=grab(v.means2D, 2 * P, torch.float32).view(P, 2),
                       conic_opacity=grab(v.conic_opacity, 4 * P, torch.float32).view(P, 4),
                       rgb=grab(v.rgb, 3 * P, torch.float32).view(P, 3),
                       clamped=grab(v.clamped, 3 * P, torch.uint8).view(P, 3),
                       tiles_touched=grab(v.tiles_touched, P, torch.int32),
                       final_T=grab(v.accum_alpha, N, torch.float32).view(H, W),
                       n_contrib=grab(v.n_contrib, N, torch.int32).view(H, W),
                       ranges=grab(v.ranges, 2 * T, torch.int32).view(T, 2),
                       point_list=grab(v.point_list, R, torch.int32) if R > 0 else torch.zeros(0, dtype=torch.int32))
        return out

    def backward(self, dL_dpix, sync=True):
        torch = self.torch
        d, a, M = self.d, self.a, self.M
        P, W, H = int(d["P"]), int(d["W"]), int(d["H"])
        dev = d["means3D"].device
        f = dict(dtype=torch.float32, device=dev)
        g = dict(means2D=torch.empty(P, 3, **f), conic=torch.empty(P, 4, **f), opacity=torch.empty(P, 1, **f),
                 colors=torch.empty(P, 3, **f), means3D=torch.empty(P, 3, **f), cov3D=torch.empty(P, 6, **f),
                 sh=torch.empty(P, M, 3, **f), scales=torch.empty(P, 3, **f), rot=torch.empty(P, 4, **f))
        torch.cuda.current_stream().synchronize()
        rc = self.lib.ref_backward(self.h, P, int(d["sh_degree"]), M, int(self.R), self._dp(d["bg"]), W, H,
                                   self._dp(d["means3D"]), self._dp(a["sh"]), self._dp(a["colors"]),
                                   self._dp(a["scales"]), float(d["scale_modifier"]), self._dp(a["rot"]),
                                   self._dp(a["cov"]), self._dp(d["viewmatrix"]), self._dp(d["projmatrix"]),
                                   self._dp(d["campos"]), float(d["tanfovx"]), float(d["tanfovy"]),
                                   self.radii.data_ptr(), dL_dpix.data_ptr(), g["means2D"].data_ptr(),
                                   g["conic"].data_ptr(), g["opacity"].data_ptr(), g["colors"].data_ptr(),
                                   g["means3D"].data_ptr(), g["cov3D"].data_ptr(),
                                   g["sh"].data_ptr() if M > 0 else 0, g["scales"].data_ptr(), g["rot"].data_ptr(),
                                   0, 1)
        if rc != 0:
            raise RuntimeError("reference backward failed: " + self.lib.ref_last_error().decode())
        if sync:
            torch.cuda.synchronize()
        return g

    def mark_visible(self, means3D, view, proj):
        torch = self.torch
        present = torch.zeros(means3D.shape[0], dtype=torch.bool, device=means3D.device)
        torch.cuda.synchronize()
        self.lib.ref_mark_visible(means3D.shape[0], means3D.data_ptr(), view.data_ptr(), proj.data_ptr(),
                                  present.data_ptr())
        torch.cuda.synchronize()
        return present
