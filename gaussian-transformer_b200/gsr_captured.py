"""Small-scene path: `rasterize_gaussians` forward + backward of a FIXED shape replayed as two CUDA graphs.

At the sizes of the transformer trainers and of BASELINE config C1 (train_transformer.py:79-81,185-187: a few 10^4
Gaussians at 800x450) a step is ~16 kernels of 5-40 us each; the reference -- and this library's `_C`-shaped entry --
spend more time on launch latency, allocation and the num_rendered host round trip
(DGR/cuda_rasterizer/rasterizer_impl.cu:281) than on the kernels.  Here all buffers are allocated once, the forward
runs through `gsr_forward_capacity` (no host synchronisation: the lists are written into a buffer of `capacity` entries
and the kernels check the device-side num_rendered against it), and both directions are captured once and replayed.

Same math, same kernels, same results as `_C.rasterize_gaussians` / `rasterize_gaussians_backward` (the image is
bit-identical; gradients differ by the atomics' summation order only, tests/test_gpu_small_scene.py).  When a view needs
more than `capacity` list entries nothing is rendered, the backward is a no-op, and `forward(check=True)` (default)
notices, doubles the buffer, re-captures and repeats the step; with check=False the caller asks `fits()` after its own
synchronisation.

    cr = CapturedRasterizer(P, M, sh_degree, W, H, tanfovx, tanfovy, bg)
    cr.set_camera(viewmatrix, projmatrix, campos)
    color, radii = cr.rasterize(means3D, sh, opacity, scales, rotations)   # autograd-aware
    loss(color).backward()
"""
import ctypes

import torch

from diff_gaussian_rasterization import _C, _lib


class CapturedRasterizer:
    def __init__(self, P, M, sh_degree, W, H, tanfovx, tanfovy, bg, scale_modifier=1.0, capacity=None,
                 device="cuda", raw_params=False):
        if P <= 0:
            raise ValueError("CapturedRasterizer needs P > 0")
        self.dev = torch.device(device if str(device) != "cuda" else "cuda:%d" % torch.cuda.current_device())
        self.P, self.M, self.D, self.W, self.H = int(P), int(M), int(sh_degree), int(W), int(H)
        self.tanfovx, self.tanfovy, self.scale_modifier = float(tanfovx), float(tanfovy), float(scale_modifier)
        self.raw_params = bool(raw_params)
        f32 = dict(dtype=torch.float32, device=self.dev)
        z = lambda *s: torch.zeros(*s, **f32)
        # static inputs (the graphs read these addresses)
        self.inp = dict(means3D=z(P, 3), sh=z(P, M, 3), opacity=z(P, 1), scales=z(P, 3), rotations=z(P, 4))
        self.inp["rotations"][:, 0] = 1.0
        self.viewmatrix, self.projmatrix, self.campos = z(4, 4), z(4, 4), z(3)
        self.bg = torch.as_tensor(bg, dtype=torch.float32).to(self.dev).reshape(3).contiguous()
        self.dL_dout = z(3, H, W)
        # static outputs
        self.out_color = z(3, H, W)
        self.radii = torch.zeros(P, dtype=torch.int32, device=self.dev)
        self.grads = dict(means3D=z(P, 3), means2D=z(P, 3), sh=z(P, M, 3), colors=z(P, 3), opacity=z(P, 1),
                          scales=z(P, 3), rot=z(P, 4), cov3D=z(P, 6), conic=z(P, 2, 2))
        gb, _, ib, wb = _lib.scratch_bytes(P, 0, W, H)
        u8 = dict(dtype=torch.uint8, device=self.dev)
        self._geom, self._image, self._work = torch.empty(gb, **u8), torch.empty(ib, **u8), torch.empty(wb, **u8)
        self._host = torch.zeros(2, dtype=torch.int64).pin_memory()
        self._done = torch.cuda.Event()
        self.capacity = int(capacity) if capacity else 0
        self._binning = None
        self._fwd_graph = self._bwd_graph = None
        self.recaptures = 0

    # ---- inputs ------------------------------------------------------------------------------------------------
    def set_camera(self, viewmatrix, projmatrix, campos):
        """Row-vector matrices as in scene/cameras.py:54-58 (world_view_transform, full_proj_transform, camera_center)."""
        self.viewmatrix.copy_(viewmatrix.reshape(4, 4))
        self.projmatrix.copy_(projmatrix.reshape(4, 4))
        self.campos.copy_(campos.reshape(3))

    def load(self, means3D=None, sh=None, opacity=None, scales=None, rotations=None):
        """Copy the given tensors into the static inputs (tensors that ARE the static inputs are left alone)."""
        for name, t in (("means3D", means3D), ("sh", sh), ("opacity", opacity), ("scales", scales),
                        ("rotations", rotations)):
            if t is None:
                continue
            dst = self.inp[name]
            if t.data_ptr() != dst.data_ptr():
                dst.copy_(t.detach().reshape(dst.shape))

    # ---- launches ----------------------------------------------------------------------------------------------
    def _args(self):
        i = self.inp
        return _C._inputs(self.P, self.D, self.M, self.W, self.H, self.tanfovx, self.tanfovy, self.scale_modifier,
                          False, False, self.bg, i["means3D"], i["sh"], None, i["opacity"], i["scales"],
                          i["rotations"], None, self.viewmatrix, self.projmatrix, self.campos,
                          raw_params=self.raw_params)

    def _launch_forward(self):
        s = torch.cuda.current_stream(self.dev).cuda_stream
        _lib.check(_lib.lib.gsr_forward_capacity(ctypes.byref(self._args()), self._geom.data_ptr(),
                                                 self._binning.data_ptr(), self.capacity, self._image.data_ptr(),
                                                 self.radii.data_ptr(), self._host.data_ptr(),
                                                 self.out_color.data_ptr(), s))

    def _launch_backward(self):
        s = torch.cuda.current_stream(self.dev).cuda_stream
        g = self.grads
        _lib.check(_lib.lib.gsr_backward(ctypes.byref(self._args()), self.radii.data_ptr(), self._geom.data_ptr(),
                                         self._binning.data_ptr(), self._image.data_ptr(), self.capacity,
                                         self.dL_dout.data_ptr(), g["means3D"].data_ptr(), g["means2D"].data_ptr(),
                                         g["sh"].data_ptr() if self.M else 0, g["colors"].data_ptr(),
                                         g["opacity"].data_ptr(), g["scales"].data_ptr(), g["rot"].data_ptr(),
                                         g["cov3D"].data_ptr(), g["conic"].data_ptr(), self._work.data_ptr(), s))

    def _measure(self):
        """num_rendered of the current static inputs through the ordinary entry point (one synchronising call)."""
        i = self.inp
        R = _C.rasterize_gaussians(self.bg, i["means3D"], None, i["opacity"], i["scales"], i["rotations"],
                                   self.scale_modifier, None, self.viewmatrix, self.projmatrix, self.tanfovx,
                                   self.tanfovy, self.H, self.W, i["sh"], self.D, self.campos, False, False,
                                   raw_params=self.raw_params)[0]
        return int(R)

    def capture(self, capacity=None):
        """(Re)build both graphs.  capacity=None: 1.5 x the num_rendered of the current static inputs + 64 K."""
        with torch.cuda.device(self.dev):
            if capacity is None and self.capacity == 0:
                R = self._measure()
                capacity = R + R // 2 + 65536
            if capacity is not None:
                self.capacity = int(min(capacity, 0x7FFFFFFF))
            self._fwd_graph = self._bwd_graph = None
            self._binning = torch.empty(_lib.scratch_bytes(self.P, self.capacity, self.W, self.H)[1],
                                        dtype=torch.uint8, device=self.dev)
            self._launch_forward()   # eager once: one-time kernel attributes are set outside the capture
            self._launch_backward()
            torch.cuda.synchronize(self.dev)
            fg, bg_ = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
            with torch.cuda.graph(fg):
                self._launch_forward()
            with torch.cuda.graph(bg_):
                self._launch_backward()
            self._fwd_graph, self._bwd_graph = fg, bg_
        return self

    # ---- replay ------------------------------------------------------------------------------------------------
    def fits(self):
        """After a synchronisation that covers the last forward: did its lists fit?  -> (ok, num_rendered)."""
        R, flags = int(self._host[0]), int(self._host[1])
        if flags & 1:
            raise RuntimeError("gsrast: point is filtered although prefiltered is set. This shouldn't happen!")
        return R <= self.capacity, R

    def forward(self, check=True):
        """Replay the forward on the current stream -> (out_color, radii): the STATIC outputs, overwritten by the
        next replay.  check=True waits for num_rendered and repeats the step with a larger buffer if needed."""
        with torch.cuda.device(self.dev):
            if self._fwd_graph is None:
                self.capture()
            self._fwd_graph.replay()
            if check:
                self._done.record(torch.cuda.current_stream(self.dev))
                self._done.synchronize()
                ok, R = self.fits()
                if not ok:
                    self.recaptures += 1
                    self.capture(capacity=2 * R + 65536)
                    self._fwd_graph.replay()
        return self.out_color, self.radii

    def backward(self, dL_dout=None):
        """Replay the backward of the last forward -> dict of the STATIC gradient tensors (means3D, means2D, sh,
        colors, opacity, scales, rot, cov3D, conic)."""
        if self._bwd_graph is None:
            raise RuntimeError("CapturedRasterizer.backward() before forward()")
        with torch.cuda.device(self.dev):
            if dL_dout is not None and dL_dout.data_ptr() != self.dL_dout.data_ptr():
                self.dL_dout.copy_(dL_dout)
            self._bwd_graph.replay()
        return self.grads

    def rasterize(self, means3D, sh, opacity, scales, rotations, check=True):
        """Autograd-aware call: -> (color (3,H,W), radii (P,)); copies of the static outputs."""
        return _CapturedFn.apply(self, check, means3D, sh, opacity, scales, rotations)


class _CapturedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, cr, check, means3D, sh, opacity, scales, rotations):
        cr.load(means3D, sh, opacity, scales, rotations)
        color, radii = cr.forward(check=check)
        ctx.cr = cr
        radii = radii.clone()
        ctx.mark_non_differentiable(radii)
        return color.clone(), radii

    @staticmethod
    def backward(ctx, dL_dcolor, _):
        g = ctx.cr.backward(dL_dcolor.contiguous())
        return (None, None, g["means3D"].clone(), g["sh"].clone(), g["opacity"].clone(), g["scales"].clone(),
                g["rot"].clone())
