"""Config C1 (33,730 COLMAP points as Gaussians, SH3, 800x450) forward + backward per iteration, device time over many
iterations (CUDA events around the whole loop, so host-side gaps count):
  entry     the `_C`-shaped entry points (allocation + the num_rendered round trip per call)
  captured  gsr_captured.CapturedRasterizer, sync-free replay (check=False)
  checked   the same with the per-step wait for num_rendered (check=True)
  autograd  CapturedRasterizer.rasterize(...) + .backward() on leaf tensors
usage: python tools/small_scene_bench.py [iters] > profiles/rNN_small_scene.json
       python tools/small_scene_bench.py --timeline FILE   (kernel timeline of 2 entry + 2 captured iterations)"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "gaussian-transformer_b200"), os.path.join(ROOT, "tests")]
import gsr_scenes as S  # noqa: E402
from diff_gaussian_rasterization import _C, _lib  # noqa: E402
from gsr_captured import CapturedRasterizer  # noqa: E402
from util import to_device  # noqa: E402

timeline = sys.argv[2] if len(sys.argv) > 2 and sys.argv[1] == "--timeline" else None
iters = int(sys.argv[1]) if len(sys.argv) > 1 and not timeline else 300
sc = S.scene_C1(os.path.join(ROOT, "tests", "golden", "data", "c1_tiramisu_points.npz"))
d = to_device(sc)
P, W, H = d["P"], d["W"], d["H"]
dpix = torch.from_numpy(S.upstream_grad(W, H, 1)).cuda()
e = torch.Tensor([])


def entry():
    R, c, radii, g, b, i = _C.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"],
                                                  1.0, e, d["viewmatrix"], d["projmatrix"], d["tanfovx"], d["tanfovy"],
                                                  H, W, d["sh"], 3, d["campos"], False, False)
    _C.rasterize_gaussians_backward(d["bg"], d["means3D"], radii, e, d["scales"], d["rotations"], 1.0, e,
                                    d["viewmatrix"], d["projmatrix"], d["tanfovx"], d["tanfovy"], dpix, d["sh"], 3,
                                    d["campos"], g, R, b, i, False)
    return R


cr = CapturedRasterizer(P, d["M"], 3, W, H, d["tanfovx"], d["tanfovy"], d["bg"])
cr.set_camera(d["viewmatrix"], d["projmatrix"], d["campos"])
cr.load(d["means3D"], d["sh"], d["opacities"], d["scales"], d["rotations"])
cr.capture()
cr.dL_dout.copy_(dpix)
leaves = [d[k].clone().requires_grad_(True) for k in ("means3D", "sh", "opacities", "scales", "rotations")]


def captured():
    cr.forward(check=False)
    cr.backward()


def checked():
    cr.forward(check=True)
    cr.backward()


def autograd():
    for t in leaves:
        t.grad = None
    color, _ = cr.rasterize(*leaves)
    color.backward(dpix)


def timed(fn, n):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = _lib.launch_count()
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3, (_lib.launch_count() - l0) / n


if timeline:
    from torch.profiler import ProfilerActivity, profile
    for fn in (entry, captured):
        for _ in range(5):
            fn()
    torch.cuda.synchronize()
    with open(timeline, "w") as f:
        for name, fn in (("entry", entry), ("captured", captured)):
            with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as tp:
                for _ in range(2):
                    fn()
                torch.cuda.synchronize()
            ev = sorted((x for x in tp.events() if x.device_type.name == "CUDA"), key=lambda x: x.time_range.start)
            t0 = ev[0].time_range.start
            f.write("# %s: start us, duration us, kernel\n" % name)
            for x in ev:
                f.write("%9.1f %8.1f  %s\n" % (x.time_range.start - t0, x.time_range.end - x.time_range.start,
                                                x.name[:90]))
    sys.exit(0)

out = {"config": "C1: %d Gaussians, SH3, %dx%d, fwd+bwd" % (P, W, H), "iters": iters, "num_rendered": entry(),
       "capacity": cr.capacity, "us_per_iter": {}, "library_launch_calls_per_iter": {}}
for name, fn in (("entry", entry), ("captured", captured), ("checked", checked), ("autograd", autograd)):
    us, launches = timed(fn, iters)
    out["us_per_iter"][name] = round(us, 1)
    out["library_launch_calls_per_iter"][name] = launches  # replays launch the captured kernels without the library
ok, R = cr.fits()
out["captured_fits"], out["recaptures"] = bool(ok), cr.recaptures

print(json.dumps(out))
