"""dev tool: config C1 through the autograd module (the call bench.py's cpu_baseline_pytorch times next to the PyTorch CPU
implementation), with the backward's PAIR instantiation on / off, before and after the CPU implementation has spun up the
host's OpenMP threads."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "gaussian-transformer_b200"), os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]
import gsr_scenes as S  # noqa: E402
from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, _lib  # noqa: E402

TENSORS = ("means3D", "scales", "rotations", "opacities", "sh", "bg", "viewmatrix", "projmatrix", "campos")
LEAVES = ("means3D", "sh", "opacities", "scales", "rotations")
sc = S.scene_C1(os.path.join(ROOT, "tests", "golden", "data", "c1_tiramisu_points.npz"))
dpix_np = S.upstream_grad(sc["W"], sc["H"], 1)
t = {k: torch.from_numpy(np.ascontiguousarray(sc[k])).cuda() for k in TENSORS}
dpix = torch.from_numpy(dpix_np).cuda()
raster = GaussianRasterizer(GaussianRasterizationSettings(
    sc["H"], sc["W"], sc["tanfovx"], sc["tanfovy"], t["bg"], 1.0, t["viewmatrix"], t["projmatrix"], sc["sh_degree"],
    t["campos"], False, False))


def one():
    leaves = {k: t[k].detach().requires_grad_(True) for k in LEAVES}
    color, _ = raster(means3D=leaves["means3D"], means2D=torch.zeros_like(leaves["means3D"]),
                      opacities=leaves["opacities"], shs=leaves["sh"], scales=leaves["scales"],
                      rotations=leaves["rotations"])
    color.backward(dpix)


def timed(n=100):
    for _ in range(5):
        one()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0 = time.time()
    a.record()
    for _ in range(n):
        one()
    b.record()
    torch.cuda.synchronize()
    return {"device_us": round(a.elapsed_time(b) / n * 1e3, 1), "wall_us": round((time.time() - w0) / n * 1e6, 1)}


out = {}
for label, v in (("pair_auto", 2), ("pair_off", 5), ("pair_auto_again", 2)):
    _lib.set_option("render_bwd", v)
    out[label] = timed()
_lib.set_option("render_bwd", 2)
_lib.profile_enable(True)
out["profiling_on"] = timed()
_lib.profile_read()
_lib.profile_enable(False)
import torch_cpu_rasterizer as T  # noqa: E402
torch.set_num_threads(os.cpu_count() or 1)
t0 = time.time()
T.fwd_bwd(sc, dpix_np)
out["cpu_pytorch_s"] = round(time.time() - t0, 2)
out["after_cpu_baseline"] = timed()
time.sleep(1.0)
out["after_cpu_baseline_1s_later"] = timed()
print(json.dumps(out))
