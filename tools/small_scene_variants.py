"""dev tool: small scenes, captured forward + backward per iteration for the one-pixel-per-thread and two-pixel blend kernels and
the backward's PAIR instantiation (gsr_set_option render_fwd / render_bwd), plus the longest tile list of each scene.
usage: python tools/small_scene_variants.py > profiles/rNN_small_scene_variants.jsonl"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "gaussian-transformer_b200"), os.path.join(ROOT, "tests")]
import gsr_scenes as S  # noqa: E402
from diff_gaussian_rasterization import _lib  # noqa: E402
from gsr_captured import CapturedRasterizer  # noqa: E402
from util import ours_forward, to_device  # noqa: E402

SCENES = {
    "c1_tiramisu_800x450": lambda: S.scene_C1(os.path.join(ROOT, "tests", "golden", "data", "c1_tiramisu_points.npz")),
    "S_33730_800x450": lambda: S.scene_S(33_730, 800, 450, seed=0),
    "S_20000_640x360_big": lambda: S.scene_S(20_000, 640, 360, seed=5, log_scale_mean=np.log(0.03)),
    "S_200000_1280x720": lambda: S.scene_S(200_000, 1280, 720, seed=1),
    "S_100000_1024x576": lambda: S.scene_S(100_000, 1024, 576, seed=2),
    "S_500000_1920x1080": lambda: S.scene_S(500_000, 1920, 1080, seed=3),
}


def timed(fn, n=200):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


for name, make in SCENES.items():
    d = to_device(make())
    fwd = ours_forward(d)
    tc = fwd["tile_count"].numpy()
    dpix = torch.from_numpy(S.upstream_grad(d["W"], d["H"], 1)).cuda()
    row = {"scene": name, "P": int(d["P"]), "tiles": int(tc.size), "num_rendered": int(fwd["num_rendered"]),
           "longest_list": int(tc.max()), "mean_list": float(tc.mean()), "us": {}}
    # 1 = one pixel per thread, 2 = two pixels; backward 5 = one list entry per step, 4 = two (PAIR)
    for vf, vb in ((1, 1), (2, 5), (2, 4)):
        _lib.set_option("render_fwd", vf)
        _lib.set_option("render_bwd", vb)
        cr = CapturedRasterizer(d["P"], d["M"], d["sh_degree"], d["W"], d["H"], d["tanfovx"], d["tanfovy"], d["bg"])
        cr.set_camera(d["viewmatrix"], d["projmatrix"], d["campos"])
        cr.load(d["means3D"], d["sh"], d["opacities"], d["scales"], d["rotations"])
        cr.capture()
        cr.dL_dout.copy_(dpix)

        def fb():
            cr.forward(check=False)
            cr.backward()

        def f():
            cr.forward(check=False)

        row["us"]["fwd%d_bwd%d" % (vf, vb)] = {"fwd+bwd": round(timed(fb), 1), "fwd": round(timed(f), 1)}
    _lib.set_option("render_fwd", 2)
    _lib.set_option("render_bwd", 2)
    print(json.dumps(row), flush=True)
