"""dev tool for ncu: config C1 forward + backward once per backward instantiation (one list entry per step, then PAIR), eager launches.
ncu --set full --clock-control none --import-source on -k regex:render_.wd2 -c 4 -o OUT python tools/small_scene_profile.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "gaussian-transformer_b200"), os.path.join(ROOT, "tests")]
import gsr_scenes as S  # noqa: E402
from diff_gaussian_rasterization import _lib  # noqa: E402
from util import ours_backward, ours_forward, to_device  # noqa: E402

d = to_device(S.scene_C1(os.path.join(ROOT, "tests", "golden", "data", "c1_tiramisu_points.npz")))
dpix = torch.from_numpy(S.upstream_grad(d["W"], d["H"], 1)).cuda()
for vf, vb in ((2, 5), (2, 4)):
    _lib.set_option("render_fwd", vf)
    _lib.set_option("render_bwd", vb)
    ours_backward(d, ours_forward(d), dpix)
torch.cuda.synchronize()
