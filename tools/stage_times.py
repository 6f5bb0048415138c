"""Per-stage device times of one forward + backward for an arbitrary scene size (dev tool):
python tools/stage_times.py P W H"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "gaussian-transformer_b200"), os.path.join(ROOT, "tests")]
import gsr_scenes as S  # noqa: E402
from diff_gaussian_rasterization import _C, _lib  # noqa: E402
from util import to_device  # noqa: E402

P, W, H = (int(a) for a in sys.argv[1:4])
sc = S.scene_S(P, W, H, seed=0)
d = to_device(sc)
dpix = torch.from_numpy(S.upstream_grad(W, H, 1)).cuda()
e = torch.Tensor([])


def fb():
    R, c, radii, g, b, i = _C.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"],
                                                  1.0, e, d["viewmatrix"], d["projmatrix"], d["tanfovx"], d["tanfovy"],
                                                  H, W, d["sh"], 3, d["campos"], False, False)
    _C.rasterize_gaussians_backward(d["bg"], d["means3D"], radii, e, d["scales"], d["rotations"], 1.0, e,
                                    d["viewmatrix"], d["projmatrix"], d["tanfovx"], d["tanfovy"], dpix, d["sh"], 3,
                                    d["campos"], g, R, b, i, False)
    return R


for _ in range(3):
    R = fb()
torch.cuda.synchronize()
_lib.profile_enable(True)
n = 5
for _ in range(n):
    fb()
torch.cuda.synchronize()
prof = _lib.profile_read()
_lib.profile_enable(False)
print(f"P={P} {W}x{H} R={R}: " + " ".join(f"{k}={v[0] / n * 1e3:.0f}us" for k, v in prof.items() if v[1]))
