"""compute-sanitizer driver: one forward + backward of a few small parity cases through the
C ABI, so that `compute-sanitizer --tool memcheck|racecheck|initcheck python tools/sanitize_cases.py` checks every
kernel of the path, including the shared-memory cursor tables of binning.cuh and the leader updates of the sort.
usage: python tools/sanitize_cases.py [case ...]   (default: s_deg3 s_huge_lists one_huge c1_tiramisu)"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "gaussian-transformer_b200"), os.path.join(ROOT, "tests")]
import cases  # noqa: E402
import gsr_scenes as S  # noqa: E402
from util import ours_backward, ours_forward, to_device  # noqa: E402

names = sys.argv[1:] or ["s_deg3", "s_huge_lists", "one_huge", "c1_tiramisu"]
for name in names:
    sc, opt = cases.case(name)
    colors, cov = cases.extras(sc, opt)
    d = to_device(sc)
    cp = torch.from_numpy(colors).cuda() if colors is not None else None
    cv = torch.from_numpy(cov).cuda() if cov is not None else None
    f = ours_forward(d, colors_precomp=cp, cov3D_precomp=cv)
    dpix = torch.from_numpy(S.upstream_grad(sc["W"], sc["H"], seed=99)).cuda()
    g = ours_backward(d, f, dpix, colors_precomp=cp, cov3D_precomp=cv)
    torch.cuda.synchronize()
    print("%s: P=%d R=%d |dL_dmeans3D|=%.4e" % (name, sc["P"], f["num_rendered"], float(g["means3D"].abs().sum())),
          flush=True)
