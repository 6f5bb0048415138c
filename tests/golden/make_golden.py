"""Generates tests/golden/*.npz by running the UNMODIFIED reference CUDA rasterizer (oracle/_ref) on a B200.

Run on the GPU box:  python tests/golden/make_golden.py [outdir]     (default: gpurun_out/golden)
then copy the files into tests/golden/ and commit them.  Each file holds the inputs and the reference's
outputs (forward, decoded internal buffers, backward) for one case of tests/cases.py.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(ROOT, "gaussian-transformer_b200"), os.path.join(ROOT, "tests")]

import cases  # noqa: E402
import gsr_scenes as S  # noqa: E402
from checkers import Reference  # noqa: E402
from util import to_device  # noqa: E402


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out, exist_ok=True)
    ref = Reference()
    for name in cases.GOLDEN:
        sc, opt = cases.case(name)
        colors, cov = cases.extras(sc, opt)
        d = to_device(sc)
        cp = torch.from_numpy(colors).cuda() if colors is not None else None
        cv = torch.from_numpy(cov).cuda() if cov is not None else None
        f = ref.forward(d, colors_precomp=cp, cov3D_precomp=cv)
        dpix = S.upstream_grad(sc["W"], sc["H"], seed=99)
        g = ref.backward(torch.from_numpy(dpix).cuda())
        rec = {("in_" + k): np.asarray(v) for k, v in sc.items()}
        if colors is not None:
            rec["in_colors_precomp"] = colors
        if cov is not None:
            rec["in_cov3D_precomp"] = cov
        rec["in_dL_dpix"] = dpix
        vis = f["radii"].cpu().numpy() > 0
        for k in ("color", "radii", "tiles_touched", "ranges", "point_list", "n_contrib", "final_T"):
            rec["ref_" + k] = f[k].cpu().numpy()
        rec["ref_num_rendered"] = np.int64(f["num_rendered"])
        # per-Gaussian internals are only defined where radii > 0 (SURVEY App. B caveat)
        for k in ("means2D", "depths", "conic_opacity", "rgb", "clamped"):
            a = f[k].cpu().numpy().copy()
            if k == "rgb" and colors is not None:
                continue
            a[~vis] = 0
            rec["ref_" + k] = a
        for k, v in g.items():
            rec["ref_grad_" + k] = v.cpu().numpy()
        path = os.path.join(out, name + ".npz")
        np.savez_compressed(path, **rec)
        print(name, "R =", f["num_rendered"], "visible =", int(vis.sum()), os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
