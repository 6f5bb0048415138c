"""BASELINE config 5: scene S with 6 M Gaussians at 3840x2160, forward only, a sweep of 16 camera
yaw angles around the cloud.  Per angle one JSON line: visible Gaussians, tile instances R, per-tile list lengths
(mean / p50 / p99 / max and a power-of-two histogram), forward time (CUDA events, median), the per-stage device times
of the library's profile hooks with the depth sort's and the binning's share, and the unmodified reference's forward
time on the same inputs (oracle/_ref, as the checker/baseline; skipped with --no-reference).

python tools/sweep_c5.py [P W H] [--yaw-max 45] [--angles 16] [--no-reference]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "gaussian-transformer_b200"), os.path.join(ROOT, "tests")]
import gsr_scenes as S  # noqa: E402
from diff_gaussian_rasterization import _C, _lib  # noqa: E402
from util import ours_forward, to_device  # noqa: E402


def median_ms(fn, n, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("size", nargs="*", type=int, default=[6_000_000, 3840, 2160])
    ap.add_argument("--yaw-max", type=float, default=45.0)
    ap.add_argument("--angles", type=int, default=16)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--no-reference", action="store_true")
    args = ap.parse_args()
    P, W, H = args.size
    sc = S.scene_S(P, W, H, seed=0)
    d = to_device(sc)
    ref = None
    if not args.no_reference:
        from checkers import Reference
        ref = Reference()
    e = torch.Tensor([])
    for yaw in np.linspace(-args.yaw_max, args.yaw_max, args.angles):
        cam = S.look_at_camera(W, H, 60.0, float(yaw))
        for k in ("viewmatrix", "projmatrix", "campos"):
            d[k] = torch.from_numpy(cam[k]).cuda()

        def fwd():
            return _C.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                          d["viewmatrix"], d["projmatrix"], d["tanfovx"], d["tanfovy"], H, W, d["sh"],
                                          3, d["campos"], False, False)
        f = ours_forward(d)
        r = f["ranges"].numpy().astype(np.int64)
        ln = r[:, 1] - r[:, 0]
        hist = np.bincount(np.where(ln > 0, np.floor(np.log2(np.maximum(ln, 1))).astype(np.int64) + 1, 0))
        row = {"yaw_deg": round(float(yaw), 2), "P": P, "W": W, "H": H, "visible": int((f["radii"] > 0).sum()),
               "num_rendered": int(f["num_rendered"]),
               "list_len": {"mean": round(float(ln.mean()), 1), "p50": int(np.percentile(ln, 50)),
                            "p99": int(np.percentile(ln, 99)), "max": int(ln.max()),
                            "hist_pow2": {("0" if i == 0 else "<%d" % (1 << i)): int(c) for i, c in enumerate(hist) if c}},
               "sum_n_contrib": int(f["n_contrib"].sum())}
        del f
        row["fwd_ms"] = round(median_ms(fwd, args.iters), 3)
        _lib.profile_enable(True)
        _lib.profile_read()
        for _ in range(args.iters):
            fwd()
        torch.cuda.synchronize()
        prof = _lib.profile_read()
        _lib.profile_enable(False)
        stages = {k: round(v[0] / args.iters, 3) for k, v in prof.items() if v[1]}
        tot = sum(stages.values())
        row["stage_ms"] = stages
        row["sort_share"] = round(sum(v for k, v in stages.items() if "sort" in k) / tot, 3)
        row["bin_share"] = round(sum(v for k, v in stages.items() if k.startswith("bin")) / tot, 3)
        if ref is not None:
            row["reference_fwd_ms"] = round(median_ms(lambda: ref.forward(d, internals=False, sync=False), 3, 1), 3)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
