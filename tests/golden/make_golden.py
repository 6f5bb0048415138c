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


def point_list_crc(pl):
    """Order-sensitive checksum of a point list: sum (i + 1) * (id + 1) mod 2^64."""
    pl = np.asarray(pl).astype(np.uint64)
    return np.uint64(((np.arange(1, pl.size + 1, dtype=np.uint64)) * (pl + np.uint64(1))).sum(dtype=np.uint64))


def digest(ref, name, out):
    """Cases too large for a full fixture (BASELINE config 1: 33,730 Gaussians at 800x450): integer per-Gaussian
    and per-tile results in full, the point list as an order-sensitive checksum, images every 4th pixel, float
    per-Gaussian arrays and gradients every 8th Gaussian.  Inputs are NOT stored: the case is rebuilt from
    tests/golden/data/c1_tiramisu_points.npz by cases.case()."""
    sc, opt = cases.case(name)
    d = to_device(sc)
    f = ref.forward(d)
    dpix = S.upstream_grad(sc["W"], sc["H"], seed=99)
    g = ref.backward(torch.from_numpy(dpix).cuda())
    rec = {}
    for k in ("radii", "tiles_touched", "ranges"):
        rec["ref_" + k] = f[k].cpu().numpy()
    rec["ref_num_rendered"] = np.int64(f["num_rendered"])
    rec["crc_point_list"] = point_list_crc(f["point_list"].cpu().numpy())
    for k in ("color", "n_contrib", "final_T"):
        rec["sub4_" + k] = f[k].cpu().numpy()[..., ::4, ::4].copy()
    vis = f["radii"].cpu().numpy() > 0
    for k in ("means2D", "depths", "conic_opacity", "rgb", "clamped"):
        a = f[k].cpu().numpy().copy()
        a[~vis] = 0
        rec["sub8_" + k] = a[::8].copy()
    for k, v in g.items():
        a = v.cpu().numpy()
        rec["sub8_grad_" + k] = a[::8].copy()
        rec["max_grad_" + k] = np.float32(np.abs(a).max()) if a.size else np.float32(0)
    path = os.path.join(out, name + "_digest.npz")
    np.savez_compressed(path, **rec)
    print(name, "digest: R =", f["num_rendered"], "visible =", int(vis.sum()), os.path.getsize(path), "bytes")


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out, exist_ok=True)
    ref = Reference()
    digest(ref, "c1_tiramisu", out)
    # the reference simple_knn on config 1's points (pins oracle.c orc_knn3, tests/test_cpu_knn.py)
    from checkers import ref_knn
    xyz = np.load(os.path.join(HERE, "data", "c1_tiramisu_points.npz"))["xyz"]
    np.save(os.path.join(out, "c1_knn_ref.npy"), ref_knn(torch.from_numpy(xyz).cuda()).cpu().numpy())
    if len(sys.argv) > 2 and sys.argv[2] == "digest-only":
        return
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
