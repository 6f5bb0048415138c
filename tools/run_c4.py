"""BASELINE config 4: the reference's `train.py`, UNMODIFIED, on its bundled `table_ds` points, run on top of this
repo's drop-ins (`diff_gaussian_rasterization`, `simple_knn`, `plyfile`, `utils.loss_utils`) or on the reference's own
extensions (`--stack reference`).

The reference tree is not part of this repo: point `--ref-root` at a checkout of it (its `train.py`, `scene/`, `utils/`,
`gaussian_renderer/`, `arguments/`, `table_ds/`).  For a GPU-box run from the build container an unmodified, git-ignored
copy goes to `baseline/_ref/gaussian-transformer` (the default).

`table_ds` ships `cameras.bin`, `points3D.ply` and 19 JPEGs but no `images.bin` (no poses), so the harness writes one
(SURVEY.md §8d, C4): one camera per JPEG on an orbit around the cloud's median looking at it, in COLMAP's binary layout
(`scene/colmap_loader.py:180-213`: u64 count; per image `<i 4d 3d i>`, the NUL-terminated name, u64 0 observations).
The bundled photographs are the targets — inconsistent with synthetic poses, so the run is a PERFORMANCE run (loss,
gradients, densify/prune/opacity-reset all execute as usual; image quality is meaningless).

Instrumentation lives here, not in `train.py`: `GaussianModel.training_setup` is wrapped to take the loop's start time.
Prints one JSON line: iterations, seconds of the training loop, iters/s, final number of Gaussians (read back from the
saved `point_cloud.ply` with whichever `plyfile` the stack uses).
"""
import argparse
import json
import math
import os
import runpy
import shutil
import struct
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "gaussian-transformer_b200")


def rotmat_to_qvec(R):
    """(w, x, y, z) of a rotation matrix, COLMAP's convention (inverse of `scene/colmap_loader.py:qvec2rotmat`)."""
    K = np.array([[R[0, 0] - R[1, 1] - R[2, 2], 0, 0, 0],
                  [R[1, 0] + R[0, 1], R[1, 1] - R[0, 0] - R[2, 2], 0, 0],
                  [R[2, 0] + R[0, 2], R[2, 1] + R[1, 2], R[2, 2] - R[0, 0] - R[1, 1], 0],
                  [R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1], R[0, 0] + R[1, 1] + R[2, 2]]]) / 3.0
    vals, vecs = np.linalg.eigh(K)
    q = vecs[[3, 0, 1, 2], np.argmax(vals)]
    return -q if q[0] < 0 else q


def orbit_poses(n, centre, radius, tilt=0.2):
    """World-to-camera (R, t) of n cameras on a circle around `centre` (COLMAP axes: x right, y down, z forward)."""
    poses = []
    for i in range(n):
        a = 2.0 * math.pi * i / n
        C = centre + radius * np.array([math.sin(a), -tilt, -math.cos(a)])
        f = centre - C
        f /= np.linalg.norm(f)
        right = np.cross(np.array([0.0, 1.0, 0.0]), f)
        right /= np.linalg.norm(right)
        down = np.cross(f, right)
        R = np.stack([right, down, f])
        poses.append((R, -R @ C))
    return poses


def write_images_bin(path, names, poses, camera_id=1):
    with open(path, "wb") as fid:
        fid.write(struct.pack("<Q", len(names)))
        for i, (name, (R, t)) in enumerate(zip(names, poses)):
            fid.write(struct.pack("<idddddddi", i + 1, *rotmat_to_qvec(R), *t, camera_id))
            fid.write(name.encode("utf-8") + b"\x00")
            fid.write(struct.pack("<Q", 0))


def make_dataset(ref_root, work, read_ply):
    src = os.path.join(ref_root, "table_ds")
    ds = os.path.join(work, "table_ds")
    shutil.rmtree(ds, ignore_errors=True)
    os.makedirs(os.path.join(ds, "sparse", "0"))
    for f in ("cameras.bin", "points3D.ply"):
        shutil.copy(os.path.join(src, "sparse", "0", f), os.path.join(ds, "sparse", "0", f))
    os.symlink(os.path.join(src, "images"), os.path.join(ds, "images"))
    names = sorted(os.listdir(os.path.join(src, "images")))
    v = read_ply(os.path.join(ds, "sparse", "0", "points3D.ply"))["vertex"]
    xyz = np.stack([v["x"], v["y"], v["z"]], 1).astype(np.float64)
    centre = np.median(xyz, 0)
    radius = 1.5 * float(np.percentile(np.linalg.norm(xyz - centre, axis=1), 90))
    write_images_bin(os.path.join(ds, "sparse", "0", "images.bin"), names, orbit_poses(len(names), centre, radius))
    return ds, len(names), len(xyz)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref-root", default=os.path.join(ROOT, "baseline", "_ref", "gaussian-transformer"))
    ap.add_argument("--stack", choices=["b200", "reference"], default="b200")
    ap.add_argument("--iterations", type=int, default=3000)
    ap.add_argument("--work", default=os.path.join(ROOT, "gpurun_out", "c4"))
    ap.add_argument("--fused", action="store_true",
                    help="b200 stack + the optional trainer-side changes of INTEGRATION.md sections 6-9, applied by "
                         "patching the imported modules (train.py itself stays byte-identical): FusedAdam, "
                         "gsr_render.render (raw parameters + split SH), gsr_densify, fused densification statistics")
    ap.add_argument("--dataset-only", action="store_true", help="write the dataset (CPU) and stop")
    ap.add_argument("--profile", default=None, help="dev: run under torch.profiler and write the kernel / CPU-op tables "
                                                    "to this file (the timing of such a run is not a result)")
    args = ap.parse_args()
    ref_root = os.path.abspath(args.ref_root)
    if args.profile:
        args.profile = os.path.abspath(args.profile)
    if not os.path.isfile(os.path.join(ref_root, "train.py")):
        sys.exit("no train.py under --ref-root %s" % ref_root)

    if args.stack == "b200":
        sys.path.insert(0, PKG)                       # every drop-in: rasterizer, simple_knn, plyfile, utils.loss_utils
    else:
        # the reference's own extensions (pip --target baseline/_ref) + only the plyfile stand-in from this repo
        only_ply = os.path.join(args.work, "_plyfile_only")
        os.makedirs(only_ply, exist_ok=True)
        if not os.path.exists(os.path.join(only_ply, "plyfile")):
            os.symlink(os.path.join(PKG, "plyfile"), os.path.join(only_ply, "plyfile"))
        sys.path[:0] = [os.path.join(ROOT, "baseline", "_ref"), only_ply]
    sys.path.insert(0, ref_root)                      # what `python train.py` puts first
    os.makedirs(args.work, exist_ok=True)
    from plyfile import PlyData
    ds, n_cams, n_points = make_dataset(ref_root, os.path.abspath(args.work), PlyData.read)
    if args.dataset_only:
        print(json.dumps({"dataset": ds, "cameras": n_cams, "points": n_points}))
        return

    import torch
    import scene.gaussian_model as gm
    import diff_gaussian_rasterization as dgr
    import simple_knn._C as knn
    import utils.loss_utils as lu
    marks = {}
    setup = gm.GaussianModel.training_setup

    def timed_setup(self, opt):
        setup(self, opt)
        torch.cuda.synchronize()
        marks["t0"], marks["P0"] = time.perf_counter(), int(self.get_xyz.shape[0])
    gm.GaussianModel.training_setup = timed_setup
    hits = {"calls": 0, "misses": 0}
    if args.stack == "b200":
        # hit rate of the speculative forward (phase 2 launched before num_rendered reaches the host): a miss is a
        # call whose binning buffer had to be re-allocated, i.e. gsr_forward_post ran a second time
        from diff_gaussian_rasterization import _lib
        post = _lib.lib.gsr_forward_post
        fwd = _lib.lib.gsr_forward

        def count_post(*a):
            hits["misses"] += 1
            return post(*a)

        def count_fwd(*a):
            hits["calls"] += 1
            return fwd(*a)
        _lib.lib.gsr_forward_post, _lib.lib.gsr_forward = count_post, count_fwd
    if args.fused:
        if args.stack != "b200":
            sys.exit("--fused needs --stack b200")
        import gaussian_renderer
        import gsr_densify
        import gsr_optim
        import gsr_render
        inner = gm.GaussianModel.training_setup

        def fused_setup(self, opt):
            inner(self, opt)
            groups = [{"params": g["params"], "lr": g["lr"], "name": g["name"]} for g in self.optimizer.param_groups]
            self.optimizer = gsr_optim.FusedAdam(groups, lr=0.0, eps=1e-15)
        gm.GaussianModel.training_setup = fused_setup
        gaussian_renderer.render = gsr_render.render                                  # section 7
        gm.GaussianModel.densify_and_prune = gsr_densify.densify_and_prune            # section 9
        gm.GaussianModel.prune_points = gsr_densify.prune_points


    out = os.path.join(os.path.abspath(args.work), "out_" + args.stack + ("_fused" if args.fused else ""))
    shutil.rmtree(out, ignore_errors=True)
    sys.argv = ["train.py", "-s", ds, "-m", out, "--iterations", str(args.iterations), "--test_iterations", "-1",
                "--save_iterations", str(args.iterations), "--quiet"]
    os.chdir(ref_root)
    if args.profile:
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as tp:
            runpy.run_path(os.path.join(ref_root, "train.py"), run_name="__main__")
        ka = tp.key_averages()
        with open(args.profile, "w") as f:
            f.write("# %s stack, %d iterations of train.py under torch.profiler\n" % (args.stack, args.iterations))
            f.write(ka.table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=70))
            f.write("\n")
            f.write(ka.table(sort_by="self_cpu_time_total", row_limit=30, max_name_column_width=70))
    else:
        runpy.run_path(os.path.join(ref_root, "train.py"), run_name="__main__")
    torch.cuda.synchronize()
    sec = time.perf_counter() - marks["t0"]
    sys.stdout = sys.__stdout__                        # safe_state(--quiet) replaced it
    ply = os.path.join(out, "point_cloud", "iteration_%d" % args.iterations, "point_cloud.ply")
    final = PlyData.read(ply)["vertex"]
    print(json.dumps({
        "workload": "C4: reference train.py unmodified, table_ds points + synthetic orbit poses, default 1.6K rescale",
        "stack": args.stack + (" + FusedAdam, gsr_render (raw parameters, split SH), gsr_densify" if args.fused else ""),
        "speculative_forward": ({"calls": hits["calls"], "misses": hits["misses"],
                                 "hit_rate": round(1.0 - hits["misses"] / max(hits["calls"], 1), 4)}
                                if args.stack == "b200" else None),
        "iterations": args.iterations, "loop_s": round(sec, 3),
        "iters_per_s": round(args.iterations / sec, 2), "cameras": n_cams, "P_start": marks["P0"],
        "P_final": int(final.count), "ply_properties": len(final.properties),
        "modules": {"rasterizer": os.path.relpath(dgr.__file__, ROOT), "simple_knn": os.path.relpath(knn.__file__, ROOT),
                    "loss_utils": os.path.relpath(lu.__file__, ROOT)},
        "device": torch.cuda.get_device_name(0)}))


if __name__ == "__main__":
    main()
