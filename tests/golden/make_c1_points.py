"""Extracts the point cloud of BASELINE config 1 (the bundled tiramisu_ds COLMAP sparse points) into a small
fixture, so that the case can run where /root/reference does not exist (the GPU box).

Run in the build container:  python tests/golden/make_c1_points.py
Reads  /root/reference/tiramisu_ds/sparse/0/points3D.ply  (binary little-endian: x y z nx ny nz float32, r g b u8)
Writes tests/golden/data/c1_tiramisu_points.npz: xyz float32 (P,3), rgb uint8 (P,3), dist2 float32 (P,) = mean
squared distance to the 3 nearest neighbours, which is what simple_knn's distCUDA2 returns
(submodules/simple-knn/simple_knn.cu:185-221), computed here with an exact k-d tree in float64.
"""
import os

import numpy as np
from scipy.spatial import cKDTree

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/tiramisu_ds/sparse/0/points3D.ply"


def read_ply(path):
    raw = open(path, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    n = int([l for l in head.decode().splitlines() if l.startswith("element vertex")][0].split()[-1])
    rec = np.dtype([("xyz", "<f4", 3), ("n", "<f4", 3), ("rgb", "u1", 3)])
    a = np.frombuffer(body, dtype=rec, count=n)
    return a["xyz"].copy(), a["rgb"].copy()


def main():
    xyz, rgb = read_ply(SRC)
    d, _ = cKDTree(xyz.astype(np.float64)).query(xyz.astype(np.float64), k=4)
    dist2 = (d[:, 1:] ** 2).mean(1).astype(np.float32)
    out = os.path.join(HERE, "data", "c1_tiramisu_points.npz")
    np.savez_compressed(out, xyz=xyz, rgb=rgb, dist2=dist2)
    print(xyz.shape, "z range", xyz[:, 2].min(), xyz[:, 2].max(), os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
