"""CPU tests of the simple_knn oracle (oracle.c orc_knn3): against an independent exact k-d tree and against the
output of the UNMODIFIED reference simple_knn on a B200 for BASELINE config 1's point cloud
(tests/golden/data/c1_knn_ref.npy, written by tests/golden/make_golden.py)."""
import os

import numpy as np
from scipy.spatial import cKDTree

HERE = os.path.dirname(os.path.abspath(__file__))
DATA = os.path.join(HERE, "golden", "data")


def _tree_knn3(pts):
    d, _ = cKDTree(pts.astype(np.float64)).query(pts.astype(np.float64), k=4)
    return (d[:, 1:] ** 2).mean(1)


def test_oracle_knn_matches_kdtree(oracle):
    rng = np.random.default_rng(0)
    for P in (4, 5, 100, 3000):
        pts = rng.normal(0, 1, (P, 3)).astype(np.float32)
        np.testing.assert_allclose(oracle.knn3(pts), _tree_knn3(pts), rtol=2e-5)


def test_oracle_knn_duplicates_and_tiny_sets(oracle):
    pts = np.array([[0, 0, 0], [0, 0, 0], [1, 0, 0], [0, 2, 0], [0, 0, 0]], dtype=np.float32)
    got = oracle.knn3(pts)
    np.testing.assert_allclose(got, [(0 + 0 + 1) / 3, (0 + 0 + 1) / 3, (1 + 1 + 1) / 3, (4 + 4 + 4) / 3, 1 / 3],
                               rtol=1e-6)
    # fewer than 4 points: best slots stay FLT_MAX (simple_knn.cu:155,181): one slot -> FLT_MAX / 3, two -> overflow
    np.testing.assert_allclose(oracle.knn3(pts[:3]), np.float32(3.4028235e38) / np.float32(3), rtol=1e-6)
    assert np.isinf(oracle.knn3(pts[:2])).all()


def test_oracle_knn_matches_reference_output_c1(oracle):
    z = np.load(os.path.join(DATA, "c1_tiramisu_points.npz"))
    ref = np.load(os.path.join(DATA, "c1_knn_ref.npy"))
    got = oracle.knn3(z["xyz"])
    np.testing.assert_array_equal(got.view(np.int32), ref.view(np.int32))  # bit-exact
    np.testing.assert_allclose(got, z["dist2"], rtol=2e-5)
