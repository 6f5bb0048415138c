"""GPU parity of simple_knn.distCUDA2 (gsr_knn_dist2) against the UNMODIFIED reference simple_knn and the oracle:
bit-exact (the search is exact and the distance arithmetic is the reference's)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _ours(pts):
    from simple_knn._C import distCUDA2
    return distCUDA2(pts)


def _sets():
    rng = np.random.default_rng(1)
    yield "c1_tiramisu", np.load(os.path.join(HERE, "golden", "data", "c1_tiramisu_points.npz"))["xyz"]
    yield "uniform_200k", rng.uniform(-3, 3, (200_000, 3)).astype(np.float32)
    clusters = rng.normal(0, 1, (50, 3)) * 5
    yield "clustered_300k", (clusters[rng.integers(0, 50, 300_000)] + rng.normal(0, 0.05, (300_000, 3))).astype(
        np.float32)
    yield "planar_100k", np.concatenate([rng.uniform(-1, 1, (100_000, 2)), np.zeros((100_000, 1))], 1).astype(
        np.float32)
    dup = rng.normal(0, 1, (5000, 3)).astype(np.float32)
    yield "duplicates", np.concatenate([dup, dup[:2000], dup[:10]])
    yield "one_box_edge", rng.normal(0, 1, (257, 3)).astype(np.float32)


@pytest.mark.parametrize("name,pts", list(_sets()), ids=[n for n, _ in _sets()])
def test_knn_matches_reference_bit_exact(name, pts):
    from checkers import REF_KNN_SO, ref_knn
    if not os.path.exists(REF_KNN_SO):
        pytest.skip("oracle/_ref/libref_knn.so not built")
    t = torch.from_numpy(pts).cuda()
    ours, ref = _ours(t).cpu().numpy(), ref_knn(t).cpu().numpy()
    np.testing.assert_array_equal(ours.view(np.int32), ref.view(np.int32))


def test_knn_matches_oracle_and_edge_sizes(oracle):
    rng = np.random.default_rng(2)
    for P in (1, 2, 3, 4, 5, 31, 256, 1000, 4097):
        pts = rng.normal(0, 1, (P, 3)).astype(np.float32)
        ours = _ours(torch.from_numpy(pts).cuda()).cpu().numpy()
        np.testing.assert_array_equal(ours.view(np.int32), oracle.knn3(pts).view(np.int32), err_msg=str(P))
    assert _ours(torch.zeros(0, 3, device="cuda")).shape == (0,)
    with pytest.raises(RuntimeError):
        _ours(torch.zeros(5, 3))


def test_knn_million_points_properties():
    """Full size (1M points, the C2 scene's means): every value positive and finite, invariant under a permutation
    of the input (exact search), equal to the reference."""
    import gsr_scenes as S
    from checkers import REF_KNN_SO, ref_knn
    pts = S.scene_S(1_000_000, 64, 64, seed=0)["means3D"]
    t = torch.from_numpy(pts).cuda()
    a = _ours(t)
    assert torch.isfinite(a).all() and (a > 0).all()
    perm = torch.randperm(t.shape[0], device="cuda", generator=torch.Generator(device="cuda").manual_seed(0))
    b = _ours(t[perm])
    assert torch.equal(b, a[perm])
    if os.path.exists(REF_KNN_SO):
        assert torch.equal(a, ref_knn(t))
