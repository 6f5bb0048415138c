"""Host logic of the data-parallel path (gsr_dp.py), world_size 2 over gloo on CPU.

The per-view renderer here is the CPU oracle (test infrastructure standing in for the CUDA rasterizer, which has no
CPU path): what is under test is the view sharding, the flat gradient buffer and the single all-reduce.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import gsr_dp
import gsr_scenes as S

P, W, H, M, DEG, N_VIEWS = 300, 96, 64, 16, 3, 5
KEYMAP = {"means3D": "means3D", "sh": "sh", "opacities": "opacity", "scales": "scales", "rotations": "rot"}


def _view_grads(orc, v):
    yaw = -15.0 + 30.0 * v / (N_VIEWS - 1)
    sc = S.scene_S(P, W, H, seed=11, M=M, sh_degree=DEG, yaw_deg=yaw)
    orc.forward(sc, internals=False)
    g = orc.backward(S.upstream_grad(W, H, seed=100 + v))
    return {k: torch.from_numpy(np.ascontiguousarray(g[o])) for k, o in KEYMAP.items()}


def _shapes():
    return {"means3D": (P, 3), "sh": (P, M, 3), "opacities": (P, 1), "scales": (P, 3), "rotations": (P, 4)}


def _worker(rank, world, port, bucket_bytes, out_dir):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from checkers import Oracle
        orc = Oracle()
        flat = gsr_dp.FlatGrads(_shapes(), "cpu", bucket_bytes=bucket_bytes)
        for step in range(2):  # the second step checks that the buffer is overwritten, not accumulated
            views = gsr_dp.dp_step(lambda v: _view_grads(orc, v), N_VIEWS, flat, rank, world)
        np.save(os.path.join(out_dir, f"flat_{rank}.npy"), flat.flat.numpy())
        assert views["sh"].shape == (P, M, 3)
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_shard_views_partition():
    for n in (0, 1, 5, 8, 13):
        for world in (1, 2, 3, 4, 8):
            blocks = [gsr_dp.shard_views(n, r, world) for r in range(world)]
            assert sum(blocks, []) == list(range(n))
            assert max(map(len, blocks)) - min(map(len, blocks)) <= 1
    with pytest.raises(ValueError):
        gsr_dp.shard_views(4, 2, 2)


@pytest.mark.parametrize("bucket_bytes", [0, 16384])
def test_dp_step_world2_matches_single_process(tmp_path, bucket_bytes):
    from checkers import Oracle
    orc = Oracle()
    want = gsr_dp.FlatGrads(_shapes(), "cpu")
    gsr_dp.dp_step(lambda v: _view_grads(orc, v), N_VIEWS, want, 0, 1)
    assert float(want.flat.abs().sum()) > 0
    mp.spawn(_worker, args=(2, _free_port(), bucket_bytes, str(tmp_path)), nprocs=2, join=True)
    for r in range(2):
        got = np.load(tmp_path / f"flat_{r}.npy")
        # same terms, different association (3 + 2 views vs 5): fp32 rounding only
        np.testing.assert_allclose(got, want.flat.numpy(), rtol=1e-5, atol=1e-9)


def test_rank_without_views_contributes_zeros():
    flat = gsr_dp.FlatGrads(_shapes(), "cpu")
    flat.flat.fill_(7.0)
    out = gsr_dp.dp_step(lambda v: pytest.fail("no view expected"), 0, flat, 0, 1)
    assert float(out["means3D"].abs().sum()) == 0.0
