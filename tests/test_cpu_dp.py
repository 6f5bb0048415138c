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


def _rebuild_numpy(means3D, campos, factors, degree, M):
    """numpy restatement of csrc/sh_exchange.cuh: sum_v basis(normalize(mean - campos_v)) (x) factors_v."""
    m, c, f = means3D.numpy().astype(np.float64), campos.numpy().astype(np.float64), factors.numpy()
    out = np.zeros((m.shape[0], M, 3))
    C1, C2, C3 = 0.4886025119029199, (1.0925484305920792, -1.0925484305920792, 0.31539156525252005,
                                       -1.0925484305920792, 0.5462742152960396), (
        -0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
        1.445305721320277, -0.5900435899266435)
    for v in range(c.shape[0]):
        d = m - c[v]
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        x, y, z = d[:, 0], d[:, 1], d[:, 2]
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        b = [np.full_like(x, S.SH_C0), -C1 * y, C1 * z, -C1 * x, C2[0] * xy, C2[1] * yz, C2[2] * (2 * zz - xx - yy),
             C2[3] * xz, C2[4] * (xx - yy), C3[0] * y * (3 * xx - yy), C3[1] * xy * z, C3[2] * y * (4 * zz - xx - yy),
             C3[3] * z * (2 * zz - 3 * xx - 3 * yy), C3[4] * x * (4 * zz - xx - yy), C3[5] * z * (xx - yy),
             C3[6] * x * (xx - 3 * yy)]
        K = (degree + 1) ** 2
        out[:, :K] += np.stack(b[:K], 1)[:, :, None] * f[v][:, None, :]
    return torch.from_numpy(out.astype(np.float32))


def _view_grads_factored(orc, v, exchange):
    """Small-parameter gradients; the view's SH factor goes to the exchange.  The factor (masked dL/dcolor) is
    recovered from the oracle's dL/dsh row 0, which is SH_C0 * factor (backward.cu:44)."""
    yaw = -15.0 + 30.0 * v / (N_VIEWS - 1)
    sc = S.scene_S(P, W, H, seed=11, M=M, sh_degree=DEG, yaw_deg=yaw)
    g = _view_grads(orc, v)
    exchange.emit(torch.from_numpy(sc["campos"].copy()), g["sh"][:, 0, :] / S.SH_C0)
    return {k: g[k] for k in gsr_dp.SMALL_LEAVES}


def _worker_factored(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from checkers import Oracle
        orc = Oracle()
        shapes = {k: v for k, v in _shapes().items() if k in gsr_dp.SMALL_LEAVES}
        flat = gsr_dp.FlatGrads(shapes, "cpu")
        means = torch.from_numpy(S.scene_S(P, W, H, seed=11, M=M, sh_degree=DEG)["means3D"])
        ex = gsr_dp.FactorExchange((N_VIEWS + world - 1) // world, P, "cpu", world)
        for step in range(2):  # the exchange buffers persist: the second step must not see the first one's
            out = gsr_dp.dp_step_factored(lambda v, e: _view_grads_factored(orc, v, e), N_VIEWS, flat, rank, world,
                                          means, DEG, M, rebuild=_rebuild_numpy, exchange=ex)
        np.save(os.path.join(out_dir, f"fact_{rank}.npy"),
                torch.cat([out[k].reshape(-1) for k in gsr_dp.LEAVES]).numpy())
    finally:
        dist.destroy_process_group()


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


def test_factor_exchange_world2_matches_plain_sum(tmp_path):
    """all-gather of 12 B/Gaussian/view SH factors + local rebuild == all-reduce of the full SH gradient."""
    from checkers import Oracle
    orc = Oracle()
    want = gsr_dp.FlatGrads(_shapes(), "cpu")
    gsr_dp.dp_step(lambda v: _view_grads(orc, v), N_VIEWS, want, 0, 1)
    mp.spawn(_worker_factored, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    ref = want.flat.numpy()
    for r in range(2):
        got = np.load(tmp_path / f"fact_{r}.npy")
        np.testing.assert_allclose(got, ref, rtol=2e-4, atol=1e-6 * np.abs(ref).max())


def test_factored_sum_without_a_process_group_matches_plain_sum():
    """world = 1, torch.distributed not initialised (tools/c3_bench.py on one GPU): the factors of the batch's views
    are kept locally and rebuilt once; same result as accumulating the full gradients."""
    from checkers import Oracle
    orc = Oracle()
    want = gsr_dp.FlatGrads(_shapes(), "cpu")
    gsr_dp.dp_step(lambda v: _view_grads(orc, v), N_VIEWS, want, 0, 1)
    flat = gsr_dp.FlatGrads({k: v for k, v in _shapes().items() if k in gsr_dp.SMALL_LEAVES}, "cpu")
    means = torch.from_numpy(S.scene_S(P, W, H, seed=11, M=M, sh_degree=DEG)["means3D"])
    assert not dist.is_initialized()
    for step in range(2):
        out = gsr_dp.dp_step_factored(lambda v, e: _view_grads_factored(orc, v, e), N_VIEWS, flat, 0, 1, means, DEG,
                                      M, rebuild=_rebuild_numpy)
    got = torch.cat([out[k].reshape(-1) for k in gsr_dp.LEAVES]).numpy()
    ref = want.flat.numpy()
    np.testing.assert_allclose(got, ref, rtol=2e-4, atol=1e-6 * np.abs(ref).max())


def test_rank_without_views_contributes_zeros():
    flat = gsr_dp.FlatGrads(_shapes(), "cpu")
    flat.flat.fill_(7.0)
    out = gsr_dp.dp_step(lambda v: pytest.fail("no view expected"), 0, flat, 0, 1)
    assert float(out["means3D"].abs().sum()) == 0.0
