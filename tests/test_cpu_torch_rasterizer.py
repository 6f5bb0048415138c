"""The pure-PyTorch CPU rasterizer (oracle/torch_cpu_rasterizer.py: BASELINE.json's second baseline, the thing
bench.py times on the host cores for config C1) against the C oracle, which is pinned on the reference's outputs
(tests/test_cpu_oracle_golden.py).  Float32 torch ops round differently from the CUDA-shaped oracle, so integers may
differ on a handful of threshold cases and values are compared with tolerances."""
import os
import sys

import numpy as np
import pytest
import torch

import cases
import gsr_scenes as S

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import torch_cpu_rasterizer as T  # noqa: E402


@pytest.mark.parametrize("name", ["s_deg3", "s_deg1_m4", "s_bg_yaw", "s_scale_mod", "s_dense_opaque"])
def test_forward_backward_match_the_c_oracle(name, oracle):
    sc, opt = cases.case(name)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    f = oracle.forward(sc)
    dpix = S.upstream_grad(sc["W"], sc["H"], seed=99)
    g = oracle.backward(dpix)
    color, radii, grads, info = T.fwd_bwd(sc, dpix)
    assert (radii.numpy() != f["radii"]).mean() <= 2e-3
    assert abs(info["num_rendered"] - f["num_rendered"]) <= 0.01 * f["num_rendered"] + 4
    diff = np.abs(color.numpy() - f["color"])
    assert (diff > 1e-4).mean() <= 2e-3 and diff.max() < 5e-2, (diff.max(), (diff > 1e-4).mean())
    assert (info["n_contrib"].numpy() != f["n_contrib"].astype(np.int32)).mean() <= 5e-3
    for got, key in zip(grads, ("means3D", "sh", "opacity", "scales", "rot")):
        want = g[key].reshape(got.shape)
        err = np.abs(got.numpy() - want).max() / max(np.abs(want).max(), 1e-30)
        assert err <= 5e-3, (key, err)


def test_all_culled_gives_background():
    sc, _ = cases.case("all_culled")
    sc["bg"] = np.array([0.2, 0.4, 0.6], np.float32)
    leaves, kw = T.scene_tensors(sc)
    color, radii, info = T.rasterize_gaussians_torch(*leaves, **kw)
    assert int(radii.max()) == 0 and info["num_rendered"] == 0
    assert torch.allclose(color, torch.tensor(sc["bg"]).view(3, 1, 1).expand_as(color))
