"""gsr_captured.CapturedRasterizer (forward + backward replayed as CUDA graphs over gsr_forward_capacity) against the
ordinary `_C`-shaped entry points on the same inputs: image, radii and num_rendered bit-identical, gradients within the
1e-3 relative bar the parity tests use (the blend's atomics sum in a different order from run to run)."""
import os

import numpy as np
import pytest
import torch

import gsr_scenes as S
from test_gpu_parity import _check_grads
from util import np_, ours_backward, ours_forward, rel_err, to_device

pytestmark = pytest.mark.gpu

GRAD_TOL = 1e-3
C1_POINTS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "data", "c1_tiramisu_points.npz")


def _scene(name):
    if name == "c1":
        return S.scene_C1(C1_POINTS)
    return S.scene_S(20_000, 640, 360, seed=5, log_scale_mean=np.log(0.03))


def _captured(d, capacity=None):
    from gsr_captured import CapturedRasterizer
    cr = CapturedRasterizer(d["P"], d["M"], d["sh_degree"], d["W"], d["H"], d["tanfovx"], d["tanfovy"], d["bg"],
                            scale_modifier=d["scale_modifier"], capacity=capacity)
    cr.set_camera(d["viewmatrix"], d["projmatrix"], d["campos"])
    cr.load(d["means3D"], d["sh"], d["opacities"], d["scales"], d["rotations"])
    return cr


@pytest.mark.parametrize("name", ["c1", "s20k"])
def test_captured_matches_entry_points(name):
    sc = _scene(name)
    d = to_device(sc)
    dL = torch.from_numpy(S.upstream_grad(d["W"], d["H"])).cuda()
    fwd = ours_forward(d)
    want = ours_backward(d, fwd, dL)
    cr = _captured(d)
    for rep in range(2):  # the second replay runs on buffers that hold the first one's results
        color, radii = cr.forward()
        g = cr.backward(dL)
        torch.cuda.synchronize()
        ok, R = cr.fits()
        assert ok and R == fwd["num_rendered"] and cr.recaptures == 0
        assert torch.equal(color, fwd["color"]), "captured forward image differs (bit-exact bar)"
        assert torch.equal(radii, fwd["radii"])
        # all nine tensors, the parity tests' bar (dL/drot of C1's isotropic Gaussians is rounding noise around an
        # exact 0 and is measured against its natural scale there)
        _check_grads({k: v.reshape(want[k].shape) for k, v in g.items()}, want, tol=GRAD_TOL, sc=sc)


def test_captured_overflow_recaptures_and_blind_backward_is_harmless():
    d = to_device(_scene("s20k"))
    dL = torch.from_numpy(S.upstream_grad(d["W"], d["H"])).cuda()
    fwd = ours_forward(d)
    R = fwd["num_rendered"]
    assert R > 4096
    cr = _captured(d, capacity=R // 4)
    # sync-free use: the lists do not fit, nothing is rendered, the backward walks no list
    cr.forward(check=False)
    g = cr.backward(dL)
    torch.cuda.synchronize()
    ok, seen = cr.fits()
    assert not ok and seen == R
    assert float(g["means3D"].abs().max()) == 0.0 and float(g["sh"].abs().max()) == 0.0
    # checked use: notices, grows the buffer, repeats the step
    color, _ = cr.forward()
    torch.cuda.synchronize()
    assert cr.recaptures == 1 and cr.capacity >= R
    assert torch.equal(color, fwd["color"])


def test_captured_autograd_wrapper():
    d = to_device(_scene("s20k"))
    dL = torch.from_numpy(S.upstream_grad(d["W"], d["H"])).cuda()
    fwd = ours_forward(d)
    want = ours_backward(d, fwd, dL)
    cr = _captured(d)
    leaves = [d[k].clone().requires_grad_(True) for k in ("means3D", "sh", "opacities", "scales", "rotations")]
    color, radii = cr.rasterize(*leaves)
    assert torch.equal(color, fwd["color"]) and torch.equal(radii, fwd["radii"])
    (color * dL).sum().backward()
    for leaf, key in zip(leaves, ("means3D", "sh", "opacity", "scales", "rot")):
        e = rel_err(np_(leaf.grad).reshape(-1), np_(want[key]).reshape(-1))
        assert e <= GRAD_TOL, (key, e)
