"""gsr_activations.activate against the torch expressions of scene/gaussian_model.py:43-51 (sigmoid / exp /
F.normalize): bit-identical values, gradients to rounding."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _raw(P, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(P, 1, device="cuda", generator=g) * 4, torch.randn(P, 3, device="cuda", generator=g) * 2 - 3,
            torch.randn(P, 4, device="cuda", generator=g))


@pytest.mark.parametrize("P", [1, 257, 100003])
def test_values_bit_identical_to_torch(P):
    from gsr_activations import activate
    o, s, q = _raw(P, 1)
    q[0] = 0.0  # zero quaternion: normalize clamps the norm at 1e-12
    a, b, c = activate(o, s, q)
    assert torch.equal(a, torch.sigmoid(o)) and torch.equal(b, torch.exp(s)) and torch.equal(c, F.normalize(q))


def test_gradients_match_autograd():
    from gsr_activations import activate
    P = 50021
    raws = _raw(P, 2)
    up = _raw(P, 3)
    got, want = [], []
    for fused in (True, False):
        o, s, q = (t.clone().requires_grad_(True) for t in raws)
        a, b, c = activate(o, s, q) if fused else (torch.sigmoid(o), torch.exp(s), F.normalize(q))
        ((a * up[0]).sum() + (b * up[1]).sum() + (c * up[2]).sum()).backward()
        (got if fused else want).extend([o.grad, s.grad, q.grad])
    for g, w in zip(got, want):
        assert (g - w).abs().max().item() <= 2e-6 * w.abs().max().item()


def test_partial_use_and_errors():
    from gsr_activations import activate
    o, s, q = (t.requires_grad_(True) for t in _raw(100, 4))
    a, b, c = activate(o, s, q)
    b.sum().backward()  # only one output used: the others get zero upstream gradients
    assert float(o.grad.abs().max()) == 0.0 and float(q.grad.abs().max()) == 0.0
    torch.testing.assert_close(s.grad, torch.exp(s.detach()), rtol=1e-6, atol=0)
    with pytest.raises(RuntimeError):
        activate(o.cpu(), s.cpu(), q.cpu())
    with pytest.raises(RuntimeError):
        activate(o, s, q[:, :3])


@pytest.mark.parametrize("P,M,deg,split", [(3001, 16, 3, False), (2500, 16, 3, True), (1200, 4, 1, False)])
def test_raw_parameter_entry_folds_the_activations(P, M, deg, split):
    """SURVEY 8f rank 3 as written: GaussianRasterizer(..., raw_params=True) takes _opacity / _scaling / _rotation
    (scene/gaussian_model.py:95-106) and runs sigmoid / exp / normalize inside preprocess_fwd, their chain rule
    inside preprocess_bwd.  Against torch activations + the ordinary entry: image, radii bit-identical (the
    activated values are the same bits), raw-parameter gradients to rounding + blend-atomics order."""
    import numpy as np
    import gsr_scenes as S
    from diff_gaussian_rasterization import GaussianRasterizer
    from util import settings_from, to_device
    sc = S.scene_S(P, 320, 208, seed=31, M=M, sh_degree=deg)
    d = to_device(sc)
    g = torch.Generator(device="cuda").manual_seed(5)
    o_raw = torch.logit(d["opacities"].clamp(1e-4, 1 - 1e-4))
    s_raw = torch.log(d["scales"])
    q_raw = d["rotations"] * (0.5 + torch.rand(P, 1, device="cuda", generator=g) * 2)  # unnormalised, as trained
    dpix = torch.from_numpy(S.upstream_grad(sc["W"], sc["H"], 3)).cuda()
    raster = GaussianRasterizer(settings_from(d))
    res = []
    for raw in (True, False):
        leaves = [t.clone().requires_grad_(True) for t in (d["means3D"], d["sh"], o_raw, s_raw, q_raw)]
        m3, sh, o, s, q = leaves
        shs = (sh[:, :1].contiguous(), sh[:, 1:].contiguous()) if split else sh
        if raw:
            color, radii = raster(means3D=m3, means2D=torch.zeros_like(m3), opacities=o, shs=shs, scales=s,
                                  rotations=q, raw_params=True)
        else:
            color, radii = raster(means3D=m3, means2D=torch.zeros_like(m3), opacities=torch.sigmoid(o), shs=shs,
                                  scales=torch.exp(s), rotations=F.normalize(q))
        color.backward(dpix)
        res.append((color.detach(), radii, [t.grad for t in leaves]))
    (c1, r1, g1), (c0, r0, g0) = res
    assert torch.equal(r1, r0) and torch.equal(c1, c0)
    for a, b, name in zip(g1, g0, ("means3D", "sh", "opacity_raw", "scaling_raw", "rotation_raw")):
        assert a.shape == b.shape and float(b.abs().max()) > 0, name
        assert float((a - b).abs().max()) <= 1e-3 * float(b.abs().max()), name
    with pytest.raises(Exception, match="raw_params needs"):
        raster(means3D=d["means3D"], means2D=d["means3D"], opacities=o_raw, shs=d["sh"],
               cov3D_precomp=torch.zeros(P, 6, device="cuda"), raw_params=True)
