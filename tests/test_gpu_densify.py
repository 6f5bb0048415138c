"""gsr_densify (compaction kernels) against the reference's own statements for prune / clone / split
(scene/gaussian_model.py:273-403), restated here in plain torch (boolean-mask indexing + torch.cat, the TEST ORACLE):
parameters, Adam moments, statistics and the optimizer wiring must come out bit-identical on seeded cases."""
import copy

import pytest
import torch
from torch import nn

pytestmark = pytest.mark.gpu

NAMES = (("xyz", "_xyz"), ("f_dc", "_features_dc"), ("f_rest", "_features_rest"), ("opacity", "_opacity"),
         ("scaling", "_scaling"), ("rotation", "_rotation"))


class Model:
    """The slice of GaussianModel the densification code touches (attribute names of scene/gaussian_model.py)."""

    def __init__(self, P, seed, with_state=True):
        g = torch.Generator(device="cuda").manual_seed(seed)
        r = lambda *s: torch.randn(*s, generator=g, device="cuda")
        self.percent_dense = 0.01
        self._xyz = nn.Parameter(r(P, 3))
        self._features_dc = nn.Parameter(r(P, 1, 3))
        self._features_rest = nn.Parameter(r(P, 15, 3))
        self._opacity = nn.Parameter(r(P, 1) * 3)
        self._scaling = nn.Parameter(r(P, 3) - 4.0)
        self._rotation = nn.Parameter(r(P, 4))
        self.xyz_gradient_accum = r(P, 1).abs() * 1e-3
        self.denom = torch.randint(0, 4, (P, 1), generator=g, device="cuda").float()
        self.max_radii2D = r(P).abs() * 30
        groups = [{"params": [getattr(self, a)], "lr": 1e-3, "name": n} for n, a in NAMES]
        self.optimizer = torch.optim.Adam(groups, lr=0.0, eps=1e-15)
        if with_state:
            for _, a in NAMES:
                getattr(self, a).grad = torch.randn(getattr(self, a).shape, generator=g, device="cuda")
            self.optimizer.step()
            self.optimizer.zero_grad(set_to_none=True)

    def tensors(self):
        out = {a: getattr(self, a).detach() for _, a in NAMES}
        out.update(accum=self.xyz_gradient_accum, denom=self.denom, radii=self.max_radii2D)
        for gname, group in zip(NAMES, self.optimizer.param_groups):
            p = group["params"][0]
            assert p is getattr(self, gname[1]), "optimizer group does not hold the model's parameter"
            st = self.optimizer.state.get(p)
            if st is not None:
                out[gname[0] + ".m"], out[gname[0] + ".v"] = st["exp_avg"], st["exp_avg_sq"]
        return out


# ---- the reference's statements (test oracle) ----------------------------------------------------------------
def ref_prune_optimizer(m, mask):  # gaussian_model.py:273-289
    out = {}
    for group in m.optimizer.param_groups:
        st = m.optimizer.state.get(group["params"][0], None)
        if st is not None:
            st["exp_avg"] = st["exp_avg"][mask]
            st["exp_avg_sq"] = st["exp_avg_sq"][mask]
            del m.optimizer.state[group["params"][0]]
            group["params"][0] = nn.Parameter(group["params"][0][mask].requires_grad_(True))
            m.optimizer.state[group["params"][0]] = st
        else:
            group["params"][0] = nn.Parameter(group["params"][0][mask].requires_grad_(True))
        out[group["name"]] = group["params"][0]
    return out


def ref_prune_points(m, mask):  # :291-305
    valid = ~mask
    t = ref_prune_optimizer(m, valid)
    for n, a in NAMES:
        setattr(m, a, t[n])
    m.xyz_gradient_accum = m.xyz_gradient_accum[valid]
    m.denom = m.denom[valid]
    m.max_radii2D = m.max_radii2D[valid]


def ref_postfix(m, d):  # :307-347
    for group in m.optimizer.param_groups:
        ext = d[group["name"]]
        st = m.optimizer.state.get(group["params"][0], None)
        if st is not None:
            st["exp_avg"] = torch.cat((st["exp_avg"], torch.zeros_like(ext)), dim=0)
            st["exp_avg_sq"] = torch.cat((st["exp_avg_sq"], torch.zeros_like(ext)), dim=0)
            del m.optimizer.state[group["params"][0]]
            group["params"][0] = nn.Parameter(torch.cat((group["params"][0], ext), dim=0).requires_grad_(True))
            m.optimizer.state[group["params"][0]] = st
        else:
            group["params"][0] = nn.Parameter(torch.cat((group["params"][0], ext), dim=0).requires_grad_(True))
        setattr(m, dict(NAMES)[group["name"]], group["params"][0])
    n = m._xyz.shape[0]
    m.xyz_gradient_accum = torch.zeros((n, 1), device="cuda")
    m.denom = torch.zeros((n, 1), device="cuda")
    m.max_radii2D = torch.zeros((n), device="cuda")


def ref_clone(m, grads, thr, extent):  # :374-387
    sel = torch.where(torch.norm(grads, dim=-1) >= thr, True, False)
    sel = torch.logical_and(sel, torch.max(torch.exp(m._scaling), dim=1).values <= m.percent_dense * extent)
    ref_postfix(m, {n: getattr(m, a)[sel] for n, a in NAMES})


def ref_split(m, grads, thr, extent, N=2):  # :349-372
    from gsr_densify import _build_rotation
    n0 = m._xyz.shape[0]
    padded = torch.zeros((n0), device="cuda")
    padded[:grads.shape[0]] = grads.squeeze()
    sel = torch.where(padded >= thr, True, False)
    sel = torch.logical_and(sel, torch.max(torch.exp(m._scaling), dim=1).values > m.percent_dense * extent)
    stds = torch.exp(m._scaling)[sel].repeat(N, 1)
    samples = torch.normal(mean=torch.zeros((stds.size(0), 3), device="cuda"), std=stds)
    rots = _build_rotation(m._rotation[sel]).repeat(N, 1, 1)
    d = {"xyz": torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + m._xyz[sel].repeat(N, 1),
         "scaling": torch.log(torch.exp(m._scaling)[sel].repeat(N, 1) / (0.8 * N)),
         "rotation": m._rotation[sel].repeat(N, 1), "f_dc": m._features_dc[sel].repeat(N, 1, 1),
         "f_rest": m._features_rest[sel].repeat(N, 1, 1), "opacity": m._opacity[sel].repeat(N, 1)}
    ref_postfix(m, d)
    ref_prune_points(m, torch.cat((sel, torch.zeros(N * int(sel.sum()), device="cuda", dtype=bool))))


def ref_densify_and_prune(m, max_grad, min_opacity, extent, max_screen_size):  # :389-403
    grads = m.xyz_gradient_accum / m.denom
    grads[grads.isnan()] = 0.0
    ref_clone(m, grads, max_grad, extent)
    ref_split(m, grads, max_grad, extent)
    prune = (torch.sigmoid(m._opacity) < min_opacity).squeeze()
    if max_screen_size:
        prune = torch.logical_or(torch.logical_or(prune, m.max_radii2D > max_screen_size),
                                 torch.exp(m._scaling).max(dim=1).values > 0.1 * extent)
    ref_prune_points(m, prune)


def _same(a, b):
    ta, tb = a.tensors(), b.tensors()
    assert ta.keys() == tb.keys()
    for k in ta:
        assert ta[k].shape == tb[k].shape, (k, ta[k].shape, tb[k].shape)
        assert torch.equal(ta[k], tb[k]), k


@pytest.mark.parametrize("P,with_state", [(1, True), (4097, True), (50_000, True), (3000, False)])
def test_prune_points_bit_identical(P, with_state):
    import gsr_densify
    a, b = Model(P, 1, with_state), Model(P, 1, with_state)
    mask = torch.rand(P, device="cuda") < 0.3
    gsr_densify.prune_points(a, mask)
    ref_prune_points(b, mask)
    _same(a, b)
    assert a._xyz.shape[0] == int((~mask).sum())


@pytest.mark.parametrize("P,size", [(20_000, 20), (20_000, None), (257, 20)])
def test_densify_and_prune_bit_identical(P, size):
    """The whole step of train.py:118-120 on a seeded model: clone, split (same random samples: torch.normal is
    called with the reference's shapes from the same generator state) and the final prune."""
    import gsr_densify
    a, b = Model(P, 7), Model(P, 7)
    extent, thr = 4.0, 2e-4
    for m, fn in ((a, gsr_densify.densify_and_prune), (b, ref_densify_and_prune)):
        torch.manual_seed(123)
        fn(m, thr, 0.005, extent, size)
    _same(a, b)
    assert a._xyz.shape[0] != P  # something happened
    # the optimizer still steps (state surgery kept torch's layout)
    for m in (a, b):
        for _, at in NAMES:
            getattr(m, at).grad = torch.ones_like(getattr(m, at))
        m.optimizer.step()
    _same(a, b)


def test_clone_and_split_counts():
    import gsr_densify
    a = Model(10_000, 3)
    grads = a.xyz_gradient_accum / a.denom
    grads[grads.isnan()] = 0.0
    P = a._xyz.shape[0]
    C = gsr_densify.densify_and_clone(a, grads, 2e-4, 4.0)
    assert a._xyz.shape[0] == P + C and float(a.denom.abs().sum()) == 0.0
    S = gsr_densify.densify_and_split(a, grads, 2e-4, 4.0)
    assert a._xyz.shape[0] == P + C + S
