"""gsr_optim.FusedAdam (csrc/adam.cuh through the C ABI) against torch.optim.Adam on the same GPU and against the
oracle, in the reference's configuration (scene/gaussian_model.py:149-167: six groups, lr=0.0 default + per-group
lr, eps=1e-15)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GROUPS = {"xyz": ((3,), 1.6e-4), "f_dc": ((1, 3), 2.5e-3), "f_rest": ((15, 3), 2.5e-3 / 20.0), "opacity": ((1,), 0.05),
          "scaling": ((3,), 5e-3), "rotation": ((4,), 1e-3)}


def _make(P, seed, misalign=False):
    g = torch.Generator(device="cuda").manual_seed(seed)
    out = {}
    for k, (shp, _) in GROUPS.items():
        n = int(np.prod((P,) + shp))
        base = torch.randn(n + 1, device="cuda", generator=g)
        t = (base[1:] if misalign else base[:n]).view((P,) + shp)
        out[k] = t
    return out


def _optim(cls, params):
    return cls([{"params": [params[k]], "lr": GROUPS[k][1], "name": k} for k in GROUPS], lr=0.0, eps=1e-15)


@pytest.mark.parametrize("P,misalign", [(10007, False), (4099, True), (1, False)])
def test_fused_adam_matches_torch(P, misalign):
    from gsr_optim import FusedAdam
    init = _make(P, 1, misalign)
    a = {k: v.detach().clone().requires_grad_(True) for k, v in init.items()}
    if misalign:  # leaves whose storage starts 4 bytes off a 16-byte boundary: the scalar path
        a = {}
        for k, v in init.items():
            buf = torch.empty(v.numel() + 1, device="cuda")
            leaf = buf[1:].view(v.shape)
            leaf.copy_(v)
            a[k] = leaf.detach().requires_grad_(True)
            assert a[k].data_ptr() % 16 != 0
    b = {k: v.detach().clone().requires_grad_(True) for k, v in init.items()}
    oa, ob = _optim(FusedAdam, a), _optim(torch.optim.Adam, b)
    gen = torch.Generator(device="cuda").manual_seed(2)
    for step in range(1, 13):
        for k in GROUPS:
            g = torch.randn(a[k].shape, device="cuda", generator=gen) * 10.0 ** ((step % 4) - 5)
            g[::5] = 0.0
            a[k].grad, b[k].grad = g.clone(), g.clone()
        if step % 3 == 0:
            for o in (oa, ob):
                o.param_groups[0]["lr"] *= 0.7
        oa.step()
        ob.step()
    for k in GROUPS:
        sa, sb = oa.state[a[k]], ob.state[b[k]]
        assert int(sa["step"]) == int(sb["step"]) == 12
        # moments: to rounding of their largest term (exp_avg cancels where gradients alternate in sign)
        torch.testing.assert_close(sa["exp_avg"], sb["exp_avg"], rtol=2e-6, atol=2e-7 * sb["exp_avg"].abs().max().item())
        torch.testing.assert_close(sa["exp_avg_sq"], sb["exp_avg_sq"], rtol=2e-6,
                                   atol=2e-7 * sb["exp_avg_sq"].abs().max().item())
        lr = GROUPS[k][1]
        # 12 updates of at most ~lr each, one rounding apiece
        assert (a[k] - b[k]).abs().max().item() <= 12 * 4e-6 * lr + 2e-7 * b[k].abs().max().item()
        assert P == 1 or (a[k] - init[k]).abs().max().item() > 0  # P == 1: its only row has zero gradients


def test_fused_adam_matches_oracle(oracle):
    from gsr_optim import FusedAdam
    P = 3001
    p0 = torch.randn(P, 3, device="cuda")
    p = p0.clone().requires_grad_(True)
    opt = FusedAdam([{"params": [p], "lr": 3e-3}], lr=0.0, eps=1e-15)
    hp, hm, hv = p0.cpu().numpy().copy(), np.zeros((P, 3), np.float32), np.zeros((P, 3), np.float32)
    for step in range(1, 6):
        g = torch.randn(P, 3, device="cuda") * 1e-3
        p.grad = g
        opt.step()
        oracle.adam_step(hp, g.cpu().numpy().copy(), hm, hv, step, 3e-3)
    # same op order in float32: identical up to the device's fused multiply-adds
    np.testing.assert_allclose(opt.state[p]["exp_avg"].cpu().numpy(), hm, rtol=1e-6, atol=2e-7 * np.abs(hm).max())
    np.testing.assert_allclose(opt.state[p]["exp_avg_sq"].cpu().numpy(), hv, rtol=1e-6, atol=2e-7 * np.abs(hv).max())
    assert np.abs(p.detach().cpu().numpy() - hp).max() <= 5 * 4e-6 * 3e-3 + 2e-7 * np.abs(hp).max()


def test_state_surgery_like_densification():
    """gaussian_model.py:307-337 cat_tensors_to_optimizer: state rows are concatenated with zeros and the parameter
    replaced; :73,:93 state_dict round trip."""
    from gsr_optim import FusedAdam
    p = torch.randn(100, 3, device="cuda").requires_grad_(True)
    opt = FusedAdam([{"params": [p], "lr": 1e-2, "name": "xyz"}], lr=0.0, eps=1e-15)
    ref_p = p.detach().clone().requires_grad_(True)
    ref = torch.optim.Adam([{"params": [ref_p], "lr": 1e-2, "name": "xyz"}], lr=0.0, eps=1e-15)
    for o, q in ((opt, p), (ref, ref_p)):
        q.grad = torch.full_like(q, 1e-3)
        o.step()
    ext = torch.randn(20, 3, device="cuda")
    new = {}
    for o, q, tag in ((opt, p, "a"), (ref, ref_p, "b")):
        group = o.param_groups[0]
        st = o.state.get(group["params"][0], None)
        st["exp_avg"] = torch.cat((st["exp_avg"], torch.zeros_like(ext)), dim=0)
        st["exp_avg_sq"] = torch.cat((st["exp_avg_sq"], torch.zeros_like(ext)), dim=0)
        del o.state[group["params"][0]]
        group["params"][0] = torch.nn.Parameter(torch.cat((group["params"][0], ext), dim=0).requires_grad_(True))
        o.state[group["params"][0]] = st
        new[tag] = group["params"][0]
        new[tag].grad = torch.full_like(new[tag], -2e-3)
        o.step()
    torch.testing.assert_close(new["a"], new["b"], rtol=1e-6, atol=1e-7)
    sd = opt.state_dict()
    opt2 = FusedAdam([{"params": [new["a"]], "lr": 1e-2, "name": "xyz"}], lr=0.0, eps=1e-15)
    opt2.load_state_dict(sd)
    assert int(opt2.state[new["a"]]["step"]) == 2


def test_errors():
    from gsr_optim import FusedAdam
    with pytest.raises(ValueError):
        FusedAdam([torch.zeros(3, device="cuda", requires_grad=True)], weight_decay=0.1)
    p = torch.zeros(3, requires_grad=True)
    p.grad = torch.ones(3)
    with pytest.raises(RuntimeError):
        FusedAdam([p]).step()


def test_densification_stats_match_reference_updates():
    """gsr_optim.densification_stats vs the boolean-mask statements of train.py:113-116 / gaussian_model.py:405-407."""
    from gsr_optim import densification_stats
    P = 100003
    g = torch.Generator(device="cuda").manual_seed(7)
    radii = torch.randint(-1, 40, (P,), device="cuda", generator=g, dtype=torch.int32).clamp_min(0)
    grad = torch.randn(P, 3, device="cuda", generator=g) * 1e-4
    state = [torch.rand(P, device="cuda", generator=g) * 30, torch.rand(P, 1, device="cuda", generator=g),
             torch.randint(0, 50, (P, 1), device="cuda", generator=g).float()]
    want = [t.clone() for t in state]
    vis = radii > 0
    want[0][vis] = torch.max(want[0][vis], radii[vis])
    want[1][vis] += torch.norm(grad[vis, :2], dim=-1, keepdim=True)
    want[2][vis] += 1
    densification_stats(radii, grad, *state)
    assert torch.equal(state[0], want[0]) and torch.equal(state[2], want[2])
    torch.testing.assert_close(state[1], want[1], rtol=1e-6, atol=0)
    assert int((state[1] != want[1]).sum()) == 0  # torch's 2-element norm is sqrt(x*x + y*y), products rounded
    with pytest.raises(RuntimeError):
        densification_stats(radii.long(), grad, *state)
