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
