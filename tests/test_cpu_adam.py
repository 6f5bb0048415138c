"""The Adam oracle (oracle.c orc_adam_step) against torch.optim.Adam itself (CPU, the reference's configuration:
scene/gaussian_model.py:163 `torch.optim.Adam(l, lr=0.0, eps=1e-15)` with per-group lr)."""
import numpy as np
import torch


def test_oracle_matches_torch_adam(oracle):
    torch.manual_seed(0)
    shapes = {"xyz": (501, 3), "f_dc": (501, 1, 3), "opacity": (501, 1), "rotation": (501, 4)}
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "opacity": 0.05, "rotation": 1e-3}
    params = {k: torch.randn(s).requires_grad_(True) for k, s in shapes.items()}
    opt = torch.optim.Adam([{"params": [params[k]], "lr": lrs[k], "name": k} for k in shapes], lr=0.0, eps=1e-15)
    mine = {k: [params[k].detach().numpy().copy(), None, np.zeros(shapes[k], np.float32),
                np.zeros(shapes[k], np.float32)] for k in shapes}
    for step in range(1, 8):
        for k, p in params.items():
            scale = 10.0 ** ((step % 3) - 4)  # gradients of very different magnitude, some exactly zero
            g = torch.randn(shapes[k]) * scale
            g[::7] = 0.0
            p.grad = g
            mine[k][1] = g.numpy().copy()
        if step == 4:
            lrs["xyz"] *= 0.5  # the trainer reschedules this group every iteration
            opt.param_groups[0]["lr"] = lrs["xyz"]
        opt.step()
        for k in shapes:
            p, g, m, v = mine[k]
            oracle.adam_step(p, g, m, v, step, lrs[k])
            st = opt.state[params[k]]
            np.testing.assert_allclose(m, st["exp_avg"].numpy(), rtol=1e-6, atol=1e-30)
            np.testing.assert_allclose(v, st["exp_avg_sq"].numpy(), rtol=1e-6, atol=1e-30)
            # one rounding of the update: relative to the step, not to the parameter
            assert np.abs(p - params[k].detach().numpy()).max() <= 4e-6 * lrs[k] + 1e-7 * np.abs(p).max()
