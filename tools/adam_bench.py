"""Times one optimiser step over the six 3DGS parameter groups (P Gaussians, 59 floats each): FusedAdam vs
torch.optim.Adam (default foreach) vs torch.optim.Adam(fused=True)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gaussian-transformer_b200"))
from gsr_optim import FusedAdam  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
SHAPES = {"xyz": (P, 3), "f_dc": (P, 1, 3), "f_rest": (P, 15, 3), "opacity": (P, 1), "scaling": (P, 3), "rotation": (P, 4)}


def make(cls, **kw):
    ps = {k: torch.randn(s, device="cuda").requires_grad_(True) for k, s in SHAPES.items()}
    for p in ps.values():
        p.grad = torch.randn_like(p) * 1e-3
    return cls([{"params": [p], "lr": 1e-3, "name": k} for k, p in ps.items()], lr=0.0, eps=1e-15, **kw)


def timed(opt, n=30, warm=5):
    for _ in range(warm):
        opt.step()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        opt.step()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


n = sum(int(torch.tensor(s).prod()) for s in SHAPES.values())
t_f = timed(make(FusedAdam))
t_e = timed(make(torch.optim.Adam))
t_t = timed(make(torch.optim.Adam, fused=True))
print(f"P={P}: FusedAdam {t_f:.0f} us ({28 * n / t_f / 1e3:.0f} GB/s of 28 B/element) | torch foreach {t_e:.0f} us "
      f"(x{t_e / t_f:.1f}) | torch fused=True {t_t:.0f} us (x{t_t / t_f:.1f})")
