"""Quick timing of ours vs the reference CUDA rasterizer on scene S (dev tool; bench.py is the contract)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "gaussian-transformer_b200"), os.path.join(ROOT, "tests")]
import gsr_scenes as S  # noqa: E402
from checkers import Reference  # noqa: E402
from util import ours_backward, ours_forward, to_device  # noqa: E402


def timeit(fn, n=int(os.environ.get('GSR_N', 10)), warm=int(os.environ.get('GSR_WARM', 3))):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))


def main():
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    W, H = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080)
    sc = S.scene_S(P, W, H, seed=0)
    d = to_device(sc)
    dpix = torch.from_numpy(S.upstream_grad(W, H, 1)).cuda()
    ref = Reference()
    f = ours_forward(d)
    r = np.asarray(f["ranges"]).astype(np.int64)
    ln = r[:, 1] - r[:, 0]
    print(f"P={P} {W}x{H} V={(f['radii'] > 0).sum().item()} R={f['num_rendered']} list mean={ln.mean():.1f} "
          f"max={ln.max()} sum_ncontrib={int(f['n_contrib'].sum())}")
    from diff_gaussian_rasterization import _C
    e = torch.Tensor([])

    def ours_f():
        return _C.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                      d["viewmatrix"], d["projmatrix"], d["tanfovx"], d["tanfovy"], H, W, d["sh"], 3,
                                      d["campos"], False, False)
    state = {}

    def ours_fb():
        R, c, radii, g, b, i = ours_f()
        _C.rasterize_gaussians_backward(d["bg"], d["means3D"], radii, e, d["scales"], d["rotations"], 1.0, e,
                                        d["viewmatrix"], d["projmatrix"], d["tanfovx"], d["tanfovy"], dpix, d["sh"],
                                        3, d["campos"], g, R, b, i, False)

    def ref_f():
        ref.forward(d, internals=False, sync=False)

    def ref_fb():
        ref.forward(d, internals=False, sync=False)
        ref.backward(dpix, sync=False)
    to, tob = timeit(ours_f), timeit(ours_fb)
    tr, trb = timeit(ref_f), timeit(ref_fb)
    print(f"ours fwd {to:.3f} ms  fwd+bwd {tob:.3f} ms | reference fwd {tr:.3f} ms  fwd+bwd {trb:.3f} ms | "
          f"speedup fwd {tr / to:.2f}x fwd+bwd {trb / tob:.2f}x")


if __name__ == "__main__":
    main()
