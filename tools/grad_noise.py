"""Dev tool: how far apart are two runs of the REFERENCE backward (nondeterministic fp32 atomics), and ours vs it?"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "gaussian-transformer_b200"), os.path.join(ROOT, "tests")]
import gsr_scenes as S  # noqa: E402
from checkers import Oracle, Reference  # noqa: E402
from util import ours_backward, ours_forward, rel_err, to_device  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
W, H = 1920, 1080
sc = S.scene_S(P, W, H, seed=0)
d = to_device(sc)
dpix = torch.from_numpy(S.upstream_grad(W, H, 1)).cuda()
ref = Reference()
ref.forward(d, internals=False)
g1 = {k: v.cpu().numpy() for k, v in ref.backward(dpix).items()}
ref.forward(d, internals=False)
g2 = {k: v.cpu().numpy() for k, v in ref.backward(dpix).items()}
f = ours_forward(d)
go = {k: v.cpu().numpy() for k, v in ours_backward(d, f, dpix).items()}
orc = Oracle()  # fp64 accumulation: the arbiter between two fp32 summation orders
orc.forward(sc, internals=False)
gt = orc.backward(dpix.cpu().numpy())
ORC = {"means2D": "means2D", "conic": "conic", "opacity": "opacity", "colors": "colors", "means3D": "means3D",
       "cov3D": "cov3D", "sh": "sh", "scales": "scales", "rot": "rot"}
for k in g1:
    a, b, o = g1[k], g2[k], go[k].reshape(g1[k].shape)
    if ORC.get(k) in gt:
        t = gt[ORC[k]].reshape(a.shape)
        print(f"{k:8s} vs fp64-accumulating oracle: ours {rel_err(o, t):.2e}  ref {rel_err(a, t):.2e}")
    i = np.unravel_index(np.abs(o - a).argmax(), a.shape)
    print(f"{k:8s} ref-vs-ref {rel_err(b, a):.2e}  ours-vs-ref {rel_err(o, a):.2e}  max|ref| {np.abs(a).max():.3e} "
          f"worst idx {i}: ours {o[i]:.6e} ref1 {a[i]:.6e} ref2 {b[i]:.6e}")
