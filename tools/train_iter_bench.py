"""One 3DGS optimisation iteration as train.py:80-128 runs it -- activations + SH concat (gaussian_model.py:95-115),
render, L1 + D-SSIM loss, backward, Adam step over the six groups -- timed with the pieces swapped one at a time:

  reference : the UNMODIFIED reference rasterizer (oracle/_ref, test infrastructure) + utils/loss_utils.py formula in
              torch + torch.optim.Adam          (what the reference's train.py does)
  raster    : libgsrast rasterizer + torch loss + torch Adam
  +loss     : ... + fused L1/SSIM loss (utils.loss_utils.l1_ssim_loss)
  +adam     : ... + gsr_optim.FusedAdam
  +split    : ... + shs=(features_dc, features_rest) read in place, no torch.cat
  +act      : ... + gsr_activations.activate (one kernel per direction)
  +stats    : ... + gsr_optim.densification_stats (no boolean-mask host sync)   (everything this repository builds)

Every configuration also keeps the per-iteration densification statistics of train.py:113-116.

Synthetic scene S (1M Gaussians, SH degree 3, 1920x1080), ground truth = a perturbed render.  Dev tool; needs a GPU.
"""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "gaussian-transformer_b200"), os.path.join(ROOT, "tests")]
import gsr_scenes as S  # noqa: E402
from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer  # noqa: E402
from gsr_activations import activate  # noqa: E402
from gsr_optim import FusedAdam, densification_stats  # noqa: E402
from utils.loss_utils import l1_ssim_loss  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
FUSED = ("raster", "+loss", "+adam", "+split", "+act", "+stats")  # cumulative
W, H, DEG, M, LAMBDA = 1920, 1080, 3, 16, 0.2


def torch_loss(img, gt, win):
    conv = lambda t: F.conv2d(t, win, padding=5, groups=3)
    x, y = img[None], gt[None]
    mu1, mu2 = conv(x), conv(y)
    mu1_sq, mu2_sq, mu12 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s11, s22, s12 = conv(x * x) - mu1_sq, conv(y * y) - mu2_sq, conv(x * y) - mu12
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim = (((2 * mu12 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s11 + s22 + C2))).mean()
    return (1.0 - LAMBDA) * torch.abs(img - gt).mean() + LAMBDA * (1.0 - ssim)


class RefRaster(torch.autograd.Function):
    """The unmodified reference rasterizer behind autograd (its own torch glue is not built here)."""

    @staticmethod
    def forward(ctx, ref, d, means3D, means2D, sh, opacities, scales, rotations):
        dd = dict(d, means3D=means3D.detach(), sh=sh.detach(), opacities=opacities.detach(), scales=scales.detach(),
                  rotations=rotations.detach())
        f = ref.forward(dd, internals=False, sync=False)
        ctx.ref = ref
        ctx.mark_non_differentiable(f["radii"])
        return f["color"], f["radii"]

    @staticmethod
    def backward(ctx, g, _g_radii):
        r = ctx.ref.backward(g.contiguous(), sync=False)
        return None, None, r["means3D"], r["means2D"], r["sh"], r["opacity"], r["scales"], r["rot"]


def main():
    from checkers import Reference
    from util import to_device
    sc = S.scene_S(P, W, H, seed=0, M=M, sh_degree=DEG)
    d = to_device(sc)
    settings = GaussianRasterizationSettings(H, W, sc["tanfovx"], sc["tanfovy"], d["bg"], 1.0, d["viewmatrix"],
                                             d["projmatrix"], DEG, d["campos"], False, False)
    raster = GaussianRasterizer(settings)
    ref = Reference()
    w = torch.from_numpy(np.load(os.path.join(ROOT, "tests", "golden", "data", "loss_window.npy"))).cuda()
    win = (w[:, None] @ w[None, :]).expand(3, 1, 11, 11).contiguous()
    with torch.no_grad():
        gt, _ = raster(means3D=d["means3D"], means2D=torch.zeros_like(d["means3D"]), opacities=d["opacities"],
                       shs=d["sh"], scales=d["scales"], rotations=d["rotations"])
        gt = (gt + 0.03 * torch.randn_like(gt)).clamp(0, 1)

    def params():  # the raw parameters of scene/gaussian_model.py
        op = d["opacities"].clamp(1e-4, 1 - 1e-4)
        raw = {"xyz": d["means3D"].clone(), "f_dc": d["sh"][:, :1].clone(), "f_rest": d["sh"][:, 1:].clone(),
               "opacity": torch.log(op / (1 - op)), "scaling": torch.log(d["scales"]), "rotation": d["rotations"].clone()}
        return {k: v.contiguous().requires_grad_(True) for k, v in raw.items()}

    def make(kind):
        p = params()
        groups = [{"params": [p[k]], "lr": lr, "name": k} for k, lr in
                  (("xyz", 1.6e-4), ("f_dc", 2.5e-3), ("f_rest", 1.25e-4), ("opacity", 0.05), ("scaling", 5e-3),
                   ("rotation", 1e-3))]
        opt = (FusedAdam if kind in FUSED[2:] else torch.optim.Adam)(groups, lr=0.0, eps=1e-15)
        stats = [torch.zeros(P, device="cuda"), torch.zeros(P, 1, device="cuda"), torch.zeros(P, 1, device="cuda")]

        def it():
            sh = (p["f_dc"], p["f_rest"]) if kind in FUSED[3:] else torch.cat((p["f_dc"], p["f_rest"]), dim=1)
            if kind in FUSED[4:]:
                opac, scl, rot = activate(p["opacity"], p["scaling"], p["rotation"])
            else:
                opac, scl = torch.sigmoid(p["opacity"]), torch.exp(p["scaling"])
                rot = F.normalize(p["rotation"])
            m2 = torch.zeros_like(p["xyz"], requires_grad=True)      # screenspace_points
            if kind == "reference":
                img, radii = RefRaster.apply(ref, d, p["xyz"], m2, sh, opac, scl, rot)
            else:
                img, radii = raster(means3D=p["xyz"], means2D=m2, opacities=opac, shs=sh, scales=scl, rotations=rot)
            if kind in FUSED[1:]:
                loss, _ = l1_ssim_loss(img, gt, LAMBDA)
            else:
                loss = torch_loss(img, gt, win)
            loss.backward()
            with torch.no_grad():                                    # train.py:113-116
                if kind == "+stats":
                    densification_stats(radii, m2.grad, *stats)
                else:
                    vis = radii > 0
                    stats[0][vis] = torch.max(stats[0][vis], radii[vis])
                    stats[1][vis] += torch.norm(m2.grad[vis, :2], dim=-1, keepdim=True)
                    stats[2][vis] += 1
            opt.step()
            opt.zero_grad(set_to_none=True)
        return it

    if os.environ.get("GSR_TIMELINE"):
        from torch.profiler import ProfilerActivity, profile
        it = make("+stats")
        for _ in range(3):
            it()
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as tp:
            it()
            it()
            torch.cuda.synchronize()
        ev = sorted((e for e in tp.events() if e.device_type.name == "CUDA"), key=lambda e: e.time_range.start)
        t0 = ev[0].time_range.start
        with open(os.environ["GSR_TIMELINE"], "w") as f:
            for e in ev:
                f.write("%9.1f %8.1f  %s\n" % (e.time_range.start - t0, e.time_range.end - e.time_range.start, e.name[:100]))
        return
    res = {}
    for kind in ("reference",) + FUSED:
        it = make(kind)
        for _ in range(3):
            it()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 20
        a.record()
        for _ in range(n):
            it()
        b.record()
        torch.cuda.synchronize()
        res[kind] = a.elapsed_time(b) / n
        print(f"{kind:10s} {res[kind]:7.3f} ms/iteration  ({1000 / res[kind]:6.1f} it/s)", flush=True)
    print(f"train.py-style iteration, P={P} {W}x{H}: {res['reference'] / res['+stats']:.2f}x the reference stack")


if __name__ == "__main__":
    main()
