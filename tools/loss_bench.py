"""Times the fused L1 + SSIM loss (forward + backward) against the reference formula in plain torch, 3x1080x1920."""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gaussian-transformer_b200"))
from utils.loss_utils import l1_ssim_loss  # noqa: E402


def torch_loss(img, gt, win, lam):
    conv = lambda t: F.conv2d(t, win, padding=5, groups=3)
    mu1, mu2 = conv(img), conv(gt)
    mu1_sq, mu2_sq, mu12 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s11, s22, s12 = conv(img * img) - mu1_sq, conv(gt * gt) - mu2_sq, conv(img * gt) - mu12
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim = (((2 * mu12 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s11 + s22 + C2))).mean()
    return (1.0 - lam) * torch.abs(img - gt).mean() + lam * (1.0 - ssim)


def timed(fn, n=50, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def main():
    H, W = 1080, 1920
    gt = torch.rand(3, H, W, device="cuda")
    img = (gt + 0.05 * torch.randn_like(gt)).clamp(0, 1).requires_grad_(True)
    w = torch.from_numpy(np.load(os.path.join(ROOT, "tests", "golden", "data", "loss_window.npy"))).cuda()
    win = (w[:, None] @ w[None, :]).expand(3, 1, 11, 11).contiguous()

    def ours():
        img.grad = None
        l1_ssim_loss(img, gt, 0.2)[0].backward()

    def ref():
        img.grad = None
        torch_loss(img[None], gt[None], win, 0.2).backward()

    if os.environ.get("GSR_TIMELINE"):
        from torch.profiler import ProfilerActivity, profile
        for _ in range(3):
            ours()
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as tp:
            ours(); ours()
            torch.cuda.synchronize()
        ev = sorted((e for e in tp.events() if e.device_type.name == "CUDA"), key=lambda e: e.time_range.start)
        for e in ev:
            print("%9.1f %8.1f  %s" % (e.time_range.start - ev[0].time_range.start, e.time_range.end - e.time_range.start, e.name[:80]))
        return
    t_o, t_r = timed(ours), timed(ref)
    px = 3 * H * W
    # algorithmic bytes: pass 1 reads img, gt and writes 3 maps; pass 2 reads 3 maps, img, gt and writes the gradient
    alg = (2 + 3 + 3 + 2 + 1) * 4 * px
    print(f"fused {t_o*1e3:.1f} us  ({alg/t_o/1e6:.0f} GB/s algorithmic)   torch {t_r*1e3:.1f} us   x{t_r/t_o:.1f}")


if __name__ == "__main__":
    main()
