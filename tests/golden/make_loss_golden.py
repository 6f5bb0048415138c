"""Generates tests/golden/data/loss_*.npz by importing the reference's own utils/loss_utils.py (CPU torch, fp32)
in the build container: values of l1_loss / ssim / the train.py:91-92 combination and their autograd gradients
w.r.t. the rendered image.  /root/reference does not exist on the GPU box, so the outputs are committed.

    python tests/golden/make_loss_golden.py
"""
import importlib.util
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/utils/loss_utils.py"


def pair(C, H, W, seed, kind):
    """A (rendered, ground truth) pair in [0,1]."""
    r = np.random.default_rng(seed)
    if kind == "noise":
        return r.random((C, H, W), dtype=np.float32), r.random((C, H, W), dtype=np.float32)
    # smooth image + a perturbed copy, some pixels left exactly equal (sign(0) = 0 in the L1 gradient)
    yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
    gt = np.stack([0.5 + 0.4 * np.sin(0.11 * xx * (c + 1) + 0.07 * yy) * np.cos(0.05 * yy * (c + 2)) for c in range(C)])
    img = np.clip(gt + 0.08 * r.standard_normal((C, H, W)), 0.0, 1.0)
    keep = r.random((C, H, W)) < 0.1
    img[keep] = gt[keep]
    return img.astype(np.float32), gt.astype(np.float32)


CASES = {  # name: (C, H, W, seed, kind, lambda_dssim)
    "loss_noise_37x53": (3, 37, 53, 1, "noise", 0.2),
    "loss_smooth_64x64": (3, 64, 64, 2, "smooth", 0.2),
    "loss_tiny_1x7x5": (1, 7, 5, 3, "noise", 0.5),       # smaller than the window in both directions
    "loss_thin_3x11x200": (3, 11, 200, 4, "smooth", 0.2),
    "loss_view_3x90x160": (3, 90, 160, 5, "smooth", 0.2),
}


def main():
    spec = importlib.util.spec_from_file_location("ref_loss_utils", REF)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    torch.manual_seed(0)
    torch.set_num_threads(4)
    for name, (C, H, W, seed, kind, lam) in CASES.items():
        img_np, gt_np = pair(C, H, W, seed, kind)
        gt = torch.from_numpy(gt_np)
        rec = {"img": img_np, "gt": gt_np, "lambda_dssim": np.float32(lam)}
        img = torch.from_numpy(img_np).requires_grad_(True)
        l1 = ref.l1_loss(img, gt)
        ss = ref.ssim(img, gt)
        loss = (1.0 - lam) * l1 + lam * (1.0 - ss)      # train.py:91-92
        loss.backward()
        rec.update(l1=l1.detach().numpy(), ssim=ss.detach().numpy(), loss=loss.detach().numpy(),
                   grad_loss=img.grad.numpy().copy())
        img.grad = None
        ref.ssim(img, gt).backward()
        rec["grad_ssim"] = img.grad.numpy().copy()
        np.savez_compressed(os.path.join(HERE, "data", name + ".npz"), **rec)
        print(name, float(l1), float(ss), float(loss))
    # batched, size_average=False (loss_utils.py:62-63)
    r = np.random.default_rng(9)
    a, b = r.random((2, 3, 20, 24), dtype=np.float32), r.random((2, 3, 20, 24), dtype=np.float32)
    v = ref.ssim(torch.from_numpy(a), torch.from_numpy(b), size_average=False)
    np.savez_compressed(os.path.join(HERE, "data", "loss_batch_2x3x20x24.npz"), img=a, gt=b, ssim=v.numpy())
    print("batch", v.numpy())
    # the window itself (loss_utils.py:23-25)
    np.save(os.path.join(HERE, "data", "loss_window.npy"), ref.gaussian(11, 1.5).numpy())


if __name__ == "__main__":
    main()
