"""Fused L1 + SSIM loss (csrc/loss.cuh through the C ABI) against the reference module's committed outputs, the
fp64 oracle, and -- at 1080p -- a plain torch fp32 restatement of utils/loss_utils.py run on the same GPU."""
import glob
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "data")
CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(DATA, "loss_*.npz")) if "batch" not in p)

VAL_RTOL = 1e-5    # loss values
GRAD_TOL = 1e-4    # of max |gradient| (fp32 cancellation in sigma = E[x^2] - mu^2, see test_cpu_loss.py)


def _torch_loss(img, gt, lam):
    """utils/loss_utils.py:17-18, :23-63 and train.py:91-92 in plain torch fp32 (fp32 reference for this kernel)."""
    w = torch.from_numpy(np.load(os.path.join(DATA, "loss_window.npy"))).to(img.device)
    C = img.size(-3)
    win = (w[:, None] @ w[None, :]).to(img.dtype).expand(C, 1, 11, 11).contiguous()
    conv = lambda t: F.conv2d(t, win, padding=5, groups=C)
    mu1, mu2 = conv(img), conv(gt)
    s11, s22, s12 = conv(img * img) - mu1 * mu1, conv(gt * gt) - mu2 * mu2, conv(img * gt) - mu1 * mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim = (((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (s11 + s22 + C2))).mean()
    l1 = (img - gt).abs().mean()
    return (1.0 - lam) * l1 + lam * (1.0 - ssim), l1, ssim


@pytest.mark.parametrize("name", CASES)
def test_loss_matches_reference_outputs(oracle, name):
    from utils.loss_utils import l1_loss, l1_ssim_loss, ssim
    d = np.load(os.path.join(DATA, name + ".npz"))
    lam = float(d["lambda_dssim"])
    gt = torch.from_numpy(d["gt"]).cuda()
    img = torch.from_numpy(d["img"]).cuda().requires_grad_(True)
    loss, l1 = l1_ssim_loss(img, gt, lam)
    loss.backward()
    want3, want_g = oracle.l1_ssim(d["img"], d["gt"], lam)
    np.testing.assert_allclose(l1.item(), want3[0], rtol=VAL_RTOL)
    np.testing.assert_allclose(loss.item(), want3[2], rtol=VAL_RTOL)
    np.testing.assert_allclose(loss.item(), float(d["loss"]), rtol=2e-5)   # the reference itself is fp32
    g = img.grad.cpu().numpy()
    assert np.abs(g - want_g).max() <= GRAD_TOL * np.abs(want_g).max()
    assert np.abs(g - d["grad_loss"]).max() <= GRAD_TOL * np.abs(want_g).max()
    # the separate entry points of the reference module
    img.grad = None
    s = ssim(img, gt)
    s.backward()
    np.testing.assert_allclose(s.item(), want3[1], rtol=VAL_RTOL)
    assert np.abs(img.grad.cpu().numpy() - d["grad_ssim"]).max() <= GRAD_TOL * np.abs(d["grad_ssim"]).max()
    img.grad = None
    a = l1_loss(img, gt)
    a.backward()
    np.testing.assert_allclose(a.item(), want3[0], rtol=VAL_RTOL)
    sign = np.sign(d["img"] - d["gt"]) / d["img"].size
    np.testing.assert_allclose(img.grad.cpu().numpy(), sign, rtol=1e-6, atol=0)


def test_ssim_batch_size_average_false():
    from utils.loss_utils import ssim
    d = np.load(os.path.join(DATA, "loss_batch_2x3x20x24.npz"))
    a, b = torch.from_numpy(d["img"]).cuda(), torch.from_numpy(d["gt"]).cuda()
    np.testing.assert_allclose(ssim(a, b, size_average=False).cpu().numpy(), d["ssim"], rtol=2e-5)
    # batch folded into channels for the global mean
    np.testing.assert_allclose(ssim(a, b).item(), d["ssim"].mean(), rtol=2e-5)


def test_identical_images_and_upstream_scale():
    from utils.loss_utils import l1_ssim_loss
    d = np.load(os.path.join(DATA, "loss_smooth_64x64.npz"))
    gt = torch.from_numpy(d["gt"]).cuda()
    img = gt.clone().requires_grad_(True)
    loss, l1 = l1_ssim_loss(img, gt, 0.2)
    (3.0 * loss).backward()
    assert l1.item() == 0.0 and abs(loss.item()) < 1e-6
    assert img.grad.abs().max().item() < 1e-7
    img2 = torch.from_numpy(d["img"]).cuda().requires_grad_(True)
    l_a, _ = l1_ssim_loss(img2, gt, 0.2)
    (3.0 * l_a).backward()
    g3 = img2.grad.clone()
    img2.grad = None
    l_b, _ = l1_ssim_loss(img2, gt, 0.2)
    l_b.backward()
    # the upstream scale enters the two weights before they are combined: one rounding of the larger term
    torch.testing.assert_close(g3, 3.0 * img2.grad, rtol=1e-6, atol=1e-6 * g3.abs().max().item())


def test_errors():
    from utils.loss_utils import l1_ssim_loss, ssim
    a = torch.rand(3, 8, 8)
    with pytest.raises(RuntimeError):
        l1_ssim_loss(a, a)                      # CPU tensors: no CPU path
    with pytest.raises(RuntimeError):
        ssim(a.cuda(), a.cuda(), window_size=7)
    with pytest.raises(RuntimeError):
        l1_ssim_loss(a.cuda(), torch.rand(3, 8, 9).cuda())


def test_1080p_vs_torch():
    """BASELINE.json's full image size.  The reference formula in torch fp32 is itself 1.4e-4 of max away from
    its fp64 evaluation here (sigma = E[x^2] - mu^2 cancels), so fp64 is the truth and fp32 torch the yardstick."""
    from utils.loss_utils import l1_ssim_loss
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    g = torch.Generator(device="cuda").manual_seed(5)
    H, W = 1080, 1920
    yy, xx = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
    gt = torch.stack([0.5 + 0.4 * torch.sin(0.013 * xx * (c + 1) + 0.02 * yy) * torch.cos(0.007 * yy * (c + 2))
                      for c in range(3)]).float()
    img = (gt + 0.05 * torch.randn(3, H, W, device="cuda", generator=g)).clamp(0, 1)
    a = img.clone().requires_grad_(True)
    loss, l1 = l1_ssim_loss(a, gt, 0.2)
    loss.backward()
    b = img.clone().requires_grad_(True)
    ref32, _, _ = _torch_loss(b[None], gt[None], 0.2)
    ref32.backward()
    c = img.double().requires_grad_(True)
    want, want_l1, _ = _torch_loss(c[None], gt.double()[None], 0.2)
    want.backward()
    np.testing.assert_allclose(l1.item(), want_l1.item(), rtol=VAL_RTOL)
    np.testing.assert_allclose(loss.item(), want.item(), rtol=VAL_RTOL)
    gmax = c.grad.abs().max().item()
    err = (a.grad - c.grad).abs().max().item()
    assert err <= GRAD_TOL * gmax
    assert err <= 1.25 * (b.grad - c.grad).abs().max().item()   # no farther from fp64 than the reference formula


def test_second_argument_is_differentiable_like_the_reference():
    """train_autoencoder.py:172 calls ssim(in_im, out_im) with the network output SECOND; the reference's functions
    are differentiable in both arguments.  Checked against the torch fp32 restatement's autograd: gradient of the
    second image alone, and of both images at once."""
    from utils.loss_utils import l1_loss, l1_ssim_loss, ssim
    g = torch.Generator().manual_seed(5)
    a = torch.rand(3, 70, 90, generator=g).cuda()
    b = (a + 0.2 * torch.rand(3, 70, 90, generator=g).cuda()).clamp(0, 1)
    for both in (False, True):
        x = a.clone().requires_grad_(both)
        y = b.clone().requires_grad_(True)
        (1.0 - ssim(x, y)).backward()
        xr = a.clone().requires_grad_(both)
        yr = b.clone().requires_grad_(True)
        _, _, s_ref = _torch_loss(xr[None], yr[None], 0.2)
        (1.0 - s_ref).backward()
        assert (y.grad - yr.grad).abs().max() <= 2 * GRAD_TOL * yr.grad.abs().max()
        if both:
            assert (x.grad - xr.grad).abs().max() <= 2 * GRAD_TOL * xr.grad.abs().max()
    y = b.clone().requires_grad_(True)
    l1_loss(a, y).backward()
    assert torch.equal(y.grad, torch.sign(b - a) / a.numel())
    y.grad = None
    loss, _ = l1_ssim_loss(a, y, 0.2)
    loss.backward()
    yr = b.clone().requires_grad_(True)
    _torch_loss(a[None], yr[None], 0.2)[0].backward()
    assert (y.grad - yr.grad).abs().max() <= 2 * GRAD_TOL * yr.grad.abs().max()
