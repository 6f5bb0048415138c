"""`utils.loss_utils`: the reference's training loss (utils/loss_utils.py) on the library's fused kernels.

Same names and argument meaning as the reference module -- `l1_loss`, `l2_loss`, `ssim` -- plus `l1_ssim_loss`,
the train.py:91-92 combination in ONE forward+backward pair of launches (csrc/loss.cuh):

    Ll1  = l1_loss(image, gt_image)
    loss = (1.0 - opt.lambda_dssim) * Ll1 + opt.lambda_dssim * (1.0 - ssim(image, gt_image))
  becomes
    loss, Ll1 = l1_ssim_loss(image, gt_image, opt.lambda_dssim)

CUDA float32 only; there is no CPU path.  Like the reference's functions, both arguments are differentiable
(train_autoencoder.py:172 passes the network output SECOND): L1 and SSIM are symmetric in their arguments, so the
gradient of the second image is the same kernel pair run with the roles swapped.
"""
import ctypes

import torch

from diff_gaussian_rasterization import _lib


def _check(img, gt):
    if not (img.is_cuda and gt.is_cuda):
        raise RuntimeError("loss_utils: tensors must be CUDA tensors (there is no CPU path)")
    if img.shape != gt.shape or img.dim() not in (3, 4):
        raise RuntimeError("loss_utils: images must have the same (C,H,W) or (B,C,H,W) shape")


def _forward(img, gt, lam):
    """img, gt: contiguous float32 (C,H,W).  Returns (out3 tensor [L1, SSIM, loss], scratch holding the derivative
    maps for the gradient pass)."""
    C, H, W = img.shape
    dev = img.device
    out3 = torch.empty(3, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        nbytes = ctypes.c_size_t()
        _lib.check(_lib.lib.gsr_loss_scratch_bytes(C, H, W, ctypes.byref(nbytes)))
        scratch = torch.empty(nbytes.value, dtype=torch.uint8, device=dev)
        _lib.check(_lib.lib.gsr_l1_ssim_loss(C, H, W, img.data_ptr(), gt.data_ptr(), float(lam), out3.data_ptr(), None,
                                             scratch.data_ptr(), torch.cuda.current_stream(dev).cuda_stream))
    return out3, scratch


class _L1SSIM(torch.autograd.Function):
    """out3 = [L1, SSIM, (1-l)*L1 + l*(1-SSIM)]; only out3[2] carries a gradient (the others are reports).  The
    gradient pass runs in backward, scaled by the upstream gradient read on the device (no extra multiply)."""

    @staticmethod
    def forward(ctx, img, gt, lam):
        x = img.detach().to(torch.float32).contiguous()
        y = gt.detach().to(torch.float32).contiguous()
        ctx.shape = x.shape
        if x.dim() == 4:  # the mean over everything is the same with batch folded into channels
            x, y = x.reshape(-1, *x.shape[2:]), y.reshape(-1, *y.shape[2:])
        out3, scratch = _forward(x, y, lam)
        ctx.lam = lam
        ctx.save_for_backward(x, y, scratch)
        return out3

    @staticmethod
    def _grad(x, y, lam, up, scratch):
        """d loss / d x with `scratch` = the derivative maps of the forward pass over (x, y)."""
        C, H, W = x.shape
        grad = torch.empty_like(x)
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib.gsr_l1_ssim_backward(C, H, W, x.data_ptr(), y.data_ptr(), float(lam), up.data_ptr(),
                                                     scratch.data_ptr(), grad.data_ptr(),
                                                     torch.cuda.current_stream(x.device).cuda_stream))
        return grad

    @staticmethod
    def backward(ctx, g_out):
        x, y, scratch = ctx.saved_tensors
        up = g_out.to(torch.float32).contiguous()[2:3]
        gx = gy = None
        if ctx.needs_input_grad[0]:
            gx = _L1SSIM._grad(x, y, ctx.lam, up, scratch).reshape(ctx.shape)
        if ctx.needs_input_grad[1]:
            # the loss is symmetric: the second image's gradient is the first's with the roles swapped (its
            # derivative maps come from one more forward pass over (y, x))
            gy = _L1SSIM._grad(y, x, ctx.lam, up, _forward(y, x, ctx.lam)[1]).reshape(ctx.shape)
        return gx, gy, None


def l1_ssim_loss(image, gt_image, lambda_dssim=0.2):
    """(loss, Ll1): train.py:91-92 in one fused forward + backward.  `Ll1` is detached (train.py only logs it)."""
    _check(image, gt_image)
    out3 = _L1SSIM.apply(image, gt_image, float(lambda_dssim))
    return out3[2], out3[0].detach()


def l1_loss(network_output, gt):
    """utils/loss_utils.py:17-18."""
    _check(network_output, gt)
    return _L1SSIM.apply(network_output, gt, 0.0)[2]


def l2_loss(network_output, gt):
    """utils/loss_utils.py:20-21 (unused by the trainers; plain torch)."""
    return ((network_output - gt) ** 2).mean()


def ssim(img1, img2, window_size=11, size_average=True):
    """utils/loss_utils.py:33-63: 11x11 Gaussian window, sigma 1.5, zero padding, per channel."""
    if window_size != 11:
        raise RuntimeError("ssim: only the reference's window_size = 11 is built")
    _check(img1, img2)
    if size_average:
        return 1.0 - _L1SSIM.apply(img1, img2, 1.0)[2]
    if img1.dim() != 4:
        raise RuntimeError("ssim: size_average=False needs a (B,C,H,W) batch")  # the reference's .mean(1) x3
    return torch.stack([1.0 - _L1SSIM.apply(img1[b], img2[b], 1.0)[2] for b in range(img1.size(0))])
