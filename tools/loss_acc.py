"""Accuracy of the fused loss gradient at 1080p: ours and torch fp32 against torch fp64."""
import os, sys
import numpy as np, torch, torch.nn.functional as F
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gaussian-transformer_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
from utils.loss_utils import l1_ssim_loss
from test_gpu_loss import _torch_loss
torch.backends.cudnn.allow_tf32 = False
H, W = 1080, 1920
g = torch.Generator(device="cuda").manual_seed(5)
yy, xx = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
gt = torch.stack([0.5 + 0.4 * torch.sin(0.013 * xx * (c + 1) + 0.02 * yy) * torch.cos(0.007 * yy * (c + 2)) for c in range(3)]).float()
img = (gt + 0.05 * torch.randn(3, H, W, device="cuda", generator=g)).clamp(0, 1)
a = img.clone().requires_grad_(True); l, _ = l1_ssim_loss(a, gt, 0.2); l.backward()
b = img.clone().requires_grad_(True); w32, _, _ = _torch_loss(b[None], gt[None], 0.2); w32.backward()
c = img.double().requires_grad_(True)
import test_gpu_loss as T
w = torch.from_numpy(np.load(os.path.join(T.DATA, "loss_window.npy"))).cuda()
win = (w[:, None] @ w[None, :]).double().expand(3, 1, 11, 11).contiguous()
conv = lambda t: F.conv2d(t, win, padding=5, groups=3)
x, y = c[None], gt.double()[None]
mu1, mu2 = conv(x), conv(y)
s11, s22, s12 = conv(x*x)-mu1*mu1, conv(y*y)-mu2*mu2, conv(x*y)-mu1*mu2
ss = (((2*mu1*mu2+1e-4)*(2*s12+9e-4))/((mu1*mu1+mu2*mu2+1e-4)*(s11+s22+9e-4))).mean()
w64 = 0.8*(x-y).abs().mean() + 0.2*(1-ss); w64.backward()
m = c.grad.abs().max().item()
print("loss ours %.9f torch32 %.9f torch64 %.9f" % (l.item(), w32.item(), w64.item()))
print("grad max %.3e  ours-64 %.3e (%.2e of max)  torch32-64 %.3e (%.2e of max)" % (m, (a.grad-c.grad).abs().max().item(), (a.grad-c.grad).abs().max().item()/m, (b.grad-c.grad).abs().max().item(), (b.grad-c.grad).abs().max().item()/m))
