"""TEST INFRASTRUCTURE ONLY (oracle): torch restatement of the trainer's image loss, for checking csrc/ssim.cu.

What it restates: utils/loss_utils.py of the reference -- ``l1_loss`` (:18-19), the 11-tap Gaussian window of sigma 1.5
(``gaussian`` :31-33, ``create_window`` :36-40), the SSIM map and its mean (``_ssim`` :54-77) -- and their combination at
train_rig.py:508-509.  Pinned against outputs of the reference's own module (tests/golden/loss_ssim.npz, made by
tests/golden/make_loss_golden.py).  Device- and dtype-agnostic; gradients come from torch autograd.  The five windowed
moments are produced by ONE depthwise convolution over a stacked tensor (same arithmetic per plane as five separate calls).
"""
import torch
import torch.nn.functional as F

C1, C2 = 0.01 ** 2, 0.03 ** 2


def window_1d(taps: int = 11, sigma: float = 1.5) -> torch.Tensor:
    """fp32 taps exp(-(i - taps//2)^2 / (2 sigma^2)), normalised (the reference builds them in fp32 as well)."""
    offs = torch.arange(taps, dtype=torch.float64) - taps // 2
    g = torch.tensor([float(torch.exp(-o * o / (2.0 * sigma * sigma))) for o in offs], dtype=torch.float32)
    return g / g.sum()


def windowed_moments(x: torch.Tensor, y: torch.Tensor, taps: int = 11):
    """(w*x, w*y, w*x^2, w*y^2, w*xy) with zero padding taps//2, per (H, W) plane; x, y: (..., H, W)."""
    g = window_1d(taps)
    w2 = torch.outer(g, g).float().to(device=x.device, dtype=x.dtype)  # fp32 outer product, then the image's dtype
    planes = torch.stack([x, y, x * x, y * y, x * y], 0).reshape(1, -1, x.shape[-2], x.shape[-1])
    k = w2.expand(planes.shape[1], 1, taps, taps).contiguous()
    out = F.conv2d(planes, k, padding=taps // 2, groups=planes.shape[1])
    return out.reshape(5, *x.shape)


def ssim(img1: torch.Tensor, img2: torch.Tensor, window_size: int = 11) -> torch.Tensor:
    mu1, mu2, xx, yy, xy = windowed_moments(img1, img2, window_size)
    var1, var2, cov = xx - mu1 * mu1, yy - mu2 * mu2, xy - mu1 * mu2
    smap = ((2 * mu1 * mu2 + C1) * (2 * cov + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (var1 + var2 + C2))
    return smap.mean()


def l1_loss(network_output: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    return (network_output - gt).abs().mean()


def image_loss(image: torch.Tensor, gt: torch.Tensor, lambda_dssim: float):
    """-> (loss, l1, ssim) with loss = (1 - lambda) l1 + lambda (1 - ssim)"""
    l1, s = l1_loss(image, gt), ssim(image, gt)
    return (1.0 - lambda_dssim) * l1 + lambda_dssim * (1.0 - s), l1, s
