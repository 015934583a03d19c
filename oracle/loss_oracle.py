"""TEST INFRASTRUCTURE ONLY (oracle): torch restatement of the trainer's image loss, for checking csrc/ssim.cu.

Follows utils/loss_utils.py of the reference: ``l1_loss`` (:18-19), ``gaussian`` (:31-33), ``create_window`` (:36-40),
``_ssim`` (:54-77), and the combination at train_rig.py:508-509.  Pinned against outputs of the reference's own module
(tests/golden/loss_ssim.npz, made by tests/golden/make_loss_golden.py).  Device- and dtype-agnostic; gradients through
torch autograd.
"""
from math import exp

import torch
import torch.nn.functional as F


def window(channel: int, dtype, device, window_size: int = 11, sigma: float = 1.5) -> torch.Tensor:
    g = torch.tensor([exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(window_size)],
                     dtype=torch.float32)
    g = (g / g.sum()).unsqueeze(1)
    w2 = g.mm(g.t()).float()[None, None]
    return w2.expand(channel, 1, window_size, window_size).contiguous().to(device=device, dtype=dtype)


def ssim(img1: torch.Tensor, img2: torch.Tensor, window_size: int = 11) -> torch.Tensor:
    ch = img1.size(-3)
    w = window(ch, img1.dtype, img1.device, window_size)
    a, b = (img1, img2) if img1.dim() == 4 else (img1[None], img2[None])
    pad = window_size // 2
    mu1 = F.conv2d(a, w, padding=pad, groups=ch)
    mu2 = F.conv2d(b, w, padding=pad, groups=ch)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    sigma1_sq = F.conv2d(a * a, w, padding=pad, groups=ch) - mu1_sq
    sigma2_sq = F.conv2d(b * b, w, padding=pad, groups=ch) - mu2_sq
    sigma12 = F.conv2d(a * b, w, padding=pad, groups=ch) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim_map = ((2 * mu1_mu2 + C1) * (2 * sigma12 + C2)) / ((mu1_sq + mu2_sq + C1) * (sigma1_sq + sigma2_sq + C2))
    return ssim_map.mean()


def l1_loss(network_output: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    return torch.abs(network_output - gt).mean()


def image_loss(image: torch.Tensor, gt: torch.Tensor, lambda_dssim: float):
    """-> (loss, l1, ssim)"""
    l1 = l1_loss(image, gt)
    s = ssim(image, gt)
    return (1.0 - lambda_dssim) * l1 + lambda_dssim * (1.0 - s), l1, s
