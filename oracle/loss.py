"""TEST INFRASTRUCTURE ONLY — the reference's image loss as a checker: utils/loss_utils.py (l1_loss :17-18, ssim :33-63) imported by file path
where /root/reference exists, and restated op by op (same torch calls, so bit-identical on one device) for the GPU box. `fused_ssim`, which
train_post.py:559 calls instead of `ssim`, is an external package absent from the reference tree and from this image; it computes this SSIM."""
import importlib.util
import os
from math import exp

import torch
import torch.nn.functional as F

REF_PATH = "/root/reference/utils/loss_utils.py"


def load_reference():
    if not os.path.exists(REF_PATH):
        return None
    spec = importlib.util.spec_from_file_location("_ref_loss_utils", REF_PATH)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def l1_loss(network_output, gt):
    return torch.abs((network_output - gt)).mean()


def ssim(img1, img2, window_size=11):
    channel = img1.size(-3)
    gauss = torch.Tensor([exp(-(x - window_size // 2) ** 2 / float(2 * 1.5 ** 2)) for x in range(window_size)])   # :23-25
    g1 = (gauss / gauss.sum()).unsqueeze(1)
    window = g1.mm(g1.t()).float().unsqueeze(0).unsqueeze(0).expand(channel, 1, window_size, window_size).contiguous().to(img1)   # :27-31
    pad = window_size // 2
    mu1 = F.conv2d(img1, window, padding=pad, groups=channel)                                                   # :43-63
    mu2 = F.conv2d(img2, window, padding=pad, groups=channel)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    sigma1_sq = F.conv2d(img1 * img1, window, padding=pad, groups=channel) - mu1_sq
    sigma2_sq = F.conv2d(img2 * img2, window, padding=pad, groups=channel) - mu2_sq
    sigma12 = F.conv2d(img1 * img2, window, padding=pad, groups=channel) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim_map = ((2 * mu1_mu2 + C1) * (2 * sigma12 + C2)) / ((mu1_sq + mu2_sq + C1) * (sigma1_sq + sigma2_sq + C2))
    return ssim_map.mean()


def training_loss(image, gt, lambda_dssim, mod=None):
    """train_post.py:558-559 with `ssim` standing in for the external fused_ssim."""
    l1 = (mod.l1_loss if mod else l1_loss)(image, gt)
    s = (mod.ssim if mod else ssim)(image, gt)
    return (1.0 - lambda_dssim) * l1 + lambda_dssim * (1.0 - s)
