"""Image losses of the training loops on the device (csrc/ssim_loss.cu): `fused_ssim(img1, img2)` with the call shape of the external package
the reference imports (train_post.py:29,559: `fused_ssim(image.unsqueeze(0), gt_image.unsqueeze(0))`, gradient to img1 only), `ssim` /
`l1_loss` with the signatures of utils/loss_utils.py, and `l1_ssim_loss(image, gt, lambda_dssim)` = the whole training loss in one
forward and one backward kernel. No CPU fallback."""
import torch

from . import check, current_stream, lib

_ws = {}


def _run(img, gt, a_l1, a_ssim, bias, want_grad):
    if img.device.type != "cuda":
        raise RuntimeError("lodgs.loss needs CUDA tensors (there is no CPU path)")
    x = img.detach().to(torch.float32).contiguous()
    y = gt.detach().to(device=x.device, dtype=torch.float32).contiguous()
    if x.shape != y.shape or x.dim() < 2:
        raise RuntimeError("lodgs.loss: images must have the same [.., H, W] shape")
    H, W = x.shape[-2], x.shape[-1]
    C = x.numel() // (H * W)
    need = int(lib.lodgs_l1_ssim_workspace_bytes(C, H, W))
    ws = _ws.get(x.device.index)
    if ws is None or ws.numel() < need:
        ws = torch.empty(need, dtype=torch.uint8, device=x.device)
        _ws[x.device.index] = ws
    out = torch.empty(3, device=x.device)
    grad = torch.empty_like(x) if want_grad else None
    check(lib.lodgs_l1_ssim_loss(x.data_ptr(), y.data_ptr(), C, H, W, float(a_l1), float(a_ssim), float(bias), out.data_ptr(),
                                 grad.data_ptr() if want_grad else None, ws.data_ptr(), ws.numel(), current_stream()))
    return out, grad


class _Loss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, img, gt, a_l1, a_ssim, bias):
        out, grad = _run(img, gt, a_l1, a_ssim, bias, img.requires_grad)
        ctx.save_for_backward(grad if grad is not None else torch.empty(0, device=img.device))
        ctx.shape = img.shape
        return out[0]

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return (grad.view(ctx.shape) * g if grad.numel() else None), None, None, None, None


def l1_ssim_loss(image, gt, lambda_dssim=0.2):
    """(1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt)): train_post.py:558-559."""
    return _Loss.apply(image, gt, 1.0 - lambda_dssim, -lambda_dssim, lambda_dssim)


def fused_ssim(img1, img2, padding="same", train=True):
    """Mean SSIM of img1 against img2 ([B,C,H,W] or [C,H,W]); differentiable in img1. Every image of a batch is its own set of channels:
    the window never mixes images or channels, so a batch is C*B planes."""
    if padding != "same":
        raise NotImplementedError("fused_ssim: only the zero 'same' padding the training loops use")
    return _Loss.apply(img1, img2, 0.0, 1.0, 0.0)


def ssim(img1, img2, window_size=11, size_average=True):
    """utils/loss_utils.py:33-41."""
    if window_size != 11 or not size_average:
        raise NotImplementedError("ssim: window 11 with size_average, as the training loops call it")
    return _Loss.apply(img1, img2, 0.0, 1.0, 0.0)


def l1_loss(network_output, gt):
    """utils/loss_utils.py:17-18."""
    return _Loss.apply(network_output, gt, 1.0, 0.0, 0.0)
