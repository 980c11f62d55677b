"""GPU: fused L1 + SSIM loss (csrc/ssim_loss.cu) against the reference's utils/loss_utils.py formulas (oracle/loss.py, pinned to
the module by tests/test_loss_pin.py) evaluated by torch on the same device: values <= 2e-6, gradients <= 1e-4 relative L2 (float
rounding of a separable against a 2-D window), at small ragged sizes and at 1080p."""
import pytest
import torch

from lodgs import loss as L
from oracle import loss as ol

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))


@pytest.mark.parametrize("C,H,W,seed", [(3, 64, 96, 0), (3, 83, 150, 1), (1, 7, 5, 2), (3, 1080, 1920, 3), (3, 16, 16, 4)])
def test_l1_ssim_loss_matches_reference_formulas(C, H, W, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    a, b = torch.rand(C, H, W, device="cuda", generator=g), torch.rand(C, H, W, device="cuda", generator=g)
    if seed == 4:
        b = a.clone()
        b[:, 3, 5] += 0.25          # identical images except one pixel: |x - y| = 0 almost everywhere (sign 0)
    x, xr = a.clone().requires_grad_(True), a.clone().requires_grad_(True)
    mine = L.l1_ssim_loss(x, b, 0.2)
    want = ol.training_loss(xr, b, 0.2)
    mine.backward(); want.backward()
    assert abs(float(mine) - float(want)) <= 2e-6
    assert rel_l2(x.grad, xr.grad) <= 1e-4
    # the pieces on their own, with the reference's call shapes
    s = L.fused_ssim(a.unsqueeze(0), b.unsqueeze(0))
    assert abs(float(s) - float(ol.ssim(a, b))) <= 2e-6 and abs(float(L.l1_loss(a, b)) - float(ol.l1_loss(a, b))) <= 1e-6
    xs, xsr = a.clone().requires_grad_(True), a.clone().requires_grad_(True)
    (2.0 * L.ssim(xs, b)).backward(); (2.0 * ol.ssim(xsr, b)).backward()
    assert rel_l2(xs.grad, xsr.grad) <= 1e-4
    # deterministic: a second evaluation gives the same bits
    assert torch.equal(L.l1_ssim_loss(a, b, 0.2), mine.detach())
