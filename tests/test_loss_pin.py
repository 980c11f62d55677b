"""Pins oracle/loss.py (checker of lodgs_l1_ssim_loss) to utils/loss_utils.py: bit-identical values and gradients on CPU where
/root/reference exists. CPU only."""
import pytest
import torch

from oracle import loss as ol


@pytest.mark.skipif(ol.load_reference() is None, reason="/root/reference absent")
@pytest.mark.parametrize("C,H,W,seed", [(3, 40, 56, 0), (1, 11, 11, 1), (3, 7, 5, 2)])
def test_restated_loss_equals_the_reference_module(C, H, W, seed):
    mod = ol.load_reference()
    g = torch.Generator().manual_seed(seed)
    a, b = torch.rand(C, H, W, generator=g), torch.rand(C, H, W, generator=g)
    x1, x2 = a.clone().requires_grad_(True), a.clone().requires_grad_(True)
    l1, l2 = ol.training_loss(x1, b, 0.2, mod), ol.training_loss(x2, b, 0.2, None)
    l1.backward(); l2.backward()
    assert torch.equal(l1, l2) and torch.equal(x1.grad, x2.grad)
    assert torch.equal(mod.ssim(a, b), ol.ssim(a, b)) and torch.equal(mod.l1_loss(a, b), ol.l1_loss(a, b))
