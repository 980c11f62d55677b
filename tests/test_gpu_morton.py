"""GPU: Morton codes (gaussian_hierarchy._C.get_morton_indices) against the CPU restatement, the committed reference outputs
and the reference kernel run side by side — bit-exact."""
import os

import numpy as np
import pytest
import torch

import cases
from gaussian_hierarchy import _C as gh
from oracle import oracle as orc
from oracle import refshim

pytestmark = pytest.mark.gpu
DEV = "cuda"
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "morton.npz")


def _mine(xyz, mn, mx):
    codes = torch.full((xyz.shape[0],), -1, dtype=torch.int64, device=DEV)
    gh.get_morton_indices(torch.from_numpy(xyz).to(DEV), torch.from_numpy(mn).to(DEV), torch.from_numpy(mx).to(DEV), codes)
    return codes.cpu().numpy()


@pytest.mark.parametrize("name,deg", [("plain", False), ("flat_z", True)])
def test_morton_bit_exact(name, deg):
    xyz, mn, mx = cases.make_morton_case(degenerate=deg)
    mine = _mine(xyz, mn, mx)
    assert np.array_equal(mine, orc.get_morton_codes(xyz, mn, mx))
    assert np.array_equal(mine, np.load(GOLD)[name])
    assert mine.min() >= 0 and len(np.unique(mine)) > (100 if deg else 10000)
    if refshim.available():
        t = [torch.from_numpy(a).to(DEV) for a in (xyz, mn, mx)]
        ref = torch.zeros(xyz.shape[0], dtype=torch.int64, device=DEV)
        torch.cuda.synchronize()
        refshim.lib().ref_get_morton_code(t[0].data_ptr(), t[1].data_ptr(), t[2].data_ptr(), ref.data_ptr(), xyz.shape[0])
        torch.cuda.synchronize()
        assert np.array_equal(mine, ref.cpu().numpy())


def test_morton_points_outside_the_box_and_order():
    """The training script passes the extent of a SUBSET (skybox points excluded), so points outside the box occur: negative and
    > 2^21 scaled coordinates keep only their low 21 bits (two's complement), like the reference's shift loop."""
    rng = np.random.default_rng(2)
    xyz = rng.uniform(-3, 3, (5000, 3)).astype(np.float32)
    mn, mx = np.array([-1, -1, -1], np.float32), np.array([1, 2, 1.5], np.float32)
    mine = _mine(xyz, mn, mx)
    assert np.array_equal(mine, orc.get_morton_codes(xyz, mn, mx))
    # inside the box the code is monotone in each axis' quantised coordinate: sorting by code groups points by octant
    inside = np.all((xyz >= mn) & (xyz < mx), 1)
    q = ((xyz[inside] - mn) / (mx - mn) * 2.0).astype(np.int64)  # top bit per axis
    top = mine[inside] >> 60
    assert np.array_equal(top, q[:, 0] | q[:, 1] << 1 | q[:, 2] << 2)
    # P = 0 is a no-op
    gh.get_morton_indices(torch.empty(0, 3, device=DEV), torch.from_numpy(mn).to(DEV), torch.from_numpy(mx).to(DEV),
                          torch.empty(0, dtype=torch.int64, device=DEV))
