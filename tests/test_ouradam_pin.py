"""Pins oracle/ouradam.py (the checker of the sparse Adam kernel and the optimizer of bench.py's reference arm) to the
reference's scene/OurAdam.py: bit-identical to the imported class where /root/reference exists, bit-identical to the golden
outputs that class produced (tests/golden/ouradam.npz, tests/golden/make_golden_adam.py) everywhere. CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import ouradam

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ouradam.npz")
CASES = [(257, 1e-15, 11), (257, 1e-8, 12)]


@pytest.mark.parametrize("rows,eps,seed", CASES)
def test_restatement_matches_golden_bitwise(rows, eps, seed):
    gold = np.load(GOLD)
    res = ouradam.run_scenario(ouradam.RestatedAdam, rows, eps, seed)
    for k, (p, m, v) in res.items():
        tag = "r%d_s%d_%s" % (rows, seed, k)
        for arr, suffix in ((p, "_p"), (m, "_m"), (v, "_v")):
            assert np.array_equal(arr.numpy().view(np.uint32), gold[tag + suffix].view(np.uint32)), tag + suffix


@pytest.mark.skipif(ouradam.load_reference() is None, reason="/root/reference absent (GPU box): the golden file stands in")
@pytest.mark.parametrize("rows,eps,seed", [(64, 1e-15, 3), (513, 1e-8, 4), (1000, 1e-15, 5)])
def test_restatement_matches_imported_reference_bitwise(rows, eps, seed):
    ref = ouradam.load_reference()
    a = ouradam.run_scenario(ref.Adam, rows, eps, seed)
    b = ouradam.run_scenario(ouradam.RestatedAdam, rows, eps, seed)
    for k in a:
        for x, y in zip(a[k], b[k]):
            assert torch.equal(x, y), k


def test_empty_relevant_updates_every_row():
    """OurAdam.py:214: `relevant.size(0) == 0` selects the all-rows form; the last step of the scenario relies on it."""
    params, steps = ouradam.scenario(32, 1e-15, 1)
    p = params["xyz"]
    opt = ouradam.RestatedAdam([{"params": [p], "lr": 1e-2}], eps=1e-15)
    before = p.detach().clone()
    p.grad = torch.ones_like(p)
    opt.step(torch.empty(0, dtype=torch.int64))
    assert (p.detach() != before).all()
