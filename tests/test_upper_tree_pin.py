"""Pins oracle/upper_tree.py (checker of lodgs_upper_tree_cut) to the reference: bit-identical to the reference's OWN
functions compiled out of scene/gaussian_model.py where /root/reference exists, and to their committed outputs
(tests/golden/upper_tree.npz) everywhere. CPU only."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
from make_golden_upper import CASES  # noqa: E402
from oracle import upper_tree as ut  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "upper_tree.npz")


def _same(a, b):
    a, b = np.asarray(a), np.asarray(b)
    return a.shape == b.shape and (np.array_equal(a.view(np.uint32), b.view(np.uint32)) if a.dtype == np.float32 else np.array_equal(a, b))


def _run(leaves, seed, vs, mult, cullit, fns=None):
    from lodgs.upper_tree import extract_frustum_planes
    t = {k: torch.from_numpy(v) for k, v in ut.make_upper_tree(leaves, seed).items()}
    campos, proj = ut.make_view(vs)
    planes = extract_frustum_planes(torch.from_numpy(proj))
    return ut.coarse_cut(t["nodes"], t["xyz"], t["min_dist_sq"], torch.from_numpy(campos), t["bounds"] if cullit else None, planes, mult, 0, fns), planes


@pytest.mark.parametrize("case", CASES)
def test_restatement_matches_golden(case):
    leaves, seed, views, mult, cullit = case
    gold = np.load(GOLD)
    for vs in views:
        r, planes = _run(leaves, seed, vs, mult, cullit)
        tag = "L%d_s%d_v%d_m%g_c%d_" % (leaves, seed, vs, mult, cullit)
        assert _same(planes.numpy(), gold[tag + "planes"])          # lodgs.upper_tree.extract_frustum_planes == the reference's
        for k, v in r.items():
            assert _same(v.numpy(), gold[tag + k]), (tag, k)


@pytest.mark.skipif(ut.reference_functions() is None, reason="/root/reference absent: the golden file stands in")
def test_restatement_matches_the_reference_functions():
    fns = ut.reference_functions()
    for leaves, seed, vs, mult, cullit in ((777, 21, 31, 1.0, True), (3000, 22, 32, 2.5, True), (1500, 23, 33, 1.0, False)):
        a, _ = _run(leaves, seed, vs, mult, cullit, fns)
        b, _ = _run(leaves, seed, vs, mult, cullit, None)
        for k in a:
            assert _same(a[k].numpy(), b[k].numpy()), k
        assert len(a["coarse_cut"]) > 0
