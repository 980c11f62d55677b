"""GPU: lodgs_upper_tree_cut (csrc/cut_upper.cu) against the outputs of the reference's own torch code
(tests/golden/upper_tree.npz) and the restatement pinned to it (oracle/upper_tree.py): cut ids, SPT ids / nodes / distances and direct
leaves bit-identical, in the reference's order; and the chain into get_spt_cut_cuda."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
from make_golden_upper import CASES  # noqa: E402
from lodgs import upper_tree as lut  # noqa: E402
from oracle import upper_tree as ut  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "upper_tree.npz")
KEYS = ("coarse_cut", "SPT_indices", "SPT_node_indices", "SPT_distances", "upper_tree_nodes_to_render")


def _bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint32) if a.dtype == np.float32 else a


def _gpu(t, campos, proj, mult, cullit):
    d = {k: torch.from_numpy(v).cuda() for k, v in t.items()}
    planes = lut.extract_frustum_planes(torch.from_numpy(proj).cuda())
    return lut.coarse_cut(d["nodes"], d["xyz"], d["min_dist_sq"], torch.from_numpy(campos).cuda(), d["bounds"] if cullit else None, planes, mult, 0)


@pytest.mark.parametrize("case", CASES)
def test_upper_cut_matches_reference_outputs(case):
    leaves, seed, views, mult, cullit = case
    gold = np.load(GOLD)
    t = ut.make_upper_tree(leaves, seed)
    for vs in views:
        campos, proj = ut.make_view(vs)
        r = _gpu(t, campos, proj, mult, cullit)
        tag = "L%d_s%d_v%d_m%g_c%d_" % (leaves, seed, vs, mult, cullit)
        for k in KEYS:
            got, want = r[k].cpu().numpy(), gold[tag + k]
            assert got.shape == want.shape and np.array_equal(_bits(got), _bits(want)), (tag, k)


@pytest.mark.parametrize("leaves,seed", [(60000, 41), (400000, 42)])
def test_upper_cut_large_vs_restatement(leaves, seed):
    t = ut.make_upper_tree(leaves, seed)
    tc = {k: torch.from_numpy(v) for k, v in t.items()}
    for vs, mult, cullit in ((51, 1.0, True), (52, 3.0, True), (53, 1.0, False)):
        campos, proj = ut.make_view(vs)
        r = _gpu(t, campos, proj, mult, cullit)
        planes = lut.extract_frustum_planes(torch.from_numpy(proj))
        w = ut.coarse_cut(tc["nodes"], tc["xyz"], tc["min_dist_sq"], torch.from_numpy(campos), tc["bounds"] if cullit else None, planes, mult, 0)
        for k in KEYS:
            assert np.array_equal(_bits(r[k].cpu().numpy()), _bits(w[k].numpy())), (vs, k)
        assert r["coarse_cut"].numel() > 100
        # and against the same torch code run where the reference runs it: on the device
        td = {k: v.cuda() for k, v in tc.items()}
        wd = ut.coarse_cut(td["nodes"], td["xyz"], td["min_dist_sq"], torch.from_numpy(campos).cuda(), td["bounds"] if cullit else None,
                           lut.extract_frustum_planes(torch.from_numpy(proj).cuda()), mult, 0)
        for k in KEYS:
            assert np.array_equal(_bits(r[k].cpu().numpy()), _bits(wd[k].cpu().numpy())), ("cuda torch", vs, k)


def test_upper_cut_feeds_the_spt_cut_without_leaving_the_device():
    """The lists of the upper cut are exactly what get_spt_cut_cuda takes (train_post.py:398): run the chain on the device."""
    import cases
    from gaussian_hierarchy import _C as gh
    ids, starts, spt_max, spt_min, sel, dist = cases.make_spt_case("ragged")
    n_tables = len(starts) - 1
    t = ut.make_upper_tree(800, 61, spt_share=0.9, zero_share=0.0)
    leaf = t["nodes"][:, 2] == 0
    has = leaf & (t["nodes"][:, 3] > 0)
    t["nodes"][has, 3] = (t["nodes"][has, 3] - 1) % n_tables          # valid table ids
    campos, proj = ut.make_view(62)
    r = _gpu(t, campos, proj, 1.0, True)
    S = r["SPT_indices"].numel()
    assert S > 0
    c = lambda a: torch.from_numpy(a).cuda()  # noqa: E731
    cut, counts = gh.get_spt_cut_cuda(S, c(ids), c(starts), c(spt_max), c(spt_min), r["SPT_indices"], r["SPT_distances"])
    from oracle import oracle as orc
    w_cut, w_counts = orc.get_spt_cut(ids, starts, spt_max, spt_min, r["SPT_indices"].cpu().numpy(), r["SPT_distances"].cpu().numpy())
    assert np.array_equal(cut.cpu().numpy(), w_cut) and np.array_equal(counts.cpu().numpy(), w_counts)
