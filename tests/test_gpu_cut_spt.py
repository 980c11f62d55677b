"""GPU: the SPT cut (gaussian_hierarchy._C.get_spt_cut_cuda) against the literal CPU restatement, the committed reference outputs
and the reference's Switching::getSPTCut run side by side — ids, order and count prefix bit-exact."""
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
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cut_spt.npz")


def _mine(arrs):
    t = [torch.from_numpy(a).to(DEV) for a in arrs]
    cut, counts = gh.get_spt_cut_cuda(len(arrs[4]), t[0], t[1], t[2], t[3], t[4], t[5])
    assert cut.dtype == torch.int32 and counts.dtype == torch.int32 and cut.is_cuda and counts.is_cuda
    return cut.cpu().numpy(), counts.cpu().numpy()


@pytest.mark.parametrize("name", list(cases.SPT_CASES))
def test_spt_cut_bit_exact_vs_oracle_golden_and_reference(name):
    arrs = cases.make_spt_case(name)
    cut, counts = _mine(arrs)
    o_cut, o_counts = orc.get_spt_cut(*arrs)
    assert len(o_cut) > 0
    assert np.array_equal(cut, o_cut) and np.array_equal(counts, o_counts)
    g = np.load(GOLD)
    assert np.array_equal(cut, g[name + "_cut"]) and np.array_equal(counts, g[name + "_counts"])
    if refshim.available():
        r_cut, r_counts = refshim.get_spt_cut(*[torch.from_numpy(a).to(DEV) for a in arrs])
        assert np.array_equal(cut, r_cut.cpu().numpy()) and np.array_equal(counts, r_counts.cpu().numpy())


def test_spt_cut_edges():
    ids, starts, mx, mn, sel, dist = cases.make_spt_case("all_tables")
    t = [torch.from_numpy(a).to(DEV) for a in (ids, starts, mx, mn)]
    # no slot selected: the training script never calls the op then; the op returns two empty tensors
    cut, counts = gh.get_spt_cut_cuda(0, *t, torch.empty(0, dtype=torch.int32, device=DEV), torch.empty(0, device=DEV))
    assert cut.numel() == 0 and counts.numel() == 0
    # only empty tables selected -> nothing to evaluate anywhere
    empty = np.nonzero(np.diff(starts) == 0)[0].astype(np.int32)
    assert len(empty) >= 2
    d = np.full(len(empty), 5.0, np.float32)
    cut, counts = _mine((ids, starts, mx, mn, empty, d))
    o_cut, o_counts = orc.get_spt_cut(ids, starts, mx, mn, empty, d)
    assert cut.size == 0 and o_cut.size == 0 and np.array_equal(counts, o_counts) and not counts.any()
    # empty tables first and last in the selection; the same table selected twice
    full = np.nonzero(np.diff(starts) > 3)[0].astype(np.int32)
    sel2 = np.concatenate([empty[:1], full[:5], full[2:4], empty[1:2]]).astype(np.int32)
    d2 = np.linspace(3.0, 45.0, len(sel2)).astype(np.float32)
    cut, counts = _mine((ids, starts, mx, mn, sel2, d2))
    o_cut, o_counts = orc.get_spt_cut(ids, starts, mx, mn, sel2, d2)
    assert np.array_equal(cut, o_cut) and np.array_equal(counts, o_counts) and cut.size > 0
    # the last table of the arrays fully selected and no padding entry behind it: the one-past-the-end probe is out of bounds
    last = np.array([len(starts) - 2], np.int32)
    n = int(starts[-1])
    assert starts[-1] > starts[-2]
    cut, counts = _mine((ids[:n], starts, mx[:n], mn[:n], last, np.zeros(1, np.float32)))
    o_cut, o_counts = orc.get_spt_cut(ids[:n], starts, mx[:n], mn[:n], last, np.zeros(1, np.float32))
    assert np.array_equal(cut, o_cut) and np.array_equal(counts, o_counts)


def test_spt_cut_rejects_cpu_tensors():
    arrs = [torch.from_numpy(a) for a in cases.make_spt_case("small")]
    with pytest.raises(RuntimeError, match="CUDA"):
        gh.get_spt_cut_cuda(len(arrs[4]), *arrs)
