"""Hierarchy files against the REFERENCE's own reader and writer (CPU only).

Two anchors: (1) golden files written by the unmodified GH/hierarchy_writer.cpp (tests/golden/ref_*.hier|dhier,
tests/golden/make_golden_hier.py) — always checked; (2) oracle/_ref/libref_hier.so (unmodified hierarchy_loader.cpp /
hierarchy_writer.cpp / traversal.cpp compiled against oracle/eigen_stub) run side by side, both directions — whenever that library is
present (the build container and, since oracle/_ref travels, the GPU box)."""
import filecmp
import os

import numpy as np
import pytest
import torch

import hier_cases
from gaussian_hierarchy import _C as gh
from oracle import refshim_hier as rh

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
need_ref = pytest.mark.skipif(not rh.available(), reason="oracle/_ref/libref_hier.so not built")
KEYS = ("pos", "shs", "alpha", "scales", "rot", "nodes", "boxes")


def _ours_to_np(t):
    pos, shs, alpha, scales, rot, nodes, boxes = t
    return {"pos": pos.numpy(), "shs": shs.reshape(shs.size(0), 48).numpy(), "alpha": alpha.reshape(-1).numpy(), "scales": scales.numpy(),
            "rot": rot.numpy(), "nodes": nodes.numpy(), "boxes": boxes.reshape(boxes.size(0), 8).numpy()}


def _write_ours(path, h, compressed):
    t = {k: torch.from_numpy(v) for k, v in h.items()}
    gh.write_hierarchy(path, t["pos"], t["shs"].view(-1, 16, 3), t["alpha"].view(-1, 1), t["scales"], t["rot"], t["nodes"],
                       t["boxes"].view(-1, 2, 4), compressed)


def _same_bits(a, b):
    return a.shape == b.shape and a.dtype == b.dtype and np.array_equal(a.view(np.uint32) if a.dtype == np.float32 else a,
                                                                        b.view(np.uint32) if b.dtype == np.float32 else b)


def _half(a):
    with np.errstate(over="ignore"):
        return a.astype(np.float16).astype(np.float32)


@pytest.mark.parametrize("name,compressed", [("ref_fp32.hier", False), ("ref_half.hier", True)])
def test_our_writer_reproduces_the_reference_file_byte_for_byte(tmp_path, name, compressed):
    f = str(tmp_path / name)
    _write_ours(f, hier_cases.static_case(), compressed)
    assert filecmp.cmp(f, os.path.join(GOLD, name), shallow=False)


@pytest.mark.parametrize("name,compressed", [("ref_fp32.hier", False), ("ref_half.hier", True)])
def test_our_loader_reads_the_reference_file(name, compressed):
    h = hier_cases.static_case()
    got = _ours_to_np(gh.load_hierarchy(os.path.join(GOLD, name)))
    for k in KEYS:
        want = h[k] if (not compressed or k in ("pos", "nodes")) else _half(h[k])
        assert _same_bits(got[k], want), k


@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_dynamic_files_both_ways_against_golden(tmp_path, deg):
    h = hier_cases.dynamic_case(deg)
    gold = os.path.join(GOLD, "ref_deg%d.dhier" % deg)
    f = str(tmp_path / "o.dhier")
    gh.write_dynamic_hierarchy(f, *(torch.from_numpy(h[k]) for k in ("pos", "shs", "alpha", "scales", "rot", "nodes")), deg)
    assert filecmp.cmp(f, gold, shallow=False)
    pos, shs, alpha, scales, rot, nodes = gh.load_dynamic_hierarchy(gold)
    assert shs.shape == (h["pos"].shape[0], (deg + 1) ** 2, 3)
    for got, k in ((pos, "pos"), (shs.reshape(shs.size(0), -1), "shs"), (alpha.reshape(-1), "alpha"), (scales, "scales"), (rot, "rot"),
                   (nodes, "nodes")):
        assert _same_bits(got.numpy(), h[k]), k


@need_ref
@pytest.mark.parametrize("compressed", [False, True])
@pytest.mark.parametrize("P,N", [(1, 1), (257, 513), (0, 0)])
def test_static_files_side_by_side_with_reference_code(tmp_path, compressed, P, N):
    h = hier_cases.static_case(P, N, seed=P + 3)
    fo, fr = str(tmp_path / "ours.hier"), str(tmp_path / "ref.hier")
    _write_ours(fo, h, compressed)
    rh.write(fr, h, compressed)
    assert filecmp.cmp(fo, fr, shallow=False)                      # both writers, same bytes
    theirs_of_ours, ours_of_theirs = rh.load(fo), _ours_to_np(gh.load_hierarchy(fr))   # their loader on our file, ours on theirs
    for k in KEYS:
        assert _same_bits(theirs_of_ours[k], ours_of_theirs[k]), k
        if not compressed:
            assert _same_bits(ours_of_theirs[k], h[k]), k


@need_ref
@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_dynamic_files_side_by_side_with_reference_code(tmp_path, deg):
    h = hier_cases.dynamic_case(deg, P=301, seed=40)
    fo, fr = str(tmp_path / "ours.dhier"), str(tmp_path / "ref.dhier")
    gh.write_dynamic_hierarchy(fo, *(torch.from_numpy(h[k]) for k in ("pos", "shs", "alpha", "scales", "rot", "nodes")), deg)
    rh.write_dynamic(fr, h, deg)
    assert filecmp.cmp(fo, fr, shallow=False)
    theirs = rh.load_dynamic(fo)
    assert theirs["sh_degree"] == deg
    pos, shs, alpha, scales, rot, nodes = gh.load_dynamic_hierarchy(fr)
    for got, k in ((pos, "pos"), (shs.reshape(shs.size(0), -1), "shs"), (alpha.reshape(-1), "alpha"), (scales, "scales"), (rot, "rot"),
                   (nodes, "nodes")):
        assert _same_bits(got.numpy(), theirs[k]) and _same_bits(theirs[k], h[k]), k


@need_ref
def test_expand_to_target_side_by_side_with_reference_code():
    import cases
    nodes, boxes, views = cases.make_cut_case(L=900, seed=4)
    depths = nodes[:, 0]
    for target in (-1, 0, 1, 2, int(np.median(depths)), int(depths.max()), int(depths.max()) + 2):
        assert np.array_equal(gh.expand_to_target(torch.from_numpy(nodes), target).numpy(), rh.expand_to_target(nodes, target)), target
