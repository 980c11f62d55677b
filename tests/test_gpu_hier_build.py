"""GPU: hierarchy construction (csrc/hier_build.cu) against the reference's own CPU creator — the UNMODIFIED
GH/mainHierarchyCreator.cpp pipeline compiled against oracle/eigen_stub (oracle/_ref/GaussianHierarchyCreator) — run on the same synthetic
PLY: node tables bit-identical (topology, pre-order ids, leaf assignment), leaf Gaussians bit-identical up to the alignment's axis
relabelling, merged Gaussians (parity UNPINNED in their last bits: Eigen's eigen-solver is replaced by a Jacobi iteration on both sides,
and gcc contracts the creator's float expressions as it pleases) to 1e-4 of their scale; plus the static-format ids against a literal
restatement of upstream populateRec."""
import os

import numpy as np
import pytest
import torch

from gaussian_hierarchy import _C as gh
from lodgs import hier_build
from oracle import ply as oply
from oracle import refshim_hier as rh

pytestmark = pytest.mark.gpu


def _cov(scales, rot):
    w, x, y, z = rot[:, 0], rot[:, 1], rot[:, 2], rot[:, 3]
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                     2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                     2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], 1).view(-1, 3, 3).double()
    return R @ torch.diag_embed(scales.double() ** 2) @ R.transpose(1, 2)


def _creator_activation(pos, shs12, op, ls, rot):
    """What GH/loader.cpp:105-124 computes for a LessRichPoint, with the SAME libm the creator binary calls (expf through ctypes) and
    float32 operations in its order, so that both sides start from identical bits."""
    import ctypes
    libm = ctypes.CDLL("libm.so.6")
    libm.expf.restype, libm.expf.argtypes = ctypes.c_float, [ctypes.c_float]
    expf = lambda a: np.array([libm.expf(float(x)) for x in a.ravel()], np.float32).reshape(a.shape)  # noqa: E731
    one = np.float32(1.0)
    opacity = one / (one + expf(-op.reshape(-1)))
    scales = expf(ls)
    sq = rot[:, 0] * rot[:, 0]
    for c in (1, 2, 3):
        sq = sq + rot[:, c] * rot[:, c]
    rn = rot / np.sqrt(sq)[:, None]
    shs = np.zeros((pos.shape[0], 48), np.float32)
    shs[:, :3] = shs12[:, :3]
    return pos, scales, rn.astype(np.float32), opacity.astype(np.float32), shs


@pytest.mark.skipif(not rh.creator_available(), reason="oracle/_ref/GaussianHierarchyCreator not built")
@pytest.mark.parametrize("n,seed", [(1, 0), (2, 1), (3, 2), (257, 3), (5000, 4), (40001, 5)])
def test_build_matches_reference_creator(tmp_path, n, seed):
    pts = oply.synth_points(n, seed=seed)
    plyf = str(tmp_path / "in.ply")
    oply.write_less_rich_ply(plyf, *pts)
    rh.run_creator(plyf, str(tmp_path) + "/ref_")
    r_pos, r_shs, r_op, r_ls, r_rot, r_nodes = gh.load_dynamic_hierarchy(str(tmp_path) + "/ref_hierarchy.dhier")
    h = hier_build.build_hierarchy(*(torch.from_numpy(a).cuda() for a in _creator_activation(*pts)), dynamic=True, static=False)
    torch.cuda.synchronize()
    N = 2 * n - 1
    assert r_nodes.shape == (N, 6) and h["dyn_nodes"].shape == (N, 6)
    assert torch.equal(h["dyn_nodes"].cpu(), r_nodes)                  # topology, ids, sibling links, leaf -> input Gaussian
    pos, op, shs, rot, sc = h["pos"].cpu(), h["opacities"].cpu(), h["shs"].cpu(), h["rotations"].cpu(), h["scales"].cpu()
    bits = lambda t: t.contiguous().view(torch.int32)  # noqa: E731
    # kernels compiled without FMA contraction evaluate the creator's expressions in its order: the merged Gaussians come out
    # BIT-IDENTICAL to the creator binary's (both sides use the same Jacobi eigen-solver, see oracle/eigen_stub)
    assert torch.equal(bits(pos), bits(r_pos)), float((pos - r_pos).abs().max())
    assert torch.equal(bits(op), bits(r_op.reshape(-1))), float((op - r_op.reshape(-1)).abs().max())
    assert torch.equal(bits(shs), bits(r_shs.reshape(N, 48)))
    assert torch.equal(bits(rot), bits(r_rot)), float((rot - r_rot).abs().max())
    assert torch.allclose(torch.log(sc), r_ls, rtol=0, atol=2e-6)       # the file holds logf(scale): libm vs torch, 1 ulp apart


@pytest.mark.skipif(not rh.creator_available(), reason="oracle/_ref/GaussianHierarchyCreator not built")
@pytest.mark.parametrize("n,seed", [(3000, 9), (100000, 10)])
def test_create_dynamic_hierarchy_writes_the_creators_file(tmp_path, n, seed):
    """The product path from a PLY (tools/create_hierarchy.py: host activation with the creator's libm calls, GPU build, .dhier writer)
    against the creator binary run on the same PLY: the two hierarchy.dhier files are byte-identical."""
    import filecmp
    plyf = str(tmp_path / "in.ply")
    oply.write_less_rich_ply(plyf, *oply.synth_points(n, seed=seed))
    rh.run_creator(plyf, str(tmp_path) + "/ref_")
    hier_build.create_dynamic_hierarchy(plyf, str(tmp_path) + "/ours_")
    torch.cuda.synchronize()
    assert filecmp.cmp(str(tmp_path) + "/ours_hierarchy.dhier", str(tmp_path) + "/ref_hierarchy.dhier", shallow=False)


def _upstream_populate(dyn):
    """Writer::makeHierarchy with the upstream child emission (GH/writer.cpp:21-96), restated literally on the dynamic
    node table: ids allocated in blocks when a node is visited; -> nodes[N,7]."""
    import sys
    sys.setrecursionlimit(10000)
    N = dyn.shape[0]
    kids = [[] for _ in range(N)]
    for i in range(1, N):
        kids[int(dyn[i, 1])].append(i)
    height = np.zeros(N, np.int64)
    for i in range(N - 1, -1, -1):
        if kids[i]:
            height[i] = max(height[c] for c in kids[i]) + 1
    out = [[-1, -1, 0, 0, 0, 0, 0]]
    count = [0]

    def rec(t, nid):
        out[nid][2] = count[0]
        count[0] += 1
        out[nid][3] = 0 if kids[t] else 1
        out[nid][4] = 1 if kids[t] else 0
        out[nid][5] = len(out)
        for _ in kids[t]:
            out.append([-1, nid, 0, 0, 0, 0, 0])
        out[nid][6] = len(kids[t])
        out[nid][0] = int(height[t])
        for k, c in enumerate(kids[t]):
            rec(c, out[nid][5] + k)
    rec(0, 0)
    return np.array(out, np.int32)


@pytest.mark.parametrize("n", [1, 2, 7, 64, 1000, 4097])
def test_static_format_ids_follow_upstream_populate_rec(n):
    pos, shs, op, ls, rot = oply.synth_points(n, seed=n)
    t = lambda a: torch.from_numpy(a).cuda()  # noqa: E731
    rt = t(rot)
    h = hier_build.build_hierarchy(t(pos), torch.exp(t(ls)), rt / rt.norm(dim=1, keepdim=True), torch.sigmoid(t(op)).reshape(-1), t(shs))
    dyn, nodes, boxes = h["dyn_nodes"].cpu().numpy(), h["nodes"].cpu().numpy(), h["boxes"].cpu().numpy()
    want = _upstream_populate(dyn)
    assert np.array_equal(nodes, want)
    # boxes: children inside parents, w = longest side for inner nodes and 0 for leaves: the hierarchy passes the traversal check
    inner = nodes[:, 6] > 0
    side = (boxes[:, 1, :3] - boxes[:, 0, :3]).max(1)
    assert np.array_equal(boxes[inner, 0, 3], side[inner]) and np.all(boxes[~inner, 0, 3] == 0) and np.array_equal(boxes[:, 0, 3], boxes[:, 1, 3])
    if n > 1:
        assert gh.prepare_hierarchy(h["nodes"], h["boxes"]) is True
    # every leaf box is its Gaussian's +-3 max-scale cube
    g = nodes[~inner, 2]
    P, S = h["pos"].cpu().numpy(), h["scales"].cpu().numpy()
    r = 3.0 * S[g].max(1, keepdims=True)
    assert np.array_equal(boxes[~inner, 0, :3], P[g] - r) and np.array_equal(boxes[~inner, 1, :3], P[g] + r)
