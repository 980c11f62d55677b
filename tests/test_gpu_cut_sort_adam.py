"""GPU: LOD cut, interpolation weights, onesweep sort and sparse Adam through the C ABI, checked against the oracle,
the golden vectors of the reference build, the reference CUDA build itself and CUB (torch.sort) respectively."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

import cases
import lodgs
from gaussian_hierarchy import _C as gh
from lodgs import synth
from oracle import oracle as orc
from oracle import refshim

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
DEV = "cuda"


def _cut(nodes_t, boxes_t, tau, v, mode):
    N = nodes_t.size(0)
    ri, pi, ni = (torch.full((N,), -7, dtype=torch.int32, device=DEV) for _ in range(3))
    K = gh.expand_to_size(nodes_t, boxes_t, float(tau), torch.from_numpy(v).to(DEV), torch.zeros(3), ri, pi, ni, mode=mode)
    return K, ri, pi, ni


@pytest.mark.parametrize("mode", [0, 1])
def test_cut_bit_exact_vs_golden_and_oracle(mode):
    g = np.load(os.path.join(GOLD, "cut.npz"))
    nodes, boxes, views = cases.make_cut_case()
    n_t, b_t = torch.from_numpy(nodes).to(DEV), torch.from_numpy(boxes).to(DEV)
    if mode == 1:
        assert gh.prepare_hierarchy(n_t, b_t), "synthetic kd-tree must be traversable"
    for vi, v in enumerate(views):
        for tau_px in (0, 3, 6, 15, 60):
            tau = np.float32(synth.tau_to_threshold(tau_px, 0.8, 1920))
            K, ri, pi, ni = _cut(n_t, b_t, tau, v, mode)
            key = "v%d_t%d" % (vi, tau_px)
            assert K == len(g[key + "_ri"]), key
            assert np.array_equal(ri[:K].cpu().numpy(), g[key + "_ri"]), key
            assert np.array_equal(pi[:K].cpu().numpy(), g[key + "_pi"]), key
            assert np.array_equal(ni[:K].cpu().numpy(), g[key + "_ni"]), key
            assert (ri[K:] == -7).all()  # nothing written past K
            ts, kids = torch.zeros(K, device=DEV), torch.zeros(K, dtype=torch.int32, device=DEV)
            gh.get_interpolation_weights(ni[:K], float(tau), n_t, b_t, torch.from_numpy(v), torch.zeros(3), ts, kids)
            assert np.array_equal(kids.cpu().numpy(), g[key + "_kids"])
            assert np.array_equal(ts.cpu().numpy().view(np.uint32), g[key + "_ts"].view(np.uint32))


@pytest.mark.parametrize("mode", [0, 1])
def test_cut_large_random_vs_oracle_and_reference(mode):
    nodes, boxes, _ = cases.make_cut_case(L=60000, seed=4)
    n_t, b_t = torch.from_numpy(nodes).to(DEV), torch.from_numpy(boxes).to(DEV)
    N = nodes.shape[0]
    rng = np.random.default_rng(9)
    for trial in range(6):
        v = (rng.uniform(-1, 1, 3) * np.array([60, 60, 30])).astype(np.float32)
        tau = np.float32(synth.tau_to_threshold(float(rng.choice([0, 3, 6, 15, 40, 200])), 0.8, 1920))
        K, ri, pi, ni = _cut(n_t, b_t, tau, v, mode)
        o_ri, o_pi, o_ni = orc.expand_to_size(nodes, boxes, tau, v)
        assert K == len(o_ri)
        assert np.array_equal(ri[:K].cpu().numpy(), o_ri) and np.array_equal(pi[:K].cpu().numpy(), o_pi)
        assert np.array_equal(ni[:K].cpu().numpy(), o_ni)
        assert (np.diff(o_ni) > 0).all()  # ascending node ids: the order the stable sort downstream depends on
        if refshim.available():
            r_ri, r_pi, r_ni = (torch.zeros(N, dtype=torch.int32, device=DEV) for _ in range(3))
            vp = torch.from_numpy(v).to(DEV)
            Kr = refshim.lib().ref_expand_to_size(N, float(tau), n_t.data_ptr(), b_t.data_ptr(), vp.data_ptr(), 0.0, 0.0, 0.0,
                                                  r_ri.data_ptr(), r_pi.data_ptr(), r_ni.data_ptr())
            torch.cuda.synchronize()
            assert Kr == K and torch.equal(r_ri[:K], ri[:K]) and torch.equal(r_pi[:K], pi[:K]) and torch.equal(r_ni[:K], ni[:K])


def test_cut_edge_cases():
    # single root leaf, viewpoint inside the box (size = FLT_MAX) and outside
    nodes = torch.tensor([[0, -1, 0, 1, 0, -1, 0]], dtype=torch.int32, device=DEV)
    boxes = torch.tensor([[[-1, -1, -1, 0.0], [1, 1, 1, 0.0]]], device=DEV)
    for mode in (0, 1):
        K, ri, pi, ni = _cut(nodes, boxes, 0.01, np.zeros(3, np.float32), mode)
        assert K == 1 and ri[0] == 0 and pi[0] == -1 and ni[0] == 0
        K, _, _, _ = _cut(nodes, boxes, 0.01, np.array([5, 0, 0], np.float32), mode)
        assert K == 0  # a leaf has w = 0 -> size 0 < tau and no parent
    # capacity overflow is reported, not written past the buffers
    nodes_np, boxes_np, views = cases.make_cut_case(L=500, seed=2)
    n_t, b_t = torch.from_numpy(nodes_np).to(DEV), torch.from_numpy(boxes_np).to(DEV)
    small = torch.zeros(10, dtype=torch.int32, device=DEV)
    with pytest.raises(lodgs.LodgsError, match="capacity"):
        gh.expand_to_size(n_t, b_t, 1e-4, torch.from_numpy(views[0]).to(DEV), torch.zeros(3), small, small.clone(), small.clone())
    # a non-monotone hierarchy is refused by the traversal check (and then served by the flat rule)
    bad = boxes_np.copy()
    bad[5, 0, 3] = bad[5, 1, 3] = 1e6
    assert gh.prepare_hierarchy(n_t, torch.from_numpy(bad).to(DEV)) is False
    # a 'monotone' verdict does not survive an in-place edit of the boxes (cache keyed on storage AND version): forced traversal mode
    # re-checks, falls back to the flat rule and returns the flat rule's cut
    b_edit = b_t.clone()
    assert gh.prepare_hierarchy(n_t, b_edit) is True
    b_edit[5, :, 3] = 1e6
    v = views[1]
    tau = np.float32(synth.tau_to_threshold(6, 0.8, 1920))
    K1, ri1, pi1, ni1 = _cut(n_t, b_edit, tau, v, 1)
    K0, ri0, pi0, ni0 = _cut(n_t, b_edit, tau, v, 0)
    want = orc.expand_to_size(nodes_np, b_edit.cpu().numpy(), tau, v)
    assert K1 == K0 == len(want[0]) and torch.equal(ri1[:K1], ri0[:K0]) and np.array_equal(ni1[:K1].cpu().numpy(), want[2])
    # overlapping child blocks (a node listed by two parents) are refused too
    n_bad = nodes_np.copy()
    inner = np.nonzero(n_bad[:, 6] > 0)[0]
    n_bad[inner[3], 5] = n_bad[inner[2], 5]           # node inner[3] now claims inner[2]'s children
    assert gh.prepare_hierarchy(torch.from_numpy(n_bad).to(DEV), b_t.clone()) is False


@pytest.mark.parametrize("n,end_bit", [(1, 45), (31, 8), (4096, 45), (4097, 45), (100003, 45), (1 << 20, 64), (777777, 33), (50000, 16)])
def test_onesweep_sort_matches_stable_sort(n, end_bit):
    g = torch.Generator(device=DEV).manual_seed(n)
    keys = torch.randint(0, 1 << 62, (n,), generator=g, device=DEV, dtype=torch.int64)
    if end_bit < 64:
        keys &= (1 << end_bit) - 1
    keys[::3] = keys[0].clone()  # many duplicates: stability matters
    vals = torch.arange(n, device=DEV, dtype=torch.int32)
    ko, vo, kt, vt = torch.empty_like(keys), torch.empty_like(vals), torch.empty_like(keys), torch.empty_like(vals)
    ws = torch.empty(lodgs.lib.lodgs_sort_workspace_bytes(n), dtype=torch.uint8, device=DEV)
    lodgs.check(lodgs.lib.lodgs_sort_pairs(keys.data_ptr(), vals.data_ptr(), ko.data_ptr(), vo.data_ptr(), kt.data_ptr(), vt.data_ptr(), n,
                                           end_bit, ws.data_ptr(), ws.numel(), lodgs.current_stream()))
    want_k, order = torch.sort(keys, stable=True)  # CUB radix sort underneath: the cross-check the north star allows
    assert torch.equal(ko, want_k)
    assert torch.equal(vo.long(), order)


def test_sort_structured_tile_depth_keys():
    n = 300000
    g = torch.Generator(device=DEV).manual_seed(1)
    tile = torch.randint(0, 8160, (n,), generator=g, device=DEV, dtype=torch.int64)
    depth = (torch.rand(n, generator=g, device=DEV) * 40 + 0.2).view(torch.int32).long()
    keys = (tile << 32) | depth
    vals = torch.arange(n, device=DEV, dtype=torch.int32)
    ko, vo, kt, vt = torch.empty_like(keys), torch.empty_like(vals), torch.empty_like(keys), torch.empty_like(vals)
    ws = torch.empty(lodgs.lib.lodgs_sort_workspace_bytes(n), dtype=torch.uint8, device=DEV)
    lodgs.check(lodgs.lib.lodgs_sort_pairs(keys.data_ptr(), vals.data_ptr(), ko.data_ptr(), vo.data_ptr(), kt.data_ptr(), vt.data_ptr(), n,
                                           45, ws.data_ptr(), ws.numel(), lodgs.current_stream()))
    assert (ko[1:] >= ko[:-1]).all() and torch.equal(ko, keys[vo.long()])
    assert torch.equal(torch.sort(vo).values, vals)  # a permutation


def _check_against(res, want, rtol_p=2e-6):
    from oracle.ouradam import LRS
    for k, (p, m, v) in res.items():
        wp, wm, wv = want[k]
        # rtol 1e-6 plus 2e-7 of the tensor's scale: 0.9 m + 0.1 g cancels for a few elements, and fused vs separate rounding shows there
        assert torch.allclose(m.cpu(), wm, rtol=1e-6, atol=2e-7 * float(wm.abs().max())), k
        assert torch.allclose(v.cpu(), wv, rtol=1e-6, atol=2e-7 * float(wv.abs().max())), k
        # a step moves a parameter by lr * O(1); its rounding (sqrt, divide, fused multiply-add) differs by ~1e-6 of the STEP, which for
        # parameters near zero is more than 2e-6 of the value
        assert torch.allclose(p.cpu(), wp, rtol=rtol_p, atol=1e-5 * LRS[k]), k


@pytest.mark.parametrize("rows,eps,seed", [(257, 1e-15, 11), (257, 1e-8, 12)])
def test_sparse_adam_matches_ouradam_golden(rows, eps, seed):
    """lodgs.optim.Adam.step(relevant) over the six parameter groups (widths 3/3/45/1/3/4), four steps (sorted list, unsorted
    list, single row, empty list = all rows) against the outputs of the reference's own scene/OurAdam.py
    (tests/golden/ouradam.npz, produced by the imported class on CPU; tests/golden/make_golden_adam.py)."""
    from lodgs.optim import Adam
    from oracle import ouradam
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ouradam.npz"))
    res = ouradam.run_scenario(Adam, rows, eps, seed, device=DEV)
    torch.cuda.synchronize()
    want = {k: tuple(torch.from_numpy(gold["r%d_s%d_%s_%s" % (rows, seed, k, x)]) for x in "pmv") for k in res}
    _check_against(res, want)


@pytest.mark.parametrize("rows,eps,seed", [(5000, 1e-15, 21), (20011, 1e-8, 22)])
def test_sparse_adam_matches_ouradam_restated_on_device(rows, eps, seed):
    """Larger row counts: the checker is oracle/ouradam.RestatedAdam on the same device (bit-pinned to scene/OurAdam.py by
    tests/test_ouradam_pin.py; where /root/reference exists the imported class itself is used)."""
    from lodgs.optim import Adam
    from oracle import ouradam
    cls, _ = ouradam.reference_or_restated()
    res = ouradam.run_scenario(Adam, rows, eps, seed, device=DEV)
    want = {k: tuple(x.cpu() for x in t) for k, t in ouradam.run_scenario(cls, rows, eps, seed, device=DEV).items()}
    _check_against(res, want)


def test_sparse_adam_leaves_unlisted_rows_untouched():
    rows, width = 5000, 45
    g = torch.Generator(device=DEV).manual_seed(width)
    p = torch.randn(rows, width, generator=g, device=DEV)
    grad = torch.randn(rows, width, generator=g, device=DEV) * 1e-3
    m, v = torch.randn(rows, width, generator=g, device=DEV) * 1e-3, torch.rand(rows, width, generator=g, device=DEV) * 1e-6
    rel = torch.randperm(rows, generator=g, device=DEV)[:1700].sort().values
    p0, m0, v0 = p.clone(), m.clone(), v.clone()
    lodgs.check(lodgs.lib.lodgs_adam_step(p.data_ptr(), grad.data_ptr(), m.data_ptr(), v.data_ptr(), rel.data_ptr(), rel.numel(), None, rows,
                                          width, 1.6e-4, 0.9, 0.999, 1e-15, 7, 1, lodgs.current_stream()))
    torch.cuda.synchronize()
    mask = torch.ones(rows, dtype=torch.bool, device=DEV)
    mask[rel] = False
    assert torch.equal(p[mask], p0[mask]) and torch.equal(m[mask], m0[mask]) and torch.equal(v[mask], v0[mask])
    assert (p[rel] != p0[rel]).any()


def test_adam_visibility_mask_no_bias_correction():
    rows, width = 3000, 3
    g = torch.Generator(device=DEV).manual_seed(5)
    p, grad = torch.randn(rows, width, generator=g, device=DEV), torch.randn(rows, width, generator=g, device=DEV)
    m, v = torch.zeros(rows, width, device=DEV), torch.zeros(rows, width, device=DEV)
    vis = torch.rand(rows, generator=g, device=DEV) < 0.3
    p0 = p.clone()
    lodgs.check(lodgs.lib.lodgs_adam_step(p.data_ptr(), grad.data_ptr(), m.data_ptr(), v.data_ptr(), None, 0, vis.data_ptr(), rows, width,
                                          1e-2, 0.9, 0.999, 1e-15, 1, 0, lodgs.current_stream()))
    # AR adam.cu:26-33 evaluates (1.0f - b1), (1.0f - b2) in float: 1-0.999f is 4.7e-5 away from 0.001
    omb1 = float(np.float32(1) - np.float32(0.9))
    omb2 = float(np.float32(1) - np.float32(0.999))
    mm, vv = omb1 * grad, omb2 * grad * grad
    want = p0 - 1e-2 * mm / (vv.sqrt() + 1e-15)
    assert torch.equal(p[~vis], p0[~vis])
    assert torch.allclose(p[vis], want[vis], rtol=1e-5, atol=1e-7)
