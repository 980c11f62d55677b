"""GPU: the cut and the interpolation weights on the dynamic hierarchy format (expand_to_size_dynamic /
get_interpolation_weights_dynamic) against the CPU oracle and the reference CUDA kernels run side by side — bit-exact
indices, kids and weights."""
import numpy as np
import pytest
import torch

import cases
from gaussian_hierarchy import _C as gh
from lodgs import synth
from oracle import oracle as orc
from oracle import refshim

pytestmark = pytest.mark.gpu
DEV = "cuda"
FILL = -7


def _dyn_case(L, seed, skybox):
    rng = np.random.default_rng(seed)
    pts = (rng.uniform(-1, 1, (L, 3)) * np.array([40.0, 40.0, 4.0])).astype(np.float32)
    rad = np.exp(rng.normal(np.log(0.05), 0.5, L)).astype(np.float32)
    nodes, boxes, gpos, grad = synth.build_hierarchy(pts, rad)
    return synth.to_dynamic_hierarchy(nodes, boxes, gpos, grad, seed=seed, skybox=skybox)


def _mine(n_t, p_t, s_t, tau, v, z):
    N = n_t.size(0)
    ri, pi, ni = (torch.full((N,), FILL, dtype=torch.int32, device=DEV) for _ in range(3))
    K = gh.expand_to_size_dynamic(n_t, p_t, s_t, float(tau), torch.from_numpy(v).to(DEV), torch.from_numpy(z), ri, pi, ni)
    return K, ri, pi, ni


@pytest.mark.parametrize("L,skybox", [(1, 0), (37, 3), (3000, 5), (50000, 100)])
def test_dynamic_cut_bit_exact_vs_oracle_and_reference(L, skybox):
    dyn, pos, sca = _dyn_case(L, 11 + L, skybox)
    n_t, p_t, s_t = torch.from_numpy(dyn).to(DEV), torch.from_numpy(pos).to(DEV), torch.from_numpy(sca).to(DEV)
    N = dyn.shape[0]
    rng = np.random.default_rng(5)
    nonempty = 0
    for trial in range(8):
        v = (rng.uniform(-1, 1, 3) * np.array([60, 60, 30])).astype(np.float32)
        z = rng.normal(size=3).astype(np.float32)
        z /= np.linalg.norm(z)
        if trial % 2 == 0:  # look at the scene centre: most nodes are inside the cone (cos(viewpoint - position, zdir) < -0.5)
            z = (-v / np.linalg.norm(v)).astype(np.float32)
        tau = np.float32(rng.choice([0.0, 0.002, 0.01, 0.05, 0.5]))
        K, ri, pi, ni = _mine(n_t, p_t, s_t, tau, v, z)
        o_ri, o_pi, o_ni = orc.expand_to_size_dynamic(dyn, pos, sca, tau, v, z, parent_fill=FILL)
        assert K == len(o_ri), (trial, K, len(o_ri))
        assert np.array_equal(ri[:K].cpu().numpy(), o_ri) and np.array_equal(ni[:K].cpu().numpy(), o_ni)
        assert np.array_equal(pi[:K].cpu().numpy(), o_pi)
        assert (ri[K:] == FILL).all() and (pi[K:] == FILL).all()
        nonempty += K > 0
        if refshim.available():
            r_ri, r_pi, r_ni = (torch.full((N,), FILL, dtype=torch.int32, device=DEV) for _ in range(3))
            vp = torch.from_numpy(v).to(DEV)
            torch.cuda.synchronize()
            Kr = refshim.lib().ref_expand_to_size_dynamic(N, float(tau), n_t.data_ptr(), p_t.data_ptr(), s_t.data_ptr(), vp.data_ptr(),
                                                          float(z[0]), float(z[1]), float(z[2]), r_ri.data_ptr(), r_pi.data_ptr(),
                                                          r_ni.data_ptr())
            torch.cuda.synchronize()
            assert Kr == K
            assert torch.equal(r_ri[:K], ri[:K]) and torch.equal(r_pi[:K], pi[:K]) and torch.equal(r_ni[:K], ni[:K])
        if K == 0:
            continue
        ts = torch.zeros(K, device=DEV)
        kids = torch.zeros(K, dtype=torch.int32, device=DEV)
        gh.get_interpolation_weights_dynamic(ni[:K], float(tau), n_t, p_t, s_t, torch.from_numpy(v), torch.from_numpy(z), ts, kids)
        o_ts, o_kids = orc.get_interpolation_weights_dynamic(o_ni, tau, dyn, pos, sca, v)
        assert np.array_equal(kids.cpu().numpy(), o_kids)
        assert np.array_equal(ts.cpu().numpy().view(np.uint32), o_ts.view(np.uint32))
        if refshim.available():
            r_ts = torch.zeros(K, device=DEV)
            r_kids = torch.zeros(K, dtype=torch.int32, device=DEV)
            idx = ni[:K].contiguous()
            torch.cuda.synchronize()
            refshim.lib().ref_get_ts_indexed_dynamic(K, idx.data_ptr(), float(tau), n_t.data_ptr(), p_t.data_ptr(), s_t.data_ptr(),
                                                     float(v[0]), float(v[1]), float(v[2]), float(z[0]), float(z[1]), float(z[2]),
                                                     r_ts.data_ptr(), r_kids.data_ptr(), None)
            torch.cuda.synchronize()
            assert torch.equal(r_kids, kids) and torch.equal(r_ts.view(torch.int32), ts.view(torch.int32))
    if L >= 3000:
        assert nonempty >= 1


def test_dynamic_cut_capacity_overflow_is_reported():
    import lodgs
    dyn, pos, sca = _dyn_case(3000, 3, 0)
    n_t, p_t, s_t = torch.from_numpy(dyn).to(DEV), torch.from_numpy(pos).to(DEV), torch.from_numpy(sca).to(DEV)
    v = np.array([0.0, 0.0, -60.0], np.float32)
    z = np.array([0.0, 0.0, 1.0], np.float32)
    K, ri, pi, ni = _mine(n_t, p_t, s_t, 0.0, v, z)
    assert K > 8
    small = torch.zeros(4, dtype=torch.int32, device=DEV)
    with pytest.raises(lodgs.LodgsError):
        gh.expand_to_size_dynamic(n_t, p_t, s_t, 0.0, torch.from_numpy(v).to(DEV), torch.from_numpy(z), small, small.clone(), small.clone())
