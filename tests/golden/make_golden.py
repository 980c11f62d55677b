"""Generates tests/golden/*.npz by running the UNMODIFIED reference CUDA kernels (oracle/_ref/libref_lodgs.so) on a GPU box:
    gpurun -- 'python tests/golden/make_golden.py'   (writes into gpurun_out/golden/, copied to tests/golden/ and committed)
Inputs are regenerated from seeds by tests/cases.py, so only reference OUTPUTS are stored."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import conftest  # noqa: E402,F401
import cases  # noqa: E402
from oracle import refshim  # noqa: E402
from lodgs import synth  # noqa: E402

OUT = os.path.join(conftest.ROOT, "gpurun_out", "golden")
os.makedirs(OUT, exist_ok=True)


def cu(d):
    return {k: (torch.from_numpy(v).cuda() if isinstance(v, np.ndarray) else v) for k, v in d.items()}


for name in cases.CASES:
    sc, cam, bg, kw, dL, dLi = cases.make_case(name)
    scg, camg, kwg = cu(sc), cu(cam), cu(kw)
    fwd = refshim.forward(scg, camg, torch.from_numpy(bg).cuda(), **kwg)
    g = refshim.backward(scg, camg, torch.from_numpy(bg).cuda(), fwd, torch.from_numpy(dL).cuda(), torch.from_numpy(dLi).cuda(), **kwg)
    out = {k: v.cpu().numpy() for k, v in fwd.items() if torch.is_tensor(v)}
    out.update({k: v.cpu().numpy() for k, v in g.items() if v is not None})
    out["R"] = np.int64(fwd["R"])
    np.savez_compressed(os.path.join(OUT, "raster_%s.npz" % name), **out)
    print(name, "R", fwd["R"], "visible", int((fwd["radii"] > 0).sum()))

nodes, boxes, views = cases.make_cut_case()
res = {}
L = refshim.lib()
n_t, b_t = torch.from_numpy(nodes).cuda(), torch.from_numpy(boxes).cuda()
N = nodes.shape[0]
for vi, v in enumerate(views):
    for tau_px in (0, 3, 6, 15, 60):
        tau = synth.tau_to_threshold(tau_px, 0.8, 1920)
        ri, pi, ni = (torch.zeros(N, dtype=torch.int32, device="cuda") for _ in range(3))
        vp = torch.from_numpy(v).cuda()
        K = L.ref_expand_to_size(N, float(tau), n_t.data_ptr(), b_t.data_ptr(), vp.data_ptr(), 0.0, 0.0, 0.0, ri.data_ptr(), pi.data_ptr(),
                                 ni.data_ptr())
        ts, kids = torch.zeros(max(K, 1), device="cuda"), torch.zeros(max(K, 1), dtype=torch.int32, device="cuda")
        if K:
            L.ref_get_ts_indexed(K, ni.data_ptr(), float(tau), n_t.data_ptr(), b_t.data_ptr(), float(v[0]), float(v[1]), float(v[2]), 0.0,
                                 0.0, 0.0, ts.data_ptr(), kids.data_ptr(), None)
        torch.cuda.synchronize()
        key = "v%d_t%d" % (vi, tau_px)
        res[key + "_ri"], res[key + "_pi"], res[key + "_ni"] = ri[:K].cpu().numpy(), pi[:K].cpu().numpy(), ni[:K].cpu().numpy()
        res[key + "_ts"], res[key + "_kids"] = ts[:K].cpu().numpy(), kids[:K].cpu().numpy()
        print(key, "K", K)
np.savez_compressed(os.path.join(OUT, "cut.npz"), **res)

# ---- dynamic-format cut (expandToSizeDynamic / getTsIndexedDynamic)
dyn, pos, sca, dviews = cases.make_dynamic_cut_case()
d_t, p_t, s_t = torch.from_numpy(dyn).cuda(), torch.from_numpy(pos).cuda(), torch.from_numpy(sca).cuda()
N = dyn.shape[0]
res = {}
for vi, (v, z) in enumerate(dviews):
    for ti, tau in enumerate(cases.DYN_TAUS):
        ri, pi, ni = (torch.full((N,), -7, dtype=torch.int32, device="cuda") for _ in range(3))
        vp = torch.from_numpy(v).cuda()
        torch.cuda.synchronize()
        K = L.ref_expand_to_size_dynamic(N, float(tau), d_t.data_ptr(), p_t.data_ptr(), s_t.data_ptr(), vp.data_ptr(), float(z[0]), float(z[1]),
                                         float(z[2]), ri.data_ptr(), pi.data_ptr(), ni.data_ptr())
        ts, kids = torch.zeros(max(K, 1), device="cuda"), torch.zeros(max(K, 1), dtype=torch.int32, device="cuda")
        if K:
            L.ref_get_ts_indexed_dynamic(K, ni.data_ptr(), float(tau), d_t.data_ptr(), p_t.data_ptr(), s_t.data_ptr(), float(v[0]), float(v[1]),
                                         float(v[2]), float(z[0]), float(z[1]), float(z[2]), ts.data_ptr(), kids.data_ptr(), None)
        torch.cuda.synchronize()
        key = "v%d_t%d" % (vi, ti)
        res[key + "_ri"], res[key + "_pi"], res[key + "_ni"] = ri[:K].cpu().numpy(), pi[:K].cpu().numpy(), ni[:K].cpu().numpy()
        res[key + "_ts"], res[key + "_kids"] = ts[:K].cpu().numpy(), kids[:K].cpu().numpy()
        print("dyn", key, "K", K)
np.savez_compressed(os.path.join(OUT, "cut_dynamic.npz"), **res)

# ---- SPT cut (Switching::getSPTCut)
res = {}
for name in cases.SPT_CASES:
    arrs = [torch.from_numpy(a).cuda() for a in cases.make_spt_case(name)]
    cut, counts = refshim.get_spt_cut(*arrs)
    res[name + "_cut"], res[name + "_counts"] = cut.cpu().numpy(), counts.cpu().numpy()
    print("spt", name, "K", cut.numel())
np.savez_compressed(os.path.join(OUT, "cut_spt.npz"), **res)

# ---- Morton codes (MORTON::getMortonCode)
res = {}
for name, deg in (("plain", False), ("flat_z", True)):
    xyz, mn, mx = cases.make_morton_case(degenerate=deg)
    t = [torch.from_numpy(a).cuda() for a in (xyz, mn, mx)]
    codes = torch.zeros(xyz.shape[0], dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    L.ref_get_morton_code(t[0].data_ptr(), t[1].data_ptr(), t[2].data_ptr(), codes.data_ptr(), xyz.shape[0])
    torch.cuda.synchronize()
    res[name] = codes.cpu().numpy()
    print("morton", name, int(codes.max()))
np.savez_compressed(os.path.join(OUT, "morton.npz"), **res)
