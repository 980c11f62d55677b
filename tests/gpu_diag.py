"""Diagnostic run on a GPU box: product path vs the reference CUDA build on the same synthetic inputs, with
mismatch statistics for every intermediate. `python tests/gpu_diag.py [P] [W] [H]`."""
import sys
import time

import numpy as np
import torch

import conftest  # noqa: F401  (sets sys.path)
import minewrap
from oracle import refshim
from lodgs import synth


def cmp_int(name, a, b):
    a, b = a.reshape(-1), b.reshape(-1)
    n = (a != b).sum().item()
    print("  %-16s %s  mismatches %d / %d" % (name, "OK  " if n == 0 else "FAIL", n, a.numel()))
    if n:
        i = (a != b).nonzero()[:5].reshape(-1).tolist()
        print("     first idx", i, "mine", a[i].tolist(), "ref", b[i].tolist())
    return n


def cmp_f(name, a, b, mask=None):
    if mask is not None:
        a, b = a[mask], b[mask]
    a, b = a.reshape(-1).double(), b.reshape(-1).double()
    bits = (a != b).sum().item()
    mx = (a - b).abs().max().item() if a.numel() else 0.0
    rel = ((a - b).norm() / (b.norm() + 1e-30)).item()
    print("  %-16s max-abs %.3e  rel-L2 %.3e  non-identical %d / %d" % (name, mx, rel, bits, a.numel()))
    return mx, rel


def run(P, W, H, lod=False, seed=0):
    print("== P=%d %dx%d lod=%s" % (P, W, H, lod))
    cam_np = synth.make_camera(W=W, H=H, fx=1200.0 * W / 1920, fy=1200.0 * W / 1920)
    sc = minewrap.to_cuda(synth.make_gaussians(P, seed=seed, tanfovx=cam_np["tanfovx"], tanfovy=cam_np["tanfovy"]))
    cam = minewrap.to_cuda(cam_np)
    bg = torch.tensor([0.1, 0.2, 0.3], device="cuda")
    kw = {}
    if lod:
        g = torch.Generator().manual_seed(seed)
        K = P // 2
        idx = torch.randperm(P, generator=g)[:K].sort().values.int()
        par = torch.randint(0, P, (K,), generator=g).int()
        par[torch.rand(K, generator=g) < 0.2] = -1
        kw = {"indices": idx.cuda(), "parent_indices": par.cuda(), "ts": torch.rand(K, generator=g).cuda(),
              "kids": torch.randint(1, 5, (K,), generator=g).int().cuda()}
    t0 = time.time()
    m = minewrap.forward(sc, cam, bg, **kw)
    print("  mine fwd %.1f ms, R=%d" % ((time.time() - t0) * 1e3, m["R"]))
    t0 = time.time()
    r = refshim.forward(sc, cam, bg, **kw)
    print("  ref  fwd %.1f ms, R=%d" % ((time.time() - t0) * 1e3, r["R"]))
    vis = r["radii"] > 0
    bad = 0
    bad += cmp_int("radii", m["radii"], r["radii"])
    bad += cmp_int("rects", m["rects"][vis], r["rects"][vis])
    bad += cmp_int("tiles_touched", m["tiles_touched"], r["tiles_touched"])
    cmp_f("depths", m["depths"], r["depths"], vis)
    cmp_f("means2D", m["means2D"], r["means2D"], vis)
    cmp_f("conic_opacity", m["conic_opacity"], r["conic_opacity"], vis)
    cmp_f("rgb", m["rgb"], r["rgb"], vis)
    cmp_f("cov3D", m["cov3D"], r["cov3D"], vis)
    if m["R"] == r["R"] and r["R"] > 0:
        bad += cmp_int("keys_sorted", m["keys"], r["keys"])
        bad += cmp_int("point_list", m["point_list"], r["point_list"])
        bad += cmp_int("ranges", m["ranges"], r["ranges"])
        bad += cmp_int("n_contrib", m["n_contrib"], r["n_contrib"])
        cmp_f("final_T", m["final_T"], r["final_T"])
        cmp_f("color", m["color"], r["color"])
        cmp_f("invdepth", m["invdepth"], r["invdepth"])
        bad += cmp_int("seen", m["seen"], r["seen"])
        torch.manual_seed(1)
        target = torch.rand(3, H, W, device="cuda")
        dL = (r["color"] - target) / (3 * H * W)
        dLi = torch.rand(1, H, W, device="cuda") / (H * W)
        gm = minewrap.backward(sc, cam, bg, m, dL, dLi, **kw)
        gr = refshim.backward(sc, cam, bg, r, dL, dLi, **kw)
        for k in ("dL_dmean2D", "dL_dcolor", "dL_dopacity", "dL_dmean3D", "dL_dcov3D", "dL_dsh", "dL_dscale", "dL_drot"):
            cmp_f(k, gm[k], gr[k])
    print("  integer mismatches total:", bad)
    return bad


if __name__ == "__main__":
    a = [int(x) for x in sys.argv[1:]]
    if a:
        run(a[0], a[1], a[2])
    else:
        run(2000, 256, 192)
        run(50000, 640, 360)
        run(20000, 320, 200, lod=True)
        run(1000000, 1920, 1080)
        run(1000000, 1920, 1080, lod=True)
