"""Diagnostic (not a test): per-gradient relative L2 of the alt-rasterizer surface against the reference AR kernels."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import conftest  # noqa: F401  (sets sys.path)
import torch
import test_gpu_alt_rasterizer as T
from oracle import refshim_ar
from alt_gaussian_rasterization import _C

for name in ("wide",):
    for aa in (True,):
        for zero_depth in (False, True):
            sc, cam, bg, dL, dLi = T._scene(name)
            if zero_depth:
                dLi = torch.zeros_like(dLi)
            rf = refshim_ar.forward(sc, cam, bg, aa)
            rg = refshim_ar.backward(sc, cam, bg, rf, dL, dLi, aa)
            e = torch.empty(0, device="cuda")
            out = _C.rasterize_gaussians(bg, sc["means3D"], e, sc["opacities"], sc["scales"], sc["rotations"], 1.0, e, cam["viewmatrix"],
                                         cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], cam["H"], cam["W"], sc["dc"], sc["shs_rest"],
                                         sc["sh_degree"], cam["campos"], False, aa, False)
            rendered, nb, color, invdepth, radii, geom, binning, img, sample = out
            grads = _C.rasterize_gaussians_backward(bg, sc["means3D"], radii, e, sc["opacities"], sc["scales"], sc["rotations"], 1.0, e,
                                                    cam["viewmatrix"], cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], dL, sc["dc"],
                                                    sc["shs_rest"], dLi, sc["sh_degree"], cam["campos"], geom, rendered, binning, img, nb,
                                                    sample, aa, False)
            names = ("dL_dmean2D", "dL_dcolor", "dL_dopacity", "dL_dmean3D", "dL_dcov3D", "dL_ddc", "dL_dsh", "dL_dscale", "dL_drot")
            print("case", name, "aa", aa, "zero_depth", zero_depth, "R", rendered, rf["R"], "img err", float((color - rf["color"]).abs().max()))
            for n, g in zip(names, grads):
                print("   %-12s rel_l2 %.3e   |ref| %.3e |mine| %.3e" % (n, T.rel_l2(g, rg[n]), float(rg[n].norm()), float(g.norm())))
            d = (grads[0] - rg["dL_dmean2D"])
            i = int(d.abs().sum(1).argmax())
            op = sc["opacities"].reshape(-1)
            bad = (d.abs().sum(1) > 1e-3 * rg["dL_dmean2D"].abs().sum(1).max())
            print("   bad rows", int(bad.sum()), "opacity of bad rows (min/mean/max)", float(op[bad].min()), float(op[bad].mean()), float(op[bad].max()),
                  "radii", radii[bad][:10].tolist())
            print("   worst row", i, grads[0][i].tolist(), rg["dL_dmean2D"][i].tolist(), "radius", int(radii[i]))

# ---- where do the rows the reference leaves at zero sit in their tile lists?
import ctypes as C
import lodgs
sc, cam, bg, dL, dLi = T._scene("wide")
rf = refshim_ar.forward(sc, cam, bg, True)
rg1 = refshim_ar.backward(sc, cam, bg, rf, dL, dLi, True)
rg2 = refshim_ar.backward(sc, cam, bg, rf, dL, dLi, True)
print("reference deterministic:", all(torch.equal(rg1[k], rg2[k]) for k in rg1))
e = torch.empty(0, device="cuda")
out = _C.rasterize_gaussians(bg, sc["means3D"], e, sc["opacities"], sc["scales"], sc["rotations"], 1.0, e, cam["viewmatrix"],
                             cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], cam["H"], cam["W"], sc["dc"], sc["shs_rest"],
                             sc["sh_degree"], cam["campos"], False, True, False)
rendered, nb, color, invdepth, radii, geom, binning, img, sample = out
grads = _C.rasterize_gaussians_backward(bg, sc["means3D"], radii, e, sc["opacities"], sc["scales"], sc["rotations"], 1.0, e,
                                        cam["viewmatrix"], cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], dL, sc["dc"],
                                        sc["shs_rest"], dLi, sc["sh_degree"], cam["campos"], geom, rendered, binning, img, nb, sample, True, False)
W, H = cam["W"], cam["H"]
o = (C.c_size_t * 3)(); lodgs.lib.lodgs_image_layout(W, H, o)
ntiles = ((W + 15) // 16) * ((H + 15) // 16)
ranges = img[o[2]:o[2] + ntiles * 8].view(torch.int32).reshape(ntiles, 2).cpu()
ncon = img[o[1]:o[1] + W * H * 4].view(torch.int32).reshape(H, W).cpu()
b = (C.c_size_t * 5)(); lodgs.lib.lodgs_binning_layout(rendered, b)
pl = binning[b[3]:b[3] + rendered * 4].view(torch.int32).cpu()
zero_ref = (rg1["dL_dmean2D"].abs().sum(1) == 0) & (grads[0].abs().sum(1) > 0)
rows = zero_ref.nonzero().reshape(-1).cpu().tolist()
print("rows zero in the reference but not here:", len(rows))
gx = (W + 15) // 16
for r in rows[:12]:
    info = []
    for t in range(ntiles):
        lo, hi = int(ranges[t, 0]), int(ranges[t, 1])
        seg = pl[lo:hi]
        pos = (seg == r).nonzero().reshape(-1)
        if pos.numel():
            ty, tx = divmod(t, gx)
            mc = int(ncon[ty * 16:(ty + 1) * 16, tx * 16:(tx + 1) * 16].max())
            info.append((t, int(pos[0]), hi - lo, mc))
    print("  gaussian", r, "radius", int(radii[r]), "(tile, position in list, list length, max n_contrib of tile):", info)

# ---- is the reference's zero right? finite difference THROUGH THE REFERENCE FORWARD on the DC colour (the loss is linear in it)
for r in rows[:4]:
    base = float((rf["color"].double() * dL.double()).sum() + (rf["invdepth"].double() * dLi.double()).sum())
    sc2 = dict(sc)
    dc2 = sc["dc"].clone()
    dc2[r, 0, 0] += 1.0
    sc2["dc"] = dc2
    f2 = refshim_ar.forward(sc2, cam, bg, True)
    pert = float((f2["color"].double() * dL.double()).sum() + (f2["invdepth"].double() * dLi.double()).sum())
    print("  gaussian", r, "dL/d dc[0] by finite difference through the reference forward: %.6e   reference backward: %.6e   here: %.6e"
          % (pert - base, float(rg1["dL_ddc"][r, 0, 0]), float(grads[5][r, 0, 0])))
