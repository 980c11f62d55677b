"""Diagnostic (not a test): WHERE does the reference's per-Gaussian backward (AR/cuda_rasterizer/backward.cu:451-656) lose
contributions? For every Gaussian whose colour gradient differs between the reference's backward and this library's, the
script replays the reference's OWN sorted lists (its point_list / ranges / n_contrib / max_contrib / sampled_T, read from its scratch
buffers through refar_internals) in torch, attributes dL_dcolor to the individual (tile, list position) instances and reports which
instances the reference's result is missing and what state the reference kernel saw for them.

    python tests/gpu_diag_ar.py [case ...]        (GPU box)
"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import conftest  # noqa: F401,E402  (sets sys.path)
import torch  # noqa: E402

import test_gpu_alt_rasterizer as T  # noqa: E402
from alt_gaussian_rasterization import _C  # noqa: E402
from oracle import refshim_ar  # noqa: E402


def view(ptr, n, dtype):
    """Copy of n elements of device memory at ptr (the reference's scratch buffers) as a torch tensor."""
    t = torch.empty(n, dtype=dtype, device="cuda")
    L = refshim_ar.lib()
    L.refar_copy.restype = C.c_int
    rc = L.refar_copy(C.c_void_p(t.data_ptr()), C.c_void_p(ptr), C.c_size_t(n * t.element_size()))
    assert rc == 0, rc
    return t


def internals(rf, W, H):
    out = (C.c_void_p * 16)()
    L = refshim_ar.lib()
    L.refar_internals.restype = None
    L.refar_internals(C.c_int(rf["P"]), C.c_int(W), C.c_int(H), C.c_int(rf["R"]), C.c_int(rf["B"]), out)
    ntiles = ((W + 15) // 16) * ((H + 15) // 16)
    P, R, B = rf["P"], rf["R"], rf["B"]
    return {"ranges": view(out[0], ntiles * 2, torch.int32).view(ntiles, 2), "point_list": view(out[1], R, torch.int32),
            "n_contrib": view(out[2], W * H, torch.int32), "max_contrib": view(out[3], ntiles, torch.int32),
            "bucket_offsets": view(out[4], ntiles, torch.int32), "bucket_to_tile": view(out[5], B, torch.int32),
            "sampled_T": view(out[6], B * 256, torch.float32).view(B, 256), "means2D": view(out[7], P * 2, torch.float32).view(P, 2),
            "conic_opacity": view(out[8], P * 4, torch.float32).view(P, 4), "rgb": view(out[9], P * 3, torch.float32).view(P, 3),
            "final_T": view(out[10], W * H, torch.float32)}


def replay_tile(st, t, W, H, dL):
    """-> (ids [n], weight [n,256] = alpha*T of every (position, pixel) the forward blended, pixel validity [256], pix ids [256])."""
    gx = (W + 15) // 16
    ty, tx = divmod(t, gx)
    lo, hi = (int(x) for x in st["ranges"][t])
    ids = st["point_list"][lo:hi].long()
    k = torch.arange(256, device="cuda")
    px, py = tx * 16 + k % 16, ty * 16 + k // 16
    valid = (px < W) & (py < H)
    pid = (py * W + px).clamp(max=W * H - 1)
    xy, co = st["means2D"][ids], st["conic_opacity"][ids]
    dx = xy[:, 0:1] - px.float()[None]
    dy = xy[:, 1:2] - py.float()[None]
    power = -0.5 * (co[:, 0:1] * dx * dx + co[:, 2:3] * dy * dy) - co[:, 1:2] * dx * dy
    alpha = torch.minimum(torch.tensor(0.99, device="cuda"), co[:, 3:4] * torch.exp(power))
    ok = (power <= 0) & (alpha >= 1.0 / 255.0)
    pos = torch.arange(hi - lo, device="cuda")[:, None]
    blended = ok & (pos < st["n_contrib"][pid][None]) & valid[None]
    a = torch.where(ok, alpha, torch.zeros_like(alpha))
    Tb = torch.cumprod(torch.cat([torch.ones(1, 256, device="cuda"), 1 - a[:-1]], 0), 0)   # T before each position
    w = torch.where(blended, a * Tb, torch.zeros_like(a))
    return ids, w, Tb, valid, pid


def main(names):
    for name in names:
        for aa in (True,):
            sc, cam, bg, dL, dLi = T._scene(name)
            W, H = cam["W"], cam["H"]
            rf = refshim_ar.forward(sc, cam, bg, aa)
            rg = refshim_ar.backward(sc, cam, bg, rf, dL, dLi, aa)
            st = internals(rf, W, H)
            e = torch.empty(0, device="cuda")
            out = _C.rasterize_gaussians(bg, sc["means3D"], e, sc["opacities"], sc["scales"], sc["rotations"], 1.0, e, cam["viewmatrix"],
                                         cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], H, W, sc["dc"], sc["shs_rest"], sc["sh_degree"],
                                         cam["campos"], False, aa, False)
            rendered, nb, color, invdepth, radii, geom, binning, img, sample = out
            grads = _C.rasterize_gaussians_backward(bg, sc["means3D"], radii, e, sc["opacities"], sc["scales"], sc["rotations"], 1.0, e,
                                                    cam["viewmatrix"], cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], dL, sc["dc"],
                                                    sc["shs_rest"], dLi, sc["sh_degree"], cam["campos"], geom, rendered, binning, img, nb,
                                                    sample, aa, False)
            mine_c, ref_c = grads[1], rg["dL_dcolor"]
            cdiff = (mine_c - ref_c).abs().sum(1)
            dropped = cdiff > 1e-3 * ref_c.abs().sum(1) + 1e-12
            rows = dropped.nonzero().reshape(-1).tolist()
            print("case %s aa %s: visible %d, Gaussians whose dL_dcolor differs %d, R %d, B %d" % (name, aa, int((radii > 0).sum()), len(rows),
                                                                                                    rf["R"], rf["B"]))
            ntiles = st["ranges"].size(0)
            lens = (st["ranges"][:, 1] - st["ranges"][:, 0])
            print("   tile list lengths: max %d, tiles with > 256 entries %d of %d; max_contrib max %d" %
                  (int(lens.max()), int((lens > 256).sum()), ntiles, int(st["max_contrib"].max())))
            dLf = dL.reshape(3, -1)
            cache = {}
            summary = {"missing_instances": 0, "pos_lt_256": 0, "pos_ge_256": 0}
            for r in rows[:40]:
                # instances of r in the reference's lists
                hits = (st["point_list"] == r).nonzero().reshape(-1)
                inst = []
                for h in hits.tolist():
                    t = int(((st["ranges"][:, 0] <= h) & (h < st["ranges"][:, 1])).nonzero().reshape(-1)[0])
                    if t not in cache:
                        cache[t] = replay_tile(st, t, W, H, dL)
                    ids, w, Tb, valid, pid = cache[t]
                    pos = h - int(st["ranges"][t, 0])
                    contrib = (w[pos][None] * dLf[:, pid]).sum(1)          # [3]
                    inst.append((t, pos, contrib))
                total = sum(c for _, _, c in inst) if inst else torch.zeros(3, device="cuda")
                miss = []
                for t, pos, c in inst:
                    if float(c.abs().sum()) > 0 and float((total - c - ref_c[r]).abs().sum()) < 0.02 * float(c.abs().sum()):
                        miss.append((t, pos))
                print("   gaussian %d: mine %s | replay of the reference's lists %s | reference backward %s" %
                      (r, ["%.3e" % x for x in mine_c[r].tolist()], ["%.3e" % x for x in total.tolist()], ["%.3e" % x for x in ref_c[r].tolist()]))
                for t, pos, c in inst:
                    lo, hi = (int(x) for x in st["ranges"][t])
                    bbm = 0 if t == 0 else int(st["bucket_offsets"][t - 1])
                    b = bbm + pos // 32
                    ids, w, Tb, valid, pid = cache[t]
                    # what the kernel read for the bucket: sampled_T of the pixels that blend this instance vs the replayed T at the bucket start
                    pix = (w[pos] > 0).nonzero().reshape(-1)
                    sT = st["sampled_T"][b][pix] if b < st["sampled_T"].size(0) else None
                    rT = Tb[(pos // 32) * 32][pix]
                    mism = int(((sT - rT).abs() > 1e-4 * rT.abs() + 1e-9).sum()) if sT is not None else -1
                    print("      tile %d pos %d of %d (bucket %d, lane %d) max_contrib %d bucket_to_tile %s  contributes %s  blending pixels %d, "
                          "sampled_T != replayed T at bucket start for %d of them%s" %
                          (t, pos, hi - lo, pos // 32, pos % 32, int(st["max_contrib"][t]), int(st["bucket_to_tile"][b]) if b < st["bucket_to_tile"].numel() else None,
                           ["%.3e" % x for x in c.tolist()], pix.numel(), mism, "   <-- MISSING in the reference backward" if (t, pos) in miss else ""))
                    if (t, pos) in miss:
                        summary["missing_instances"] += 1
                        summary["pos_lt_256" if pos < 256 else "pos_ge_256"] += 1
            print("   summary", summary)


if __name__ == "__main__":
    main(sys.argv[1:] or ["wide"])
