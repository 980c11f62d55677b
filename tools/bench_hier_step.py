#!/usr/bin/env python3
"""BASELINE.json configs[2]: single-chunk training step WITH LOD interpolation on a synthetic hierarchy —
cut (expand_to_size) -> interpolation weights -> forward with in-kernel LOD blending (render indices, parent indices, t, sibling
counts through the plugin call) -> backward -> sparse Adam on the visible rows — at 1920x1080, tau in {0,3,6,15} px, next to the
reference CUDA (oracle/_ref: Switching::expandToSize / getTsIndexed and the hierarchy rasterizer, unmodified) on the same inputs.
Images are compared bit for bit, K and num_rendered exactly. Times are wall clock per phase around synchronised loops (both
arms end the cut and the forward with a count on the host).

  python tools/bench_hier_step.py [--k 21] [--iters 10]      -> one JSON line
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "hierarchical-lod-gaussians_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

W, H, DEG, M = 1920, 1080, 3, 16


def make_scene(k, dev):
    """Hierarchy of 2^k leaves in a 400x400x20 m slab and one Gaussian per node: mean = box centre, scales = box sides / 6
    (leaf boxes are point +- 3 radii, so a leaf gets its radius back), random rotations, opacities and SH."""
    import torch
    from lodgs import synth
    nodes, boxes = synth.build_hierarchy_torch(k, dev, extent=(200.0, 200.0, 10.0), radius_mean=0.1)
    N = nodes.size(0)
    g = torch.Generator(device=dev).manual_seed(7)
    lo, hi = boxes[:, 0, :3], boxes[:, 1, :3]
    sc = {"means3D": ((lo + hi) * 0.5).contiguous(), "scales": ((hi - lo) / 6.0).clamp_min(1e-3).contiguous(),
          "rotations": torch.nn.functional.normalize(torch.randn(N, 4, device=dev, generator=g), dim=1).contiguous(),
          "opacities": torch.sigmoid(torch.randn(N, 1, device=dev, generator=g) * 2.0).contiguous()}
    shs = torch.randn(N, M, 3, device=dev, generator=g) * 0.1
    shs[:, 0] = torch.randn(N, 3, device=dev, generator=g) * 0.5 / 0.2821 + 0.5 / 0.2821
    sc["shs"] = shs.contiguous()
    return nodes, boxes, sc


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--k", type=int, default=21)
    ap.add_argument("--iters", type=int, default=10)
    args = ap.parse_args()
    import numpy as np
    import torch
    import lodgs
    from diff_gaussian_rasterization import _C as dgr
    from gaussian_hierarchy import _C as gh
    from lodgs import synth
    from oracle import refshim
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    nodes, boxes, sc = make_scene(args.k, dev)
    N = nodes.size(0)
    cam_np = synth.look_at_camera(eye=(-150.0, 20.0, 18.0), target=(0.0, 0.0, 0.0), up=(0.0, 0.0, 1.0), W=W, H=H, fx=1200.0, fy=1200.0)
    cam = {k: (torch.from_numpy(np.ascontiguousarray(v, dtype=np.float32)).to(dev) if isinstance(v, np.ndarray) else v) for k, v in cam_np.items()}
    campos_cpu = cam["campos"].cpu()
    bg = torch.zeros(3, device=dev)
    target = torch.rand(3, H, W, device=dev, generator=torch.Generator(device=dev).manual_seed(1))
    ei, ef, z3 = torch.empty(0, dtype=torch.int32, device=dev), torch.empty(0, device=dev), torch.zeros(3)
    ri, pi, ni = (torch.empty(N, dtype=torch.int32, device=dev) for _ in range(3))
    ts, kids = torch.empty(N, device=dev), torch.empty(N, dtype=torch.int32, device=dev)
    gh.prepare_hierarchy(nodes, boxes)
    ref = refshim.lib() if refshim.available() else None
    if ref is not None:
        rri, rpi, rni = (torch.empty(N, dtype=torch.int32, device=dev) for _ in range(3))
        rts, rkids = torch.empty(N, device=dev), torch.empty(N, dtype=torch.int32, device=dev)
    ptr = lambda t: None if t is None or t.numel() == 0 else t.data_ptr()  # noqa: E731
    # Adam state for the five parameter tensors (learning rates x 1e-3 as in bench.py: the scene does not drift)
    keys = ("means3D", "shs", "opacities", "scales", "rotations")
    lrs = {"means3D": 1.6e-7, "shs": 2.5e-6, "opacities": 5e-5, "scales": 5e-6, "rotations": 1e-6}
    grad_of = {"means3D": 3, "shs": 5, "opacities": 2, "scales": 6, "rotations": 7}
    mom = {k: (torch.zeros_like(sc[k]), torch.zeros_like(sc[k])) for k in keys}
    step_no = [0]

    def sync():
        torch.cuda.synchronize()

    def ours_cut(tau):
        K = gh.expand_to_size(nodes, boxes, tau, cam["campos"], z3, ri, pi, ni)
        return K

    def ours_weights(tau, K):
        gh.get_interpolation_weights(ni[:K], tau, nodes, boxes, campos_cpu, z3, ts, kids)

    def ours_fwd_bwd(K):
        fa = (bg, ri[:K], pi[:K], ts[:K], kids[:K], sc["means3D"], ef, sc["opacities"], sc["scales"], sc["rotations"], 1.0, ef,
              cam["viewmatrix"], cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], H, W, sc["shs"], DEG, cam["campos"], False, False, True)
        R, color, radii, geom, binning, img, invdepth, seen = dgr.rasterize_gaussians(*fa)
        dL = (color - target) * (2.0 / color.numel())
        ba = (bg, ri[:K], pi[:K], ts[:K], kids[:K], sc["means3D"], radii, ef, sc["opacities"], sc["scales"], sc["rotations"], 1.0, ef,
              cam["viewmatrix"], cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], dL, torch.zeros_like(invdepth), sc["shs"], DEG,
              cam["campos"], geom, R, binning, img, False)
        grads = dgr.rasterize_gaussians_backward(*ba)
        return R, color, radii, grads

    def ours_adam(K, radii, grads):
        import ctypes as C
        relevant = ri[:K][radii > 0].long()
        step_no[0] += 1
        n = len(keys)
        P4 = C.c_void_p * n
        gs = [grads[grad_of[k]].contiguous() for k in keys]
        lodgs.check(lodgs.lib.lodgs_adam_step_multi(
            n, P4(*[sc[k].data_ptr() for k in keys]), P4(*[g.data_ptr() for g in gs]), P4(*[mom[k][0].data_ptr() for k in keys]),
            P4(*[mom[k][1].data_ptr() for k in keys]), relevant.data_ptr(), relevant.numel(), None, (C.c_int64 * n)(*[N] * n),
            (C.c_int32 * n)(*[sc[k].numel() // N for k in keys]), (C.c_double * n)(*[lrs[k] for k in keys]), (C.c_double * n)(*[0.9] * n),
            (C.c_double * n)(*[0.999] * n), (C.c_double * n)(*[1e-15] * n), (C.c_int64 * n)(*[step_no[0]] * n), 1, lodgs.current_stream()))
        return relevant.numel()

    def ref_cut(tau):
        return ref.ref_expand_to_size(N, float(tau), nodes.data_ptr(), boxes.data_ptr(), cam["campos"].data_ptr(), 0.0, 0.0, 0.0,
                                      rri.data_ptr(), rpi.data_ptr(), rni.data_ptr())

    def ref_weights(tau, K):
        v = campos_cpu
        ref.ref_get_ts_indexed(K, rni.data_ptr(), float(tau), nodes.data_ptr(), boxes.data_ptr(), float(v[0]), float(v[1]), float(v[2]), 0.0,
                               0.0, 0.0, rts.data_ptr(), rkids.data_ptr(), None)

    def ref_fwd_bwd(K):
        # allocations + zero fills as RasterizeGaussiansCUDA / ...BackwardCUDA do (HR/rasterize_points.cu:73-92,178-187)
        color = torch.zeros(3, H, W, device=dev)
        invdepth = torch.zeros(1, H, W, device=dev)
        radii = torch.zeros(K, dtype=torch.int32, device=dev)
        rects = torch.zeros(K, 2, dtype=torch.int32, device=dev)
        seen = torch.zeros(K, dtype=torch.int32, device=dev)
        R = ref.ref_forward(K, DEG, M, ptr(bg), W, H, ptr(rri[:K]), ptr(rpi[:K]), ptr(rts[:K]), ptr(rkids[:K]), ptr(sc["means3D"]), ptr(sc["shs"]),
                            None, ptr(sc["opacities"]), ptr(sc["scales"]), 1.0, ptr(sc["rotations"]), None, ptr(cam["viewmatrix"]),
                            ptr(cam["projmatrix"]), ptr(cam["campos"]), cam["tanfovx"], cam["tanfovy"], 0, ptr(color), ptr(invdepth), ptr(radii),
                            ptr(rects), 0, ptr(seen))
        dL = (color - target) * (2.0 / color.numel())
        dLi = torch.zeros_like(invdepth)
        zz = lambda *s: torch.zeros(*s, device=dev)  # noqa: E731
        g2, gc, go, gcol, gi = zz(N, 3), zz(N, 2, 2), zz(N, 1), zz(N, 3), zz(N, 1)
        g3, gcov, gsh, gs, gr = zz(N, 3), zz(N, 6), zz(N, M, 3), zz(N, 3), zz(N, 4)
        ref.ref_backward(K, DEG, M, R, ptr(bg), W, H, ptr(rri[:K]), ptr(rpi[:K]), ptr(rts[:K]), ptr(rkids[:K]), ptr(sc["means3D"]), ptr(sc["shs"]),
                         None, ptr(sc["opacities"]), ptr(sc["scales"]), 1.0, ptr(sc["rotations"]), None, ptr(cam["viewmatrix"]),
                         ptr(cam["projmatrix"]), ptr(cam["campos"]), cam["tanfovx"], cam["tanfovy"], ptr(radii), ptr(dL), ptr(dLi), ptr(g2), ptr(gc),
                         ptr(go), ptr(gcol), ptr(gi), ptr(g3), ptr(gcov), ptr(gsh), ptr(gs), ptr(gr), 0)
        return R, color, radii, (g2, gcol, go, g3, gcov, gsh, gs, gr)

    def timed(fn, iters):
        fn()
        sync()
        t0 = time.perf_counter()
        for _ in range(iters):
            fn()
        sync()
        return (time.perf_counter() - t0) * 1e3 / iters

    out = {"nodes": N, "gaussians": N, "view": "%dx%d" % (W, H), "iters": args.iters, "taus": {}}
    for tau_px in (0.0, 3.0, 6.0, 15.0):
        tau = synth.tau_to_threshold(tau_px, cam["tanfovx"], W)
        K = ours_cut(tau)
        ours_weights(tau, K)
        R, color, radii, grads = ours_fwd_bwd(K)
        sync()
        ent = {"K": int(K), "num_rendered": int(R), "visible": int((radii > 0).sum()),
               "nonfinite_grads_ours": int(sum((~torch.isfinite(g)).sum() for g in grads))}
        if ref is not None:   # parity first, on untouched parameters (the Adam steps below move them)
            Kr = ref_cut(tau)
            ref_weights(tau, Kr)
            Rr, rcolor, rradii, rgrads = ref_fwd_bwd(Kr)
            sync()
            rel = lambda a, b: float((a - b).norm() / b.norm().clamp_min(1e-30))  # noqa: E731
            ent["nonfinite_grads_reference"] = int(sum((~torch.isfinite(g)).sum() for g in rgrads))
            if os.environ.get("HIER_DEBUG"):
                print("DEBUG R", R, Rr, "color sums", float(color.sum()), float(rcolor.sum()), "radii>0", int((radii > 0).sum()), int((rradii > 0).sum()),
                      "norms ours", [float(g.norm()) for g in grads], "ref", [float(g.norm()) for g in rgrads], file=sys.stderr)
            ent["bit_exact"] = {"K": bool(Kr == K), "cut": bool(torch.equal(ri[:K], rri[:K]) and torch.equal(pi[:K], rpi[:K])),
                                "weights": bool(torch.equal(ts[:K], rts[:K]) and torch.equal(kids[:K], rkids[:K])),
                                "num_rendered": bool(Rr == R), "radii": bool(torch.equal(radii, rradii)), "image": bool(torch.equal(color, rcolor))}
            ent["grad_rel_l2"] = {n: round(rel(grads[i], rgrads[i]), 7) for n, i in (("dL_dopacity", 2), ("dL_dsh", 5), ("dL_dscale", 6),
                                                                                    ("dL_drot", 7), ("dL_dmean3D", 3))}
            del rgrads, rcolor
        if os.environ.get("HIER_DEBUG"):
            lodgs.profile(True)
            for _ in range(5):
                ours_fwd_bwd(K)
            sync()
            prof = lodgs.profile_dump()
            lodgs.profile(False)
            for name, (ms, n) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
                print("   %-28s %4.1f/iter %.4f ms/iter" % (name, n / 5, ms / 5), file=sys.stderr)
        o = {"cut_ms": timed(lambda: ours_cut(tau), args.iters), "weights_ms": timed(lambda: ours_weights(tau, K), args.iters),
             "fwd_bwd_ms": timed(lambda: ours_fwd_bwd(K), args.iters)}
        R, color, radii, grads = ours_fwd_bwd(K)
        o["adam_ms"] = timed(lambda: ours_adam(K, radii, grads), args.iters)

        def full_step():
            k2 = ours_cut(tau)
            ours_weights(tau, k2)
            r2 = ours_fwd_bwd(k2)
            ours_adam(k2, r2[2], r2[3])
        o["step_ms"] = timed(full_step, args.iters)
        ent["ours"] = {k: round(v, 4) for k, v in o.items()}
        if ref is not None:
            r = {"cut_ms": timed(lambda: ref_cut(tau), args.iters), "weights_ms": timed(lambda: ref_weights(tau, Kr), args.iters),
                 "fwd_bwd_ms": timed(lambda: ref_fwd_bwd(Kr), max(3, args.iters // 2))}
            ent["reference"] = {k: round(v, 4) for k, v in r.items()}
        out["taus"][str(tau_px)] = ent
        print("tau %.0f px: %s" % (tau_px, json.dumps(ent)), file=sys.stderr, flush=True)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
