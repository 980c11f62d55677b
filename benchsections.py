"""benchsections.py — the other BASELINE.json configurations, measured next to the headline workload A by bench.py and printed as extra
sections of its one JSON line (both arms):

  hier_step  configs[2]  single-chunk training step WITH LOD interpolation, 4 999 999-node hierarchy, tau in {0,3,6,15} px, 1080p
  eval_path  configs[3]  per-frame cut + weights + forward over the first 1000 poses of the reference's cameras.json, 49 999 999-node
                         hierarchy, the path sharded contiguously over the ranks (no collective)
  dp_train   configs[4]  multi-view training of one chunk through the alt-rasterizer surface, one view per rank, gradient exchange over NCCL
  creator    configs[0]  hierarchy construction (GPU builder next to the reference's CPU creator binary) and the CPU LOD cut (tau sweep)

Hierarchies are built on the device by lodgs.hier_build (kd-tree exactly as PointbasedKdTreeGenerator, merged Gaussians by the
ClusterMerger formulas, node ids in upstream populateRec order) from seeded synthetic leaves, so every rank and both arms see identical
bits. The `ours` arm runs the product path (plugin modules over liblodgs.so); the `reference` arm runs the reference's own kernels
(oracle/_ref: Switching::expandToSize / getTsIndexed, the hierarchy rasterizer through ITS torch extension, the alt-rasterizer, its
adamUpdateCUDA, scene/OurAdam.py's update) on the same inputs. Only bench.py imports this module.
"""
import math
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
W, H, DEG, M = 1920, 1080, 3, 16
TAUS = (0.0, 3.0, 6.0, 15.0)
HIER_STEP_LEAVES = 2_500_000      # N = 4 999 999 nodes (SURVEY §8d scene B)
EVAL_PATH_LEAVES = 25_000_000     # N = 49 999 999 nodes (scene C)
LRS = {"means3D": 1.6e-7, "shs": 2.5e-6, "opacities": 5e-5, "scales": 5e-6, "rotations": 1e-6}   # training rates x 1e-3: the scene does not drift
GRAD_OF = {"means3D": 3, "shs": 5, "opacities": 2, "scales": 6, "rotations": 7}
KEYS = ("means3D", "shs", "opacities", "scales", "rotations")


def _sync():
    torch.cuda.synchronize()


def _max_over_ranks(x, world, dev):
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    return float(t.item())


def _timed_ms(fn, iters, warm=1):
    for _ in range(warm):
        fn()
    _sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    _sync()
    return e0.elapsed_time(e1) / iters


# ------------------------------------------------------------------------------------------------ scenes
def synth_leaves(L, half_extent, dev, seed):
    """Leaf Gaussians of scenes B / C: positions uniform in the slab, scales as workload A's x 5 (log-normal around 0.1 m),
    random rotations, opacities sigmoid(N(0, 2^2)), SH as workload A. Generated on the device from a seeded generator."""
    g = torch.Generator(device=dev).manual_seed(seed)
    ext = torch.tensor(half_extent, device=dev, dtype=torch.float32)
    pos = (torch.rand(L, 3, device=dev, generator=g) * 2 - 1) * ext
    scales = torch.exp(torch.randn(L, 3, device=dev, generator=g) * 0.6 + math.log(0.1))
    rot = torch.nn.functional.normalize(torch.randn(L, 4, device=dev, generator=g), dim=1)
    opac = torch.sigmoid(torch.randn(L, device=dev, generator=g) * 2.0)
    shs = torch.randn(L, M, 3, device=dev, generator=g) * 0.1
    shs[:, 0] = (torch.randn(L, 3, device=dev, generator=g) * 0.5 + 0.5) / 0.2821
    return pos, scales, rot, opac, shs.view(L, 3 * M)


def build_scene(L, half_extent, dev, seed=7):
    """-> (nodes[N,7], boxes[N,2,4], parameter dict with one Gaussian per node, build milliseconds)"""
    from lodgs import hier_build
    leaves = synth_leaves(L, half_extent, dev, seed)
    _sync()
    t0 = time.perf_counter()
    h = hier_build.build_hierarchy(*leaves, dynamic=False, static=True)
    _sync()
    ms = (time.perf_counter() - t0) * 1e3
    del leaves
    N = h["pos"].size(0)
    sc = {"means3D": h["pos"], "scales": h["scales"], "rotations": h["rotations"], "opacities": h["opacities"].clamp(max=1.0).view(N, 1).contiguous(),
          "shs": h["shs"].view(N, M, 3)}
    return h["nodes"], h["boxes"], sc, ms


def _dev_cam(c, dev):
    d = {k: (torch.from_numpy(np.ascontiguousarray(v, dtype=np.float32)).to(dev) if isinstance(v, np.ndarray) else v) for k, v in c.items()}
    d["campos_cpu"] = d["campos"].cpu()
    return d


def camera_path(n, half_extent, dev, lo=0, hi=None):
    """The first n poses of the reference's cameras.json (tests/golden/camera_path_1000.npz, extracted by tests/golden/make_camera_path.py:
    camera centres and camera-to-world rotations of a street-level capture, z up), mapped affinely into the slab: the path's bounding
    rectangle is centred and scaled uniformly to 80 % of the slab's width, the camera height put 2 m above the mid-plane; rotations are
    kept. Intrinsics are the benchmark's (1080p, fx = fy = 1200)."""
    from lodgs import synth
    f = np.load(os.path.join(ROOT, "tests", "golden", "camera_path_1000.npz"))
    pos, rot = f["position"][:n].astype(np.float64), f["rotation"][:n].astype(np.float64)
    c = 0.5 * (pos.min(0) + pos.max(0))
    s = 0.8 * 2.0 * half_extent[0] / float((pos.max(0) - pos.min(0))[:2].max())
    p = (pos - c) * s
    p[:, 2] = 2.0 + (pos[:, 2] - c[2])
    cams = []
    for i in range(lo, n if hi is None else hi):
        R = rot[i]
        cams.append(_dev_cam(synth.make_camera(W=W, H=H, fx=1200.0, fy=1200.0, R=R, T=-R.T @ p[i]), dev))
    return cams


# ------------------------------------------------------------------------------------------------ arms (one hierarchical frame)
class OursHier:
    def __init__(self, nodes, boxes, sc, dev):
        from diff_gaussian_rasterization import _C as dgr
        from gaussian_hierarchy import _C as gh
        self.dgr, self.gh, self.nodes, self.boxes, self.sc, self.dev = dgr, gh, nodes, boxes, sc, dev
        N = nodes.size(0)
        self.N = N
        self.ri, self.pi, self.ni = (torch.empty(N, dtype=torch.int32, device=dev) for _ in range(3))
        self.ts, self.kids = torch.empty(N, device=dev), torch.empty(N, dtype=torch.int32, device=dev)
        self.bg, self.ef, self.z3 = torch.zeros(3, device=dev), torch.empty(0, device=dev), torch.zeros(3)
        gh.prepare_hierarchy(nodes, boxes)

    def cut(self, cam, tau):
        return self.gh.expand_to_size(self.nodes, self.boxes, tau, cam["campos"], self.z3, self.ri, self.pi, self.ni)

    def weights(self, cam, tau, K):
        self.gh.get_interpolation_weights(self.ni[:K], tau, self.nodes, self.boxes, cam["campos_cpu"], self.z3, self.ts, self.kids)

    def forward(self, cam, K, do_depth):
        s = self.sc
        return self.dgr.rasterize_gaussians(self.bg, self.ri[:K], self.pi[:K], self.ts[:K], self.kids[:K], s["means3D"], self.ef, s["opacities"],
                                            s["scales"], s["rotations"], 1.0, self.ef, cam["viewmatrix"], cam["projmatrix"], cam["tanfovx"],
                                            cam["tanfovy"], H, W, s["shs"], DEG, cam["campos"], False, False, do_depth)

    def backward(self, cam, K, fwd, dL):
        s = self.sc
        R, color, radii, geom, binning, img, invdepth, seen = fwd
        return self.dgr.rasterize_gaussians_backward(self.bg, self.ri[:K], self.pi[:K], self.ts[:K], self.kids[:K], s["means3D"], radii, self.ef,
                                                     s["opacities"], s["scales"], s["rotations"], 1.0, self.ef, cam["viewmatrix"],
                                                     cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], dL, torch.zeros_like(invdepth), s["shs"],
                                                     DEG, cam["campos"], geom, R, binning, img, False)

    def lists(self, K):
        return self.ri[:K], self.pi[:K], self.ni[:K], self.ts[:K], self.kids[:K]


def _ref_ext():
    """The reference's own torch extension (HR/rasterize_points.cu + ext.cpp, built by oracle/Makefile `refext`), or None."""
    d = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.exists(os.path.join(d, "ref_dgr", "_C.so")):
        return None
    if d not in sys.path:
        sys.path.insert(0, d)
    from ref_dgr import _C
    return _C


class RefHier(OursHier):
    """Same protocol over the reference's kernels: the cut through oracle/ref_shim.cu (Switching::expandToSize / getTsIndexed take raw
    pointers; their torch binding needs Eigen-typed loaders), the rasterizer through the reference's own `_C.rasterize_gaussians`."""

    def __init__(self, nodes, boxes, sc, dev):
        from oracle import refshim
        self.lib, self.ext = refshim.lib(), _ref_ext()
        if self.ext is None:
            raise FileNotFoundError("oracle/_ref/ref_dgr/_C.so")
        self.nodes, self.boxes, self.sc, self.dev = nodes, boxes, sc, dev
        N = nodes.size(0)
        self.N = N
        self.ri, self.pi, self.ni = (torch.empty(N, dtype=torch.int32, device=dev) for _ in range(3))
        self.ts, self.kids = torch.empty(N, device=dev), torch.empty(N, dtype=torch.int32, device=dev)
        self.bg, self.ef = torch.zeros(3, device=dev), torch.empty(0, device=dev)

    def cut(self, cam, tau):
        return self.lib.ref_expand_to_size(self.N, float(tau), self.nodes.data_ptr(), self.boxes.data_ptr(), cam["campos"].data_ptr(), 0.0, 0.0, 0.0,
                                           self.ri.data_ptr(), self.pi.data_ptr(), self.ni.data_ptr())

    def weights(self, cam, tau, K):
        v = cam["campos_cpu"]
        self.lib.ref_get_ts_indexed(K, self.ni.data_ptr(), float(tau), self.nodes.data_ptr(), self.boxes.data_ptr(), float(v[0]), float(v[1]),
                                    float(v[2]), 0.0, 0.0, 0.0, self.ts.data_ptr(), self.kids.data_ptr(), None)

    def forward(self, cam, K, do_depth):
        s = self.sc
        return self.ext.rasterize_gaussians(self.bg, self.ri[:K], self.pi[:K], self.ts[:K], self.kids[:K], s["means3D"], self.ef, s["opacities"],
                                            s["scales"], s["rotations"], 1.0, self.ef, cam["viewmatrix"], cam["projmatrix"], cam["tanfovx"],
                                            cam["tanfovy"], H, W, s["shs"], DEG, cam["campos"], False, False, do_depth)

    def backward(self, cam, K, fwd, dL):
        s = self.sc
        R, color, radii, geom, binning, img, invdepth, seen = fwd
        return self.ext.rasterize_gaussians_backward(self.bg, self.ri[:K], self.pi[:K], self.ts[:K], self.kids[:K], s["means3D"], radii, self.ef,
                                                     s["opacities"], s["scales"], s["rotations"], 1.0, self.ef, cam["viewmatrix"],
                                                     cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], dL, torch.zeros_like(invdepth), s["shs"],
                                                     DEG, cam["campos"], geom, R, binning, img, False)


def _make_arm(impl, nodes, boxes, sc, dev):
    return OursHier(nodes, boxes, sc, dev) if impl == "ours" else RefHier(nodes, boxes, sc, dev)


# ------------------------------------------------------------------------------------------------ B: hierarchical training step
def hier_step(impl, dev, iters=10, leaves=HIER_STEP_LEAVES, check=True):
    """cut -> weights -> forward with in-kernel LOD blending -> backward -> sparse Adam on the visible rows, per tau."""
    from lodgs import synth
    half = (200.0, 200.0, 10.0)
    nodes, boxes, sc, build_ms = build_scene(leaves, half, dev)
    N = nodes.size(0)
    cam = _dev_cam(synth.look_at_camera(eye=(-150.0, 20.0, 18.0), target=(0.0, 0.0, 0.0), up=(0.0, 0.0, 1.0), W=W, H=H, fx=1200.0, fy=1200.0), dev)
    target = torch.rand(3, H, W, device=dev, generator=torch.Generator(device=dev).manual_seed(1))
    arm = _make_arm(impl, nodes, boxes, sc, dev)
    other = None
    if check and impl == "ours":
        try:
            other = RefHier(nodes, boxes, sc, dev)
        except (FileNotFoundError, OSError):
            other = None
    # optimizer: lodgs.optim.Adam (ours) / scene/OurAdam.py's update (reference; the imported class where the reference tree exists)
    params = {k: torch.nn.Parameter(sc[k]) for k in KEYS}
    if impl == "ours":
        from lodgs.optim import Adam
        opt, opt_name = Adam([{"params": [params[k]], "lr": LRS[k]} for k in KEYS], lr=0.0, eps=1e-15), "lodgs.optim.Adam (lodgs_adam_step_multi)"
    else:
        from oracle import ouradam
        cls, which = ouradam.reference_or_restated()
        opt, opt_name = cls([{"params": [params[k]], "lr": LRS[k]} for k in KEYS], lr=0.0, eps=1e-15), "scene/OurAdam.py Adam.step(relevant) [%s]" % which

    def fwd_bwd(K):
        f = arm.forward(cam, K, True)
        dL = (f[1] - target) * (2.0 / f[1].numel())
        return f, arm.backward(cam, K, f, dL)

    def adam(K, f, grads):
        relevant = arm.ri[:K][f[2] > 0].long()
        for k in KEYS:
            params[k].grad = grads[GRAD_OF[k]]
        opt.step(relevant)

    def step(tau):
        K = arm.cut(cam, tau)
        arm.weights(cam, tau, K)
        f, g = fwd_bwd(K)
        adam(K, f, g)

    out = {"impl": impl, "nodes": N, "gaussians": N, "leaves": leaves, "view": "%dx%d" % (W, H), "iters": iters, "build_ms": round(build_ms, 1),
           "id_order": "upstream populateRec (children in contiguous blocks, blocks in pre-order)", "optimizer": opt_name, "taus": {}}
    for tau_px in TAUS:
        tau = synth.tau_to_threshold(tau_px, cam["tanfovx"], W)
        K = arm.cut(cam, tau)
        arm.weights(cam, tau, K)
        f, g = fwd_bwd(K)
        _sync()
        ent = {"K": int(K), "num_rendered": int(f[0]), "visible": int((f[2] > 0).sum())}
        if other is not None:    # parity on untouched parameters, before the Adam steps below move them
            Kr = other.cut(cam, tau)
            other.weights(cam, tau, Kr)
            fr = other.forward(cam, Kr, True)
            gr = other.backward(cam, Kr, fr, (fr[1] - target) * (2.0 / fr[1].numel()))
            _sync()
            same = Kr == K
            rel = lambda a, b: float((a - b).norm() / b.norm().clamp_min(1e-30))  # noqa: E731
            ent["bit_exact"] = {"K": bool(same), "cut": bool(same and all(torch.equal(a, b) for a, b in zip(arm.lists(K)[:3], other.lists(K)[:3]))),
                                "weights": bool(same and torch.equal(arm.ts[:K], other.ts[:K]) and torch.equal(arm.kids[:K], other.kids[:K])),
                                "num_rendered": bool(fr[0] == f[0]), "radii": bool(same and torch.equal(f[2], fr[2])),
                                "image": bool(torch.equal(f[1], fr[1]))}
            ent["grad_rel_l2"] = {n: round(rel(g[i], gr[i]), 7) for n, i in (("dL_dopacity", 2), ("dL_dsh", 5), ("dL_dscale", 6), ("dL_drot", 7),
                                                                             ("dL_dmean3D", 3))}
            del fr, gr
        ph = {"cut_ms": _timed_ms(lambda: arm.cut(cam, tau), iters), "weights_ms": _timed_ms(lambda: arm.weights(cam, tau, K), iters),
              "fwd_bwd_ms": _timed_ms(lambda: fwd_bwd(K), iters)}
        f, g = fwd_bwd(K)
        ph["adam_ms"] = _timed_ms(lambda: adam(K, f, g), iters)
        ph["step_ms"] = _timed_ms(lambda: step(tau), iters)
        ph["train_iters_per_s"] = 1e3 / ph["step_ms"]
        ph["cut_flat_equiv_GBps"] = (60.0 * N + 12.0 * K) / (ph["cut_ms"] * 1e-3) / 1e9
        ent.update({k: round(v, 4) for k, v in ph.items()})
        out["taus"]["%g" % tau_px] = ent
    if impl == "ours":
        import lodgs
        tau = synth.tau_to_threshold(3.0, cam["tanfovx"], W)
        lodgs.profile(True)
        for _ in range(5):
            step(tau)
        _sync()
        prof = lodgs.profile_dump()
        lodgs.profile(False)
        out["per_kernel_ms_tau3"] = {name: round(ms / 5, 4) for name, (ms, n) in sorted(prof.items(), key=lambda kv: -kv[1][0])}
    return out


# ------------------------------------------------------------------------------------------------ C: evaluation over a camera path
def eval_path(impl, rank, world, dev, frames=1000, tau_px=3.0, leaves=EVAL_PATH_LEAVES, check_every=100):
    from lodgs import parallel, synth
    side = 200.0 * math.sqrt(leaves / HIER_STEP_LEAVES)     # scene B's density
    half = (side, side, 10.0)
    nodes, boxes, sc, build_ms = build_scene(leaves, half, dev)
    N = nodes.size(0)
    lo, hi = parallel.shard_range(frames, rank, world)
    cams = camera_path(frames, half, dev, lo, hi)
    arm = _make_arm(impl, nodes, boxes, sc, dev)
    other = None
    if impl == "ours" and check_every:
        try:
            other = RefHier(nodes, boxes, sc, dev)
        except (FileNotFoundError, OSError):
            other = None

    def frame(a, cam, phases=None):
        tau = synth.tau_to_threshold(tau_px, cam["tanfovx"], W)
        t0 = time.perf_counter()
        K = a.cut(cam, tau)
        if phases is not None:
            _sync(); t1 = time.perf_counter()
        a.weights(cam, tau, K)
        if phases is not None:
            _sync(); t2 = time.perf_counter()
        f = a.forward(cam, K, False)
        if phases is not None:
            _sync(); t3 = time.perf_counter()
            phases["cut_ms"] += (t1 - t0) * 1e3; phases["weights_ms"] += (t2 - t1) * 1e3; phases["forward_ms"] += (t3 - t2) * 1e3
        return K, f

    chk = {"checked": 0, "cut": 0, "weights": 0, "num_rendered": 0, "image": 0}
    Ks, Rs = [], []
    for j, cam in enumerate(cams):
        if j % max(1, check_every):
            continue
        K, f = frame(arm, cam)
        Ks.append(int(K)); Rs.append(int(f[0]))
        if other is not None:
            Kr, fr = frame(other, cam)
            _sync()
            same = Kr == K
            chk["checked"] += 1
            chk["cut"] += int(same and all(torch.equal(a, b) for a, b in zip(arm.lists(K)[:3], other.lists(K)[:3])))
            chk["weights"] += int(same and torch.equal(arm.ts[:K], other.ts[:K]) and torch.equal(arm.kids[:K], other.kids[:K]))
            chk["num_rendered"] += int(fr[0] == f[0])
            chk["image"] += int(torch.equal(f[1], fr[1]))
            del fr
    _sync()
    n_ph = min(10, len(cams))
    ph = {"cut_ms": 0.0, "weights_ms": 0.0, "forward_ms": 0.0}
    for cam in cams[:n_ph]:
        frame(arm, cam, ph)
    per_kernel = None
    if impl == "ours" and rank == 0:
        import lodgs
        lodgs.profile(True)
        for cam in cams[:n_ph]:
            frame(arm, cam)
        _sync()
        prof = lodgs.profile_dump()
        lodgs.profile(False)
        per_kernel = {name: round(ms / n_ph, 4) for name, (ms, n) in sorted(prof.items(), key=lambda kv: -kv[1][0])}
    for cam in cams[:3]:
        frame(arm, cam)
    _sync()
    if world > 1:
        torch.distributed.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for cam in cams:
        frame(arm, cam)
    e1.record()
    _sync()
    t = _max_over_ranks(e0.elapsed_time(e1) * 1e-3, world, dev)
    c = torch.tensor([chk[k] for k in ("checked", "cut", "weights", "num_rendered", "image")] + [min(Ks), -max(Ks), min(Rs), -max(Rs)],
                     dtype=torch.int64, device=dev)
    if world > 1:
        torch.distributed.all_reduce(c[:5])
        torch.distributed.all_reduce(c[5:], op=torch.distributed.ReduceOp.MIN)
    c = c.tolist()
    out = {"impl": impl, "metric": "eval frames/s (cut + weights + forward, 1080p)", "n_gpus": world, "frames": frames, "frames_per_rank": len(cams),
           "nodes": N, "gaussians": N, "leaves": leaves, "tau_px": tau_px, "scaling": "strong", "build_ms": round(build_ms, 1),
           "camera_path": "first %d poses of the reference's cameras.json mapped into the slab" % frames,
           "K_min_max": [c[5], -c[6]], "num_rendered_min_max": [c[7], -c[8]],
           "frames_per_s": round(frames / t, 2), "ms_per_frame_per_gpu": round(t * 1e3 / max(1, len(cams)), 4),
           "mpix_per_s": round(frames * W * H / t / 1e6, 1), "phases_ms_rank0": {k: round(v / n_ph, 4) for k, v in ph.items()},
           "lod_cut_ms_per_frame": round(ph["cut_ms"] / n_ph, 4)}
    if per_kernel:
        out["per_kernel_ms"] = per_kernel
    if other is not None:
        out["bit_exact_frames"] = dict(zip(("checked", "cut", "weights", "num_rendered", "image"), c[:5]))
    return out


# ------------------------------------------------------------------------------------------------ D: data-parallel multi-view training
DP_LRS = {"means3D": 1.6e-7, "dc": 2.5e-6, "shs_rest": 1.25e-7, "opacities": 5e-5, "scales": 5e-6, "rotations": 1e-6}


def dp_train(impl, rank, world, dev, steps=20, warmup=5, n_gauss=1_000_000, mode="auto"):
    """Per step and rank: forward + backward of the rank's own 1080p view of the replicated scene through the alt-rasterizer surface,
    exchange of the parameter gradients, sparse Adam on identical sums (replicas are checked identical at the end)."""
    import torch.distributed as dist
    from lodgs import parallel, synth
    cam_np = synth.make_camera(W=W, H=H, fx=1200.0, fy=1200.0, T=(0.25 * (rank - 0.5 * (world - 1)), 0.0, 0.0))
    scn = synth.make_gaussians(n_gauss, seed=0, tanfovx=cam_np["tanfovx"], tanfovy=cam_np["tanfovy"])
    cam = _dev_cam(cam_np, dev)
    shs = torch.from_numpy(scn["shs"]).to(dev)
    par = {"means3D": torch.from_numpy(scn["means3D"]).to(dev), "dc": shs[:, :1].contiguous(), "shs_rest": shs[:, 1:].contiguous(),
           "opacities": torch.from_numpy(scn["opacities"]).to(dev), "scales": torch.from_numpy(scn["scales"]).to(dev),
           "rotations": torch.from_numpy(scn["rotations"]).to(dev)}
    target = torch.rand(3, H, W, device=dev, generator=torch.Generator(device=dev).manual_seed(100 + rank))
    bg = torch.zeros(3, device=dev)
    names = list(DP_LRS)
    order = ["opacities", "shs_rest", "means3D", "dc", "scales", "rotations"]
    N = n_gauss
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    if impl == "ours":
        from lodgs import dp
        trainer = dp.MultiViewTrainer(par, DP_LRS, cam, bg, W, H, DEG, rank, world, mode=mode)
        step = lambda timed: trainer.step(target, timed)  # noqa: E731
        mode_used = trainer.mode
    else:
        from oracle import refshim_ar as rar
        if not rar.available():
            return {"impl": "reference", "unavailable": "oracle/_ref/libref_ar.so not built"}
        Lb = rar.lib()
        sc = dict(par, sh_degree=DEG)
        mom = {k: (torch.zeros_like(par[k]), torch.zeros_like(par[k])) for k in names}
        gname = {"means3D": "dL_dmean3D", "dc": "dL_ddc", "shs_rest": "dL_dsh", "opacities": "dL_dopacity", "scales": "dL_dscale", "rotations": "dL_drot"}
        mode_used = "allreduce (the reference has no multi-GPU path: its kernels inside the same exchange)"

        def step(timed):
            e = [ev() for _ in range(4)] if timed else None
            if timed:
                e[0].record()
            fwd = rar.forward(sc, cam, bg, True)
            dL = (fwd["color"] - target) * (2.0 / fwd["color"].numel())
            g = rar.backward(sc, cam, bg, fwd, dL, torch.zeros_like(fwd["invdepth"]), True)
            if timed:
                e[1].record()
            nbytes = parallel.allreduce_gradients([g[gname[k]] for k in names])
            if timed:
                e[2].record()
            vis = (g["dL_dopacity"] != 0).reshape(-1)
            for k in names:
                Lb.refar_adam(rar.ptr(par[k]), rar.ptr(g[gname[k]]), rar.ptr(mom[k][0]), rar.ptr(mom[k][1]), rar.ptr(vis), DP_LRS[k], 0.9, 0.999,
                              1e-15, N, par[k].numel() // N)
            if timed:
                e[3].record()
            return e, nbytes
    for _ in range(max(3, warmup)):
        step(False)
    _sync()
    if world > 1:
        dist.barrier()
    t0, t1 = ev(), ev()
    t0.record()
    marks = []
    nbytes = 0
    for _ in range(steps):
        e, nbytes = step(True)
        marks.append(e)
    t1.record()
    _sync()
    if world > 1:
        dist.barrier()
    phase = {"fwd_bwd_ms": 0.0, "exchange_ms": 0.0, "adam_ms": 0.0}
    for e in marks:
        phase["fwd_bwd_ms"] += e[0].elapsed_time(e[1]); phase["exchange_ms"] += e[1].elapsed_time(e[2]); phase["adam_ms"] += e[2].elapsed_time(e[3])
    t = torch.tensor([t0.elapsed_time(t1)] + [phase[k] for k in phase], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    live = trainer.par if impl == "ours" else par
    cs = torch.stack([live[k].detach().double().sum() for k in names])
    lo_, hi_ = cs.clone(), cs.clone()
    if world > 1:
        dist.all_reduce(lo_, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi_, op=dist.ReduceOp.MAX)
    ms = float(t[0]) / steps
    ex_ms = float(t[2]) / steps
    return {"impl": impl, "metric": "multi-view chunk training views/s (fwd + bwd + gradient exchange + sparse Adam, 1080p, alt-rasterizer surface)",
            "n_gpus": world, "steps": steps, "gaussians": N, "ms_per_step": round(ms, 4), "views_per_s": round(world * 1e3 / ms, 2),
            "mpix_per_s": round(world * W * H / ms / 1e3, 1), "scaling": "weak", "exchange": mode_used,
            "phases_ms_max_over_ranks": {k: round(float(v) / steps, 4) for k, v in zip(phase, t[1:])},
            "exchange_bytes_per_rank": int(nbytes),
            "exchange_bus_GBps": round(2 * (world - 1) / world * nbytes / (ex_ms * 1e-3) / 1e9, 1) if world > 1 and ex_ms > 0 else None,
            "replicas_identical": bool(torch.equal(lo_, hi_))}


# ------------------------------------------------------------------------------------------------ configs[0]: creator + CPU cut
def creator_and_cpu_cut(dev, sample_points=200_000, cut_leaves=HIER_STEP_LEAVES, cpu_budget_s=20.0):
    """GPU hierarchy construction next to the reference's CPU creator (oracle/_ref/GaussianHierarchyCreator: the UNMODIFIED
    mainHierarchyCreator.cpp pipeline, single-threaded as the reference is) on the same seeded PLY — a bounded sample, since the bundled
    input.ply is absent from the reference tree — and the CPU LOD cut (oracle/cut_oracle.c, the restated kernels) on the 5 M-node
    hierarchy, tau sweep, single-threaded and over all host cores, next to the GPU cut."""
    import tempfile
    from gaussian_hierarchy import _C as gh
    from lodgs import hier_build, synth
    from oracle import oracle as orc
    from oracle import ply as oply
    from oracle import refshim_hier as rh
    out = {"sample_points": sample_points}
    with tempfile.TemporaryDirectory() as td:
        plyf = os.path.join(td, "in.ply")
        oply.write_less_rich_ply(plyf, *oply.synth_points(sample_points, seed=1, extent=(400.0, 400.0, 20.0), scale_mean=0.1))
        hier_build.create_dynamic_hierarchy(plyf, os.path.join(td, "warm_"))
        _sync()
        t0 = time.perf_counter()
        h = hier_build.create_dynamic_hierarchy(plyf, os.path.join(td, "ours_"))
        _sync()
        out["gpu_create_s"] = round(time.perf_counter() - t0, 4)
        recs = hier_build.read_less_rich_ply(plyf)
        inp = hier_build.activate_and_filter(recs, dev)
        out["gpu_build_only_ms"] = round(_timed_ms(lambda: hier_build.build_hierarchy(*inp, dynamic=True, static=False), 3), 3)
        if rh.creator_available():
            out["cpu_reference_creator_s"] = round(rh.run_creator(plyf, os.path.join(td, "ref_")), 3)
            out["cpu_reference_creator"] = {"cores": 1, "kind": "reference", "binary": "oracle/_ref/GaussianHierarchyCreator (unmodified sources, Eigen stand-in)"}
            import filecmp
            out["dhier_files_byte_identical"] = bool(filecmp.cmp(os.path.join(td, "ours_hierarchy.dhier"), os.path.join(td, "ref_hierarchy.dhier"),
                                                                 shallow=False))
    # CPU cut tau sweep on the 5 M-node hierarchy
    nodes, boxes, sc, _ = build_scene(cut_leaves, (200.0, 200.0, 10.0), dev)
    N = nodes.size(0)
    nn, bb = nodes.cpu().numpy(), boxes.cpu().numpy()
    cam = synth.look_at_camera(eye=(-150.0, 20.0, 18.0), target=(0.0, 0.0, 0.0), up=(0.0, 0.0, 1.0), W=W, H=H, fx=1200.0, fy=1200.0)
    vp = cam["campos"]
    ri, pi, ni = (torch.empty(N, dtype=torch.int32, device=dev) for _ in range(3))
    bufs = tuple(np.zeros(N, np.int32) for _ in range(4))
    sweep = {}
    budget = cpu_budget_s / (2 * len(TAUS))
    for tau_px in TAUS:
        tau = synth.tau_to_threshold(tau_px, cam["tanfovx"], W)
        ent = {}
        t0, reps = time.perf_counter(), 0
        while reps < 1 or (time.perf_counter() - t0 < budget and reps < 20):
            r1 = orc.expand_to_size(nn, bb, tau, vp)
            reps += 1
        ent["cpu_1_thread_ms"] = round((time.perf_counter() - t0) * 1e3 / reps, 2)
        t0, reps = time.perf_counter(), 0
        while reps < 1 or (time.perf_counter() - t0 < budget and reps < 50):
            rk = orc.expand_to_size_omp(nn, bb, tau, vp, bufs)
            reps += 1
        ent["cpu_all_cores_ms"] = round((time.perf_counter() - t0) * 1e3 / reps, 2)
        ent["cpu_threads"] = int(rk[3])
        vpd = torch.from_numpy(vp).to(dev)
        K = gh.expand_to_size(nodes, boxes, tau, vpd, torch.zeros(3), ri, pi, ni)
        ent["gpu_ms"] = round(_timed_ms(lambda: gh.expand_to_size(nodes, boxes, tau, vpd, torch.zeros(3), ri, pi, ni), 10), 4)
        ent["K"] = int(K)
        ent["identical"] = bool(K == len(r1[0]) and np.array_equal(ri[:K].cpu().numpy(), r1[0]) and np.array_equal(rk[0], r1[0])
                                and np.array_equal(ni[:K].cpu().numpy(), r1[2]))
        sweep["%g" % tau_px] = ent
    out["lod_cut_tau_sweep"] = {"nodes": N, "host_cores": os.cpu_count(), "taus": sweep}
    return out
