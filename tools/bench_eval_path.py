#!/usr/bin/env python3
"""BASELINE.json configs[3]: hierarchy evaluation over a camera path (render_hierarchy.py:render_set) — per frame
LOD cut (expand_to_size) -> interpolation weights -> forward render with in-kernel LOD blending at 1920x1080, on a synthetic
hierarchy of 2^(k+1)-1 nodes (one Gaussian per node), the camera path sharded in contiguous ranges over the ranks
(lodgs.parallel.shard_range), hierarchy and Gaussians replicated, NO data-path collective. The reference arm (oracle/_ref:
Switching::expandToSize / getTsIndexed and the hierarchy rasterizer, unmodified) runs the same frames on the same GPU; every
`--check-every`-th frame the cut lists, weights, num_rendered and the IMAGE of both arms are compared bit for bit.

  python tools/bench_eval_path.py [--k 24] [--frames 1000] [--tau 3]                          (1 GPU)
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29531 \
      tools/bench_eval_path.py --frames 1000                                                  (N GPUs)

Per rank the loop is timed by wall clock between two device synchronisations (every frame already ends its cut and its forward
with a count on the host); the job's time is the MAX over ranks, value = frames of all ranks / that time. One JSON line on rank 0.
"""
import argparse
import json
import math
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "hierarchical-lod-gaussians_b200"), ROOT, os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

W, H, DEG, M = 1920, 1080, 3, 16


def camera_path(n, dev):
    """n cameras on a closed orbit around the 400x400x20 m slab, looking at its centre from 18 m above its mid plane; the radius
    breathes between 120 and 180 m so that the cut size varies along the path as it does on a real fly-through."""
    import numpy as np
    import torch
    from lodgs import synth
    cams = []
    for i in range(n):
        a = 2.0 * math.pi * i / n
        r = 150.0 + 30.0 * math.sin(3.0 * a)
        c = synth.look_at_camera(eye=(r * math.cos(a), r * math.sin(a), 18.0), target=(0.0, 0.0, 0.0), up=(0.0, 0.0, 1.0), W=W, H=H,
                                 fx=1200.0, fy=1200.0)
        d = {k: (torch.from_numpy(np.ascontiguousarray(v, dtype=np.float32)).to(dev) if isinstance(v, np.ndarray) else v)
             for k, v in c.items()}
        d["campos_cpu"] = d["campos"].cpu()
        cams.append(d)
    return cams


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--k", type=int, default=24, help="2^k leaves, 2^(k+1)-1 nodes (24: 33.5 M, 25: 67.1 M)")
    ap.add_argument("--frames", type=int, default=1000)
    ap.add_argument("--tau", type=float, default=3.0)
    ap.add_argument("--check-every", type=int, default=25)
    ap.add_argument("--no-reference", action="store_true")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from bench_hier_step import make_scene
    from diff_gaussian_rasterization import _C as dgr
    from gaussian_hierarchy import _C as gh
    from lodgs import parallel, synth
    from oracle import refshim
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    nodes, boxes, sc = make_scene(args.k, dev)
    N = nodes.size(0)
    lo, hi = parallel.shard_range(args.frames, rank, world)
    cams = camera_path(args.frames, dev)[lo:hi]
    bg = torch.zeros(3, device=dev)
    ef, z3 = torch.empty(0, device=dev), torch.zeros(3)
    ri, pi, ni = (torch.empty(N, dtype=torch.int32, device=dev) for _ in range(3))
    ts, kids = torch.empty(N, device=dev), torch.empty(N, dtype=torch.int32, device=dev)
    gh.prepare_hierarchy(nodes, boxes)
    ref = None if args.no_reference or not refshim.available() else refshim.lib()
    if ref is not None:
        rri, rpi, rni = (torch.empty(N, dtype=torch.int32, device=dev) for _ in range(3))
        rts, rkids = torch.empty(N, device=dev), torch.empty(N, dtype=torch.int32, device=dev)
    ptr = lambda t: None if t is None or t.numel() == 0 else t.data_ptr()  # noqa: E731

    def ours_frame(cam, phases=None):
        tau = synth.tau_to_threshold(args.tau, cam["tanfovx"], W)
        t0 = time.perf_counter()
        K = gh.expand_to_size(nodes, boxes, tau, cam["campos"], z3, ri, pi, ni)
        if phases is not None:
            torch.cuda.synchronize(); t1 = time.perf_counter()
        gh.get_interpolation_weights(ni[:K], tau, nodes, boxes, cam["campos_cpu"], z3, ts, kids)
        if phases is not None:
            torch.cuda.synchronize(); t2 = time.perf_counter()
        R, color, radii, _g, _b, _i, _d, _s = dgr.rasterize_gaussians(
            bg, ri[:K], pi[:K], ts[:K], kids[:K], sc["means3D"], ef, sc["opacities"], sc["scales"], sc["rotations"], 1.0, ef, cam["viewmatrix"],
            cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], H, W, sc["shs"], DEG, cam["campos"], False, False, False)
        if phases is not None:
            torch.cuda.synchronize(); t3 = time.perf_counter()
            phases["cut_ms"] += (t1 - t0) * 1e3; phases["weights_ms"] += (t2 - t1) * 1e3; phases["forward_ms"] += (t3 - t2) * 1e3
        return K, R, color

    def ref_frame(cam, phases=None):
        tau = float(synth.tau_to_threshold(args.tau, cam["tanfovx"], W))
        v = cam["campos_cpu"]
        t0 = time.perf_counter()
        K = ref.ref_expand_to_size(N, tau, nodes.data_ptr(), boxes.data_ptr(), cam["campos"].data_ptr(), 0.0, 0.0, 0.0, rri.data_ptr(),
                                   rpi.data_ptr(), rni.data_ptr())
        if phases is not None:
            torch.cuda.synchronize(); t1 = time.perf_counter()
        ref.ref_get_ts_indexed(K, rni.data_ptr(), tau, nodes.data_ptr(), boxes.data_ptr(), float(v[0]), float(v[1]), float(v[2]), 0.0, 0.0, 0.0,
                               rts.data_ptr(), rkids.data_ptr(), None)
        if phases is not None:
            torch.cuda.synchronize(); t2 = time.perf_counter()
        # allocations + zero fills as RasterizeGaussiansCUDA does (HR/rasterize_points.cu:73-92)
        color, invdepth = torch.zeros(3, H, W, device=dev), torch.zeros(1, H, W, device=dev)
        radii, rects = torch.zeros(K, dtype=torch.int32, device=dev), torch.zeros(K, 2, dtype=torch.int32, device=dev)
        seen = torch.zeros(K, dtype=torch.int32, device=dev)
        R = ref.ref_forward(K, DEG, M, ptr(bg), W, H, ptr(rri[:K]), ptr(rpi[:K]), ptr(rts[:K]), ptr(rkids[:K]), ptr(sc["means3D"]), ptr(sc["shs"]),
                            None, ptr(sc["opacities"]), ptr(sc["scales"]), 1.0, ptr(sc["rotations"]), None, ptr(cam["viewmatrix"]),
                            ptr(cam["projmatrix"]), ptr(cam["campos"]), cam["tanfovx"], cam["tanfovy"], 0, ptr(color), ptr(invdepth), ptr(radii),
                            ptr(rects), 0, ptr(seen))
        if phases is not None:
            torch.cuda.synchronize(); t3 = time.perf_counter()
            phases["cut_ms"] += (t1 - t0) * 1e3; phases["weights_ms"] += (t2 - t1) * 1e3; phases["forward_ms"] += (t3 - t2) * 1e3
        return K, R, color

    def job_time(fn):
        """max over ranks of the wall time of this rank's share between two barriers + device synchronisations"""
        for cam in cams[:3]:
            fn(cam)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for cam in cams:
            fn(cam)
        torch.cuda.synchronize()
        t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- parity on a subset of this rank's frames (before any timing), and the size of the work
    chk = {"frames": 0, "cut": 0, "weights": 0, "num_rendered": 0, "image": 0}
    Ks, Rs = [], []
    for j, cam in enumerate(cams):
        if j % args.check_every:
            continue
        K, R, color = ours_frame(cam)
        Ks.append(int(K)); Rs.append(int(R))
        if ref is not None:
            Kr, Rr, rcolor = ref_frame(cam)
            torch.cuda.synchronize()
            chk["frames"] += 1
            chk["cut"] += int(Kr == K and torch.equal(ri[:K], rri[:K]) and torch.equal(pi[:K], rpi[:K]) and torch.equal(ni[:K], rni[:K]))
            chk["weights"] += int(Kr == K and torch.equal(ts[:K], rts[:K]) and torch.equal(kids[:K], rkids[:K]))
            chk["num_rendered"] += int(Rr == R)
            chk["image"] += int(torch.equal(color, rcolor))
    torch.cuda.synchronize()
    # ---- phases (synchronised after each, first 20 frames of the share), then the job
    n_ph = min(20, len(cams))
    ph_o = {"cut_ms": 0.0, "weights_ms": 0.0, "forward_ms": 0.0}
    for cam in cams[:n_ph]:
        ours_frame(cam, ph_o)
    if os.environ.get("EVAL_DEBUG") and rank == 0:   # per-kernel device times of this library over 10 frames (lodgs.profile)
        import lodgs
        lodgs.profile(True)
        for cam in cams[:10]:
            ours_frame(cam)
        torch.cuda.synchronize()
        prof = lodgs.profile_dump()
        lodgs.profile(False)
        for name, (ms, n) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
            print("   %-28s %4.1f/frame %.4f ms/frame" % (name, n / 10, ms / 10), file=sys.stderr)
    t_ours = job_time(ours_frame)
    out = {"metric": "eval frames/s (cut + weights + forward, 1080p)", "n_gpus": world, "frames": args.frames, "frames_per_rank": len(cams),
           "nodes": N, "gaussians": N, "tau_px": args.tau, "scaling": "strong", "data": "synthetic",
           "K_min_max": [min(Ks), max(Ks)], "num_rendered_min_max": [min(Rs), max(Rs)],
           "ours": {"frames_per_s": round(args.frames / t_ours, 2), "ms_per_frame_per_gpu": round(t_ours * 1e3 / max(1, len(cams)), 4),
                    "mpix_per_s": round(args.frames * W * H / t_ours / 1e6, 1), "phases_ms": {k: round(v / n_ph, 4) for k, v in ph_o.items()}}}
    if ref is not None:
        ph_r = {"cut_ms": 0.0, "weights_ms": 0.0, "forward_ms": 0.0}
        for cam in cams[:n_ph]:
            ref_frame(cam, ph_r)
        t_ref = job_time(ref_frame)
        out["reference"] = {"frames_per_s": round(args.frames / t_ref, 2), "ms_per_frame_per_gpu": round(t_ref * 1e3 / max(1, len(cams)), 4),
                            "mpix_per_s": round(args.frames * W * H / t_ref / 1e6, 1), "phases_ms": {k: round(v / n_ph, 4) for k, v in ph_r.items()}}
        c = torch.tensor([chk[k] for k in ("frames", "cut", "weights", "num_rendered", "image")], dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_reduce(c)
        out["bit_exact_frames"] = dict(zip(("checked", "cut", "weights", "num_rendered", "image"), c.tolist()))
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
