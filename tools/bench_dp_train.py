#!/usr/bin/env python3
"""BASELINE.json configs[4]: multi-view training of ONE chunk, views split across the GPUs, through the alt-rasterizer surface
(alt_gaussian_rasterization: dc + shs, antialiasing on — configs/kbuffer.json) — per step and per rank: forward + backward of the
rank's own 1920x1080 view of the replicated 1 M-Gaussian scene, NCCL sum all-reduce of the six parameter gradients in place
(lodgs.parallel.allreduce_gradients_async, 59 floats per Gaussian = 236 MB; each parameter's Adam update waits only for its own
gradient, so updates overlap the rest of the exchange; --no-overlap finishes the exchange first), visibility = dL_dopacity != 0 on the reduced gradient
(train_single.py:171-172), SparseGaussianAdam.step(visibility, N) on identical sums, so the replicas stay identical (checked at the end
with an all-reduce of the parameter checksums' min and max).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29541 tools/bench_dp_train.py
  python tools/bench_dp_train.py            (one GPU: no exchange)

Timing: CUDA events around K steps after W warm-up steps, barrier + synchronize on both sides, MAX over ranks. `--reference` times the
reference's alt-rasterizer (oracle/_ref/libref_ar.so, unmodified kernels incl. its adamUpdateCUDA) in the same loop with the same
all-reduce. One JSON line on rank 0.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "hierarchical-lod-gaussians_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

W, H, DEG = 1920, 1080, 3
# train_single.py / arguments/__init__.py learning rates x 1e-3: the scene (and with it the timed work) does not drift
LRS = {"means3D": 1.6e-7, "dc": 2.5e-6, "shs_rest": 1.25e-7, "opacities": 5e-5, "scales": 5e-6, "rotations": 1e-6}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--P", type=int, default=1_000_000)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--reference", action="store_true")
    ap.add_argument("--no-overlap", action="store_true", help="finish the whole all-reduce before the optimizer starts")
    ap.add_argument("--coalesce", action="store_true", help="with --no-overlap: the six in-place all-reduces inside one NCCL group")
    ap.add_argument("--flat-bucket", action="store_true", help="with --no-overlap: pack all gradients into one bucket, one NCCL call")
    args = ap.parse_args()
    import numpy as np
    import torch
    import torch.distributed as dist
    from alt_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, SparseGaussianAdam
    from lodgs import parallel, synth
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    # the scene is replicated (same seed on every rank); the view is the rank's own: cameras 0.25 m apart on a line
    cam_np = synth.make_camera(W=W, H=H, fx=1200.0, fy=1200.0, T=(0.25 * (rank - 0.5 * (world - 1)), 0.0, 0.0))
    scn = synth.make_gaussians(args.P, seed=0, tanfovx=cam_np["tanfovx"], tanfovy=cam_np["tanfovy"])
    cam = {k: (torch.from_numpy(np.ascontiguousarray(v, dtype=np.float32)).to(dev) if isinstance(v, np.ndarray) else v) for k, v in cam_np.items()}
    shs = torch.from_numpy(scn["shs"]).to(dev)
    par = {"means3D": torch.from_numpy(scn["means3D"]).to(dev), "dc": shs[:, :1].contiguous(), "shs_rest": shs[:, 1:].contiguous(),
           "opacities": torch.from_numpy(scn["opacities"]).to(dev), "scales": torch.from_numpy(scn["scales"]).to(dev),
           "rotations": torch.from_numpy(scn["rotations"]).to(dev)}
    target = torch.rand(3, H, W, device=dev, generator=torch.Generator(device=dev).manual_seed(100 + rank))
    bg = torch.zeros(3, device=dev)
    names = list(LRS)
    # all-reduce issue order = optimizer group order: dL_dopacity first (the visibility mask needs it), the largest tensor next, so the
    # remaining 56 MB travel while Adam runs over the 180 MB of shs_rest
    order = ["opacities", "shs_rest", "means3D", "dc", "scales", "rotations"]
    N = args.P
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    phase = {"fwd_bwd_ms": 0.0, "allreduce_ms": 0.0, "adam_ms": 0.0}

    if not args.reference:
        par = {k: torch.nn.Parameter(v) for k, v in par.items()}
        opt = SparseGaussianAdam([{"params": [par[k]], "lr": LRS[k], "name": k} for k in order], lr=0.0, eps=1e-15)
        rs = GaussianRasterizationSettings(H, W, cam["tanfovx"], cam["tanfovy"], bg, 1.0, cam["viewmatrix"], cam["projmatrix"], DEG, cam["campos"],
                                           False, False, True)
        rast = GaussianRasterizer(rs)
        means2D = torch.zeros(N, 3, device=dev, requires_grad=True)

        def step(timed):
            e = [ev() for _ in range(4)] if timed else None
            if timed:
                e[0].record()
            color, radii, invd = rast(par["means3D"], means2D, par["opacities"], dc=par["dc"], shs=par["shs_rest"], scales=par["scales"],
                                      rotations=par["rotations"])
            loss = ((color - target) ** 2).mean()
            loss.backward()
            if timed:
                e[1].record()
            if args.no_overlap:
                nbytes = parallel.allreduce_gradients([par[k].grad for k in names], inplace_min_bytes=(1 << 62) if args.flat_bucket else parallel.INPLACE_MIN_BYTES,
                                                        coalesce=args.coalesce)
                works = {}
            else:
                handles, nbytes = parallel.allreduce_gradients_async([par[k].grad for k in order])
                works = {par[k]: h for k, h in zip(order, handles)}
                if handles[0] is not None:
                    handles[0].wait()
            if timed:
                e[2].record()   # overlapped mode: only dL_dopacity has arrived here; the rest of the exchange is inside adam_ms
            vis = (par["opacities"].grad != 0).reshape(-1)
            opt.step_overlapped(vis, N, works)
            opt.zero_grad(set_to_none=True)
            means2D.grad = None
            if timed:
                e[3].record()
            return e, nbytes
    else:
        from oracle import refshim_ar as rar
        if not rar.available():
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libref_ar.so not built"}))
            return
        L = rar.lib()
        sc = dict(par, sh_degree=DEG)
        mom = {k: (torch.zeros_like(par[k]), torch.zeros_like(par[k])) for k in names}
        gname = {"means3D": "dL_dmean3D", "dc": "dL_ddc", "shs_rest": "dL_dsh", "opacities": "dL_dopacity", "scales": "dL_dscale", "rotations": "dL_drot"}

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
                L.refar_adam(rar.ptr(par[k]), rar.ptr(g[gname[k]]), rar.ptr(mom[k][0]), rar.ptr(mom[k][1]), rar.ptr(vis), LRS[k], 0.9, 0.999, 1e-15,
                             N, par[k].numel() // N)
            if timed:
                e[3].record()
            return e, nbytes

    for _ in range(max(3, args.warmup)):
        step(False)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0, t1 = ev(), ev()
    t0.record()
    marks = []
    for _ in range(args.steps):
        e, nbytes = step(True)
        marks.append(e)
    t1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    for e in marks:
        phase["fwd_bwd_ms"] += e[0].elapsed_time(e[1]); phase["allreduce_ms"] += e[1].elapsed_time(e[2]); phase["adam_ms"] += e[2].elapsed_time(e[3])
    t = torch.tensor([t0.elapsed_time(t1)] + [phase[k] for k in phase], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    # replicas identical? min and max over ranks of each parameter's checksum must agree
    cs = torch.stack([par[k].detach().double().sum() for k in names])
    lo, hi = cs.clone(), cs.clone()
    if world > 1:
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    ms = float(t[0]) / args.steps
    out = {"metric": "multi-view chunk training, views/s (fwd + bwd + grad all-reduce + sparse Adam, 1080p, alt-rasterizer surface)",
           "impl": "reference" if args.reference else "ours", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
           "gaussians": N, "ms_per_step": round(ms, 4), "views_per_s": round(world * 1e3 / ms, 2),
           "mpix_per_s": round(world * W * H / ms / 1e3, 1), "scaling": "weak", "data": "synthetic",
           "phases_ms_max_over_ranks": {k: round(float(v) / args.steps, 4) for k, v in zip(phase, t[1:])},
           "allreduce_bytes": int(nbytes), "allreduce_bus_GBps": (round(2 * (world - 1) / world * nbytes / (float(t[2]) / args.steps * 1e-3) / 1e9, 1)
                                                                  if world > 1 and float(t[2]) > 0 and (args.reference or args.no_overlap) else None),
           "overlap": bool(not args.reference and not args.no_overlap), "flat_bucket": bool(args.flat_bucket), "coalesce": bool(args.coalesce),
           "replicas_identical": bool(torch.equal(lo, hi))}
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
