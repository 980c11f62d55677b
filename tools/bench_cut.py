#!/usr/bin/env python3
"""LOD-cut ms/frame (BASELINE.json metric, configs B/C): gaussian_hierarchy._C.expand_to_size + get_interpolation_weights on
a synthetic 2^k-leaf hierarchy over a camera path, tau in {0,3,6,15} px, next to the reference CUDA cut (oracle/_ref,
Switching::expandToSize, unmodified) and the CPU restatement (oracle/cut_oracle.c) on the host cores.

  python tools/bench_cut.py [--k 21] [--frames 50]       -> one JSON line
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "hierarchical-lod-gaussians_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def run(k=21, frames=50, cpu_frames=2, dev=None, with_ref=True, with_cpu=True):
    import numpy as np
    import torch
    from gaussian_hierarchy import _C as gh
    from lodgs import synth
    dev = dev or torch.device("cuda", torch.cuda.current_device())
    nodes, boxes = synth.build_hierarchy_torch(k, dev)
    N = nodes.size(0)
    monotone = gh.prepare_hierarchy(nodes, boxes)
    ri, pi, ni = (torch.empty(N, dtype=torch.int32, device=dev) for _ in range(3))
    ts, kids = torch.empty(N, device=dev), torch.empty(N, dtype=torch.int32, device=dev)
    # camera path: a fly-over at 30 m height across the slab
    ang = np.linspace(0, 2 * np.pi, frames, endpoint=False)
    path = np.stack([120 * np.cos(ang), 120 * np.sin(ang), np.full(frames, 30.0)], 1).astype(np.float32)
    vps = [torch.from_numpy(p).to(dev) for p in path]
    out = {"nodes": N, "frames": frames, "monotone": bool(monotone), "taus": {}}
    ref = None
    if with_ref:
        from oracle import refshim
        if refshim.available():
            ref = refshim.lib()
            ref.ref_expand_to_size.restype = C.c_int
            ref.ref_expand_to_size.argtypes = [C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float,
                                               C.c_void_p, C.c_void_p, C.c_void_p]
            rri, rpi, rni = (torch.empty(N, dtype=torch.int32, device=dev) for _ in range(3))
    for tau_px in (0.0, 3.0, 6.0, 15.0):
        tau = synth.tau_to_threshold(tau_px, 0.8, 1920)
        ent = {}
        for mode, name in ((0, "flat"), (1, "traversal")):
            if mode == 1 and not monotone:
                continue
            for _ in range(3):
                K = gh.expand_to_size(nodes, boxes, tau, vps[0], torch.zeros(3), ri, pi, ni, mode=mode)
            torch.cuda.synchronize()
            def timed(with_weights):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                tot = 0
                for f in range(frames):
                    K = gh.expand_to_size(nodes, boxes, tau, vps[f], torch.zeros(3), ri, pi, ni, mode=mode)
                    if with_weights:
                        gh.get_interpolation_weights(ni[:K], tau, nodes, boxes, torch.from_numpy(path[f]), torch.zeros(3), ts, kids)
                    tot += K
                e1.record()
                torch.cuda.synchronize()
                return e0.elapsed_time(e1) / frames, tot
            ms_cut, Ks = timed(False)
            ms_all, _ = timed(True)
            ent[name] = {"cut_ms_per_frame": round(ms_cut, 4), "cut_plus_weights_ms_per_frame": round(ms_all, 4), "mean_K": Ks // frames,
                         "flat_equiv_GBps": round((60.0 * N + 12.0 * Ks / frames) / (ms_cut * 1e-3) / 1e9, 1)}
        if ref is not None:
            for _ in range(2):
                ref.ref_expand_to_size(N, tau, nodes.data_ptr(), boxes.data_ptr(), vps[0].data_ptr(), 0.0, 0.0, 1.0, rri.data_ptr(),
                                       rpi.data_ptr(), rni.data_ptr())
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            same = True
            for f in range(frames):
                Kr = ref.ref_expand_to_size(N, tau, nodes.data_ptr(), boxes.data_ptr(), vps[f].data_ptr(), 0.0, 0.0, 1.0, rri.data_ptr(),
                                            rpi.data_ptr(), rni.data_ptr())
            torch.cuda.synchronize()
            ent["reference_cuda"] = {"ms_per_frame": round((time.perf_counter() - t0) * 1e3 / frames, 4), "K_last": int(Kr)}
            # bit-exactness of the last frame against the reference
            K = gh.expand_to_size(nodes, boxes, tau, vps[frames - 1], torch.zeros(3), ri, pi, ni, mode=0)
            same = (K == Kr) and torch.equal(ri[:K], rri[:K]) and torch.equal(pi[:K], rpi[:K]) and torch.equal(ni[:K], rni[:K])
            if monotone:
                K2 = gh.expand_to_size(nodes, boxes, tau, vps[frames - 1], torch.zeros(3), ri, pi, ni, mode=1)
                same = same and (K2 == Kr) and torch.equal(ri[:K2], rri[:K2]) and torch.equal(ni[:K2], rni[:K2])
            ent["bit_exact_vs_reference"] = bool(same)
        out["taus"]["%g" % tau_px] = ent
    if with_cpu:
        from oracle import oracle as orc
        nn, bb = nodes.cpu().numpy(), boxes.cpu().numpy()
        t0 = time.perf_counter()
        for f in range(cpu_frames):
            orc.expand_to_size(nn, bb, synth.tau_to_threshold(3.0, 0.8, 1920), path[f])
        out["cpu_port_tau3"] = {"ms_per_frame": round((time.perf_counter() - t0) * 1e3 / cpu_frames, 2), "frames": cpu_frames,
                                "cores": 1, "kind": "port (oracle/cut_oracle.c, sequential scan+scatter)"}
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--k", type=int, default=21)
    ap.add_argument("--frames", type=int, default=50)
    a = ap.parse_args()
    print(json.dumps(run(a.k, a.frames)))
