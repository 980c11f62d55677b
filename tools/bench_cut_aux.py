#!/usr/bin/env python3
"""ms/call of the two cut ops SURVEY §8f adds — the dynamic-format cut (expand_to_size_dynamic + get_interpolation_weights_dynamic)
and the SPT cut (get_spt_cut_cuda) — next to the reference's own CUDA implementations (oracle/_ref, unmodified) on the same
inputs, results compared bit for bit. Both sides end every call with the count on the host, so wall clock around a
synchronised loop is the honest per-call time.

  python tools/bench_cut_aux.py [--leaves 200000] [--tables 200000] [--calls 20]   -> one JSON line
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


def _time(fn, calls):
    import torch
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(calls):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3 / calls


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--leaves", type=int, default=200000)
    ap.add_argument("--tables", type=int, default=200000)
    ap.add_argument("--calls", type=int, default=20)
    args = ap.parse_args()
    import numpy as np
    import torch
    from gaussian_hierarchy import _C as gh
    from lodgs import synth
    from oracle import refshim
    dev = torch.device("cuda", 0)
    ref = refshim.lib() if refshim.available() else None
    out = {}

    # ---- dynamic-format cut
    rng = np.random.default_rng(3)
    pts = (rng.uniform(-1, 1, (args.leaves, 3)) * np.array([200.0, 200.0, 10.0])).astype(np.float32)
    rad = np.exp(rng.normal(np.log(0.1), 0.5, args.leaves)).astype(np.float32)
    nodes, boxes, gpos, grad = synth.build_hierarchy(pts, rad)
    dyn, pos, sca = synth.to_dynamic_hierarchy(nodes, boxes, gpos, grad, seed=3, skybox=100)
    n_t, p_t, s_t = (torch.from_numpy(a).to(dev) for a in (dyn, pos, sca))
    N = dyn.shape[0]
    vp = torch.tensor([120.0, 0.0, 30.0], device=dev)
    z = torch.tensor([-0.97, 0.0, -0.24])
    z = z / z.norm()
    ri, pi, ni = (torch.full((N,), -7, dtype=torch.int32, device=dev) for _ in range(3))
    ts, kids = torch.empty(N, device=dev), torch.empty(N, dtype=torch.int32, device=dev)
    dres = {"nodes": N, "taus": {}}
    for tau in (0.002, 0.01, 0.05):
        def mine():
            K = gh.expand_to_size_dynamic(n_t, p_t, s_t, tau, vp, z, ri, pi, ni)
            gh.get_interpolation_weights_dynamic(ni[:K], tau, n_t, p_t, s_t, vp, z, ts, kids)
            return K
        K = mine()
        ent = {"K": int(K), "ours_ms": round(_time(mine, args.calls), 4)}
        if ref is not None:
            rri, rpi, rni = (torch.full((N,), -7, dtype=torch.int32, device=dev) for _ in range(3))
            rts, rkids = torch.empty(N, device=dev), torch.empty(N, dtype=torch.int32, device=dev)
            vh = vp.cpu()

            def theirs():
                Kr = ref.ref_expand_to_size_dynamic(N, float(tau), n_t.data_ptr(), p_t.data_ptr(), s_t.data_ptr(), vp.data_ptr(), float(z[0]),
                                                    float(z[1]), float(z[2]), rri.data_ptr(), rpi.data_ptr(), rni.data_ptr())
                if Kr:
                    ref.ref_get_ts_indexed_dynamic(Kr, rni.data_ptr(), float(tau), n_t.data_ptr(), p_t.data_ptr(), s_t.data_ptr(), float(vh[0]),
                                                   float(vh[1]), float(vh[2]), float(z[0]), float(z[1]), float(z[2]), rts.data_ptr(),
                                                   rkids.data_ptr(), None)
                return Kr
            Kr = theirs()
            torch.cuda.synchronize()
            ent["bit_exact"] = bool(Kr == K and torch.equal(ri[:K], rri[:K]) and torch.equal(ni[:K], rni[:K]) and torch.equal(pi[:K], rpi[:K])
                                    and torch.equal(ts[:K], rts[:K]) and torch.equal(kids[:K], rkids[:K]))
            ent["reference_ms"] = round(_time(theirs, args.calls), 4)
        dres["taus"][str(tau)] = ent
    out["dynamic_cut"] = dres

    # ---- SPT cut
    rng = np.random.default_rng(4)
    T = args.tables
    lens = rng.integers(1, 129, T)
    starts = np.zeros(T + 1, np.int32)
    starts[1:] = np.cumsum(lens)
    total = int(starts[-1])
    # per-run descending max distances: sort (run id, -value) pairs
    run = np.repeat(np.arange(T), lens)
    mx = rng.uniform(0.5, 60.0, total).astype(np.float32)
    order = np.lexsort((-mx, run))
    mx = mx[order]
    mx[starts[:-1]] = 1e12
    mn = np.minimum(mx, rng.uniform(0.0, 40.0, total).astype(np.float32))
    ids = rng.integers(1, 50_000_000, total).astype(np.int32)
    pad = lambda a, v: np.concatenate([a, np.array([v], a.dtype)])  # noqa: E731
    t_ids, t_st, t_mx, t_mn = (torch.from_numpy(a).to(dev) for a in (pad(ids, 0), starts, pad(mx, 0), pad(mn, 0)))
    sres = {"tables": T, "entries": total, "selections": {}}
    for share in (0.05, 0.5):
        S = int(T * share)
        sel = torch.from_numpy(rng.permutation(T)[:S].astype(np.int32)).to(dev)
        dist = torch.from_numpy(rng.uniform(0.2, 70.0, S).astype(np.float32)).to(dev)

        def mine():
            return gh.get_spt_cut_cuda(S, t_ids, t_st, t_mx, t_mn, sel, dist)
        cut, counts = mine()
        ent = {"slots": S, "K": int(cut.numel()), "ours_ms": round(_time(mine, args.calls), 4)}
        if ref is not None:
            def theirs():
                return refshim.get_spt_cut(t_ids, t_st, t_mx, t_mn, sel, dist)
            rcut, rcounts = theirs()
            ent["bit_exact"] = bool(torch.equal(cut, rcut) and torch.equal(counts, rcounts))
            ent["reference_ms"] = round(_time(theirs, args.calls), 4)
        sres["selections"][str(share)] = ent
    out["spt_cut"] = sres
    print(json.dumps(out))


if __name__ == "__main__":
    main()
