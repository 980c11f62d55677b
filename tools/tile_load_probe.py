#!/usr/bin/env python3
"""How uneven is the per-tile work of the blend kernels on a view, and what would another CTA order buy?
For one frame of the eval path (tools/bench_eval_path.py scene, 2^k leaves) it reads the tile ranges and n_contrib of the forward,
takes work(tile) = number of list entries the tile's CTA walks (max over its pixels of the last contributor's position, i.e. where
the whole tile is saturated, else the list length) and list-schedules the tiles on `slots` CTA slots in three orders: tile id (what
the grid does), longest first, and the lower bound max(sum / slots, longest tile). Prints one JSON line.

  python tools/tile_load_probe.py [--k 24] [--slots 1184]
"""
import argparse
import heapq
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "hierarchical-lod-gaussians_b200"), ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def makespan(work, slots):
    h = [0.0] * slots
    heapq.heapify(h)
    for w in work:
        heapq.heappush(h, heapq.heappop(h) + w)
    return max(h)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--k", type=int, default=24)
    ap.add_argument("--slots", type=int, default=148 * 8)
    ap.add_argument("--frame", type=int, default=0)
    ap.add_argument("--tau", type=float, default=3.0)
    args = ap.parse_args()
    import torch
    import minewrap
    from bench_eval_path import camera_path, W, H
    from bench_hier_step import make_scene
    from gaussian_hierarchy import _C as gh
    from lodgs import synth
    dev = torch.device("cuda", 0)
    nodes, boxes, sc = make_scene(args.k, dev)
    N = nodes.size(0)
    cam = camera_path(1000, dev)[args.frame]
    cam["H"], cam["W"] = H, W
    ri, pi, ni = (torch.empty(N, dtype=torch.int32, device=dev) for _ in range(3))
    ts, kids = torch.empty(N, device=dev), torch.empty(N, dtype=torch.int32, device=dev)
    tau = synth.tau_to_threshold(args.tau, cam["tanfovx"], W)
    K = gh.expand_to_size(nodes, boxes, tau, cam["campos"], torch.zeros(3), ri, pi, ni)
    gh.get_interpolation_weights(ni[:K], tau, nodes, boxes, cam["campos_cpu"], torch.zeros(3), ts, kids)
    sc = dict(sc, sh_degree=3)
    f = minewrap.forward(sc, cam, torch.zeros(3, device=dev), indices=ri[:K], parent_indices=pi[:K], ts=ts[:K], kids=kids[:K], do_depth=False)
    rng = f["ranges"].long()
    length = (rng[:, 1] - rng[:, 0])
    gx, gy = (W + 15) // 16, (H + 15) // 16
    nc = f["n_contrib"].long()
    pad = torch.zeros(gy * 16, gx * 16, dtype=torch.long, device=dev)
    pad[:H, :W] = nc
    walked = pad.view(gy, 16, gx, 16).amax(dim=(1, 3)).reshape(-1)           # last contributor over the tile's pixels
    unsat = (f["final_T"] >= 1e-4)
    padu = torch.zeros(gy * 16, gx * 16, dtype=torch.bool, device=dev)
    padu[:H, :W] = unsat
    any_unsat = padu.view(gy, 16, gx, 16).any(dim=3).any(dim=1).reshape(-1)
    work = torch.where(any_unsat, length, torch.minimum(walked, length)).double().cpu().tolist()  # an unsaturated pixel walks the whole list
    order_lpt = sorted(work, reverse=True)
    s = args.slots
    out = {"K": int(K), "num_rendered": int(f["R"]), "tiles": len(work), "slots": s, "sum_work": sum(work), "max_tile": max(work),
           "mean_tile": sum(work) / len(work), "p99_tile": sorted(work)[int(0.99 * len(work))],
           "list_len_max": int(length.max()), "list_len_sum": int(length.sum()),
           "makespan_tile_order": makespan(work, s), "makespan_longest_first": makespan(order_lpt, s),
           "lower_bound": max(sum(work) / s, max(work))}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
