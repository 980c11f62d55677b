#!/usr/bin/env python3
"""Where do ours and the reference differ on the hierarchical frame of tools/bench_hier_step.py? (diagnostic)"""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "hierarchical-lod-gaussians_b200"), ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)
import numpy as np
import torch
from gaussian_hierarchy import _C as gh
from lodgs import synth
from oracle import refshim
import bench_hier_step as B
import minewrap

k = int(sys.argv[1]) if len(sys.argv) > 1 else 18
dev = torch.device("cuda", 0)
nodes, boxes, sc = B.make_scene(k, dev)
sc["sh_degree"] = 3
N = nodes.size(0)
cam_np = synth.look_at_camera(eye=(-150.0, 20.0, 18.0), target=(0.0, 0.0, 0.0), up=(0.0, 0.0, 1.0), W=1920, H=1080, fx=1200.0, fy=1200.0)
cam = minewrap.to_cuda(cam_np)
bg = torch.zeros(3, device=dev)
ri, pi, ni = (torch.empty(N, dtype=torch.int32, device=dev) for _ in range(3))
ts, kids = torch.empty(N, device=dev), torch.empty(N, dtype=torch.int32, device=dev)
for tau_px in (3.0, 15.0):
    tau = synth.tau_to_threshold(tau_px, cam["tanfovx"], 1920)
    K = gh.expand_to_size(nodes, boxes, tau, cam["campos"], torch.zeros(3), ri, pi, ni)
    gh.get_interpolation_weights(ni[:K], tau, nodes, boxes, cam["campos"].cpu(), torch.zeros(3), ts, kids)
    kw = {"indices": ri[:K].clone(), "parent_indices": pi[:K].clone(), "ts": ts[:K].clone(), "kids": kids[:K].clone()}
    m = minewrap.forward(sc, cam, bg, **kw)
    r = refshim.forward(sc, cam, bg, **kw)
    print("tau", tau_px, "K", K, "R mine", m["R"], "ref", r["R"])
    bad = (m["radii"] != r["radii"]).nonzero().reshape(-1)
    print("  radii mismatches", bad.numel(), "of", K, " tiles_touched mismatches", int((m["tiles_touched"] != r["tiles_touched"]).sum()))
    for j in bad[:8].tolist():
        g, p = int(ri[j]), int(pi[j])
        print("   j", j, "gauss", g, "parent", p, "t", float(ts[j]), "kids", int(kids[j]), "radii mine/ref", int(m["radii"][j]), int(r["radii"][j]),
              "scale", sc["scales"][g].tolist(), "pscale", sc["scales"][p].tolist() if p >= 0 else None,
              "depth mine/ref", float(m["depths"][j]), float(r["depths"][j]), "tiles", int(m["tiles_touched"][j]), int(r["tiles_touched"][j]),
              "rect mine", m["rects"][j].tolist(), "ref", r["rects"][j].tolist())
    if m["R"] == r["R"]:
        print("  image equal", bool(torch.equal(m["color"], r["color"])))
