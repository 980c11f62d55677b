#!/usr/bin/env python3
"""Diagnostic: where does the GPU-built hierarchy differ from the reference creator's on a given synthetic PLY?"""
import os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "hierarchical-lod-gaussians_b200"), ROOT]
import numpy as np, torch
from gaussian_hierarchy import _C as gh
from lodgs import hier_build
from oracle import ply as oply, refshim_hier as rh
n = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
with tempfile.TemporaryDirectory() as td:
    plyf = os.path.join(td, "in.ply")
    pts = oply.synth_points(n, seed=1, extent=(400.0, 400.0, 20.0), scale_mean=0.1)
    oply.write_less_rich_ply(plyf, *pts)
    rh.run_creator(plyf, td + "/ref_")
    h = hier_build.create_dynamic_hierarchy(plyf, td + "/ours_")
    a = gh.load_dynamic_hierarchy(td + "/ref_hierarchy.dhier"); b = gh.load_dynamic_hierarchy(td + "/ours_hierarchy.dhier")
    names = ("pos", "shs", "alpha", "scales", "rot", "nodes")
    for nm, x, y in zip(names, a, b):
        d = (x != y)
        print(nm, "differing rows:", int(d.reshape(d.shape[0], -1).any(1).sum()), "of", x.shape[0])
    nd = (a[5] != b[5]).any(1).nonzero().reshape(-1)
    print("first differing nodes", nd[:10].tolist())
    for i in nd[:4].tolist():
        print(i, "ref", a[5][i].tolist(), "ours", b[5][i].tolist())
    pos = pts[0]
    for ax in range(3):
        u, c = np.unique(pos[:, ax], return_counts=True)
        print("axis", ax, "coordinates shared by >1 point:", int((c > 1).sum()))
