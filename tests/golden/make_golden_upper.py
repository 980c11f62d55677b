"""tests/golden/upper_tree.npz: outputs of the reference's OWN cut_hierarchy_on_condition / frustum_cull_spheres / extract_frustum_planes
(compiled out of /root/reference/scene/gaussian_model.py by oracle/upper_tree.reference_functions) on the seeded trees and views of
oracle/upper_tree.py, as train_post.py:329-356 calls them. Run in the build container: python tests/golden/make_golden_upper.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import upper_tree as ut  # noqa: E402

CASES = [(300, 1, (3, 4), 1.0, True), (5000, 2, (5, 6, 7), 1.0, True), (5000, 2, (8,), 4.0, True), (2000, 3, (9, 10), 1.0, False), (1, 4, (11,), 1.0, True)]

if __name__ == "__main__":
    fns = ut.reference_functions()
    assert fns is not None
    out = {}
    for leaves, seed, views, mult, cullit in CASES:
        t = {k: torch.from_numpy(v) for k, v in ut.make_upper_tree(leaves, seed).items()}
        for vs in views:
            campos, proj = ut.make_view(vs)
            planes = fns["extract_frustum_planes"](None, torch.from_numpy(proj))
            r = ut.coarse_cut(t["nodes"], t["xyz"], t["min_dist_sq"], torch.from_numpy(campos), t["bounds"] if cullit else None, planes, mult, 0, fns)
            tag = "L%d_s%d_v%d_m%g_c%d_" % (leaves, seed, vs, mult, cullit)
            for k, v in r.items():
                out[tag + k] = v.numpy()
            out[tag + "planes"] = planes.numpy()
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "upper_tree.npz"), **out)
    print("wrote", len(out), "arrays; cut sizes", [out[k].shape[0] for k in out if k.endswith("coarse_cut")])
