#!/usr/bin/env python3
"""Runs ONE bench section alone (no headline workload in front of it), for ncu:  profile_sections.py hier_step|eval_path [small]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "hierarchical-lod-gaussians_b200"), ROOT]
import torch  # noqa: E402

import benchsections as bs  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
which = sys.argv[1]
if which == "hier_step":
    out = bs.hier_step("ours", dev, iters=2, leaves=2_500_000, check=False)
else:
    out = bs.eval_path("ours", 0, 1, dev, frames=12, leaves=25_000_000, check_every=10 ** 9)
print(json.dumps({k: v for k, v in out.items() if not isinstance(v, dict)}))
