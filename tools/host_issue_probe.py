#!/usr/bin/env python3
"""Host-side issue time of one end-to-end step (bench.py's pipeline) and of one resident step: how much CPU time per step the Python
plugin path needs, i.e. how slow a host may get before the GPU-side figures become host-bound."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "hierarchical-lod-gaussians_b200"), ROOT]
import torch  # noqa: E402

import benchlib  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
sc, cam, target = benchlib.make_inputs(0, 1_000_000)
arm = benchlib.OursArm(sc, cam, target, dev)
out = {}
for name, fn in (("resident", arm.step_resident), ("e2e", arm.step_e2e)):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    n = 30
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    t1 = time.perf_counter()          # everything is enqueued (the step blocks only on the 4-byte instance count)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    out[name] = {"host_issue_ms_per_step": round((t1 - t0) * 1e3 / n, 3), "wall_ms_per_step": round((t2 - t0) * 1e3 / n, 3)}
print(json.dumps(out))
