#!/usr/bin/env python3
"""Per-kernel device time of bench.py's train iteration (fwd + bwd + sparse Adam) from the library's own stage timers, next to the
wall-clock per iteration."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "hierarchical-lod-gaussians_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)
import torch  # noqa: E402

import benchlib  # noqa: E402
import lodgs  # noqa: E402
from bench import make_inputs  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
sc, cam, target = make_inputs(0, 1_000_000)
arm = benchlib.OursArm(sc, cam, target, dev)
arm.train_init()
for _ in range(5):
    arm.step_train()
torch.cuda.synchronize()
lodgs.profile(True)
N = 20
t0 = time.perf_counter()
for _ in range(N):
    arm.step_train()
torch.cuda.synchronize()
wall = (time.perf_counter() - t0) / N * 1e3
prof = lodgs.profile_dump()
lodgs.profile(False)
tot = 0.0
for name, (ms, n) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
    print("%-28s %6.1f launches/iter  %.4f ms/iter" % (name, n / N, ms / N))
    tot += ms / N
print("sum of stages %.4f ms/iter, wall %.4f ms/iter" % (tot, wall))
