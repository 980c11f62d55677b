#!/usr/bin/env python3
"""Timeline of bench.py's end-to-end step (H2D / kernels / D2H streams) for both arms: per step the start and end of each phase
relative to the first step's H2D, from CUDA events. Shows whether the copies of neighbouring steps overlap the kernels."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "hierarchical-lod-gaussians_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)
import torch  # noqa: E402

import benchlib  # noqa: E402
from bench import make_inputs  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    sc, cam, target = make_inputs(0, 1_000_000)
    for name, cls in (("ours", benchlib.OursArm), ("reference", benchlib.RefArm)):
        try:
            arm = cls(sc, cam, target, dev)
        except FileNotFoundError:
            continue
        arm.ev_timing = True
        for _ in range(4):
            arm.step_e2e()
        torch.cuda.synchronize()
        marks = []
        orig = arm.fwd_bwd

        def wrapped(d, t, _orig=orig, _marks=marks):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            out = _orig(d, t)
            b.record()
            _marks.append((a, b))
            return out
        arm.fwd_bwd = wrapped
        t0 = torch.cuda.Event(enable_timing=True)
        t0.record()
        ins, outs = [], []
        for _ in range(6):
            arm.step_e2e()
            pp = arm.pipe
            slot = (pp["i"] - 1) & 1
            ins.append(pp["ev_in"][slot])
            outs.append(pp["ev_out"][slot])
        torch.cuda.synchronize()
        print(name)
        for i, ((a, b), ei, eo) in enumerate(zip(marks, ins, outs)):
            def t(e):
                try:
                    return "%7.2f" % t0.elapsed_time(e)
                except Exception:  # noqa: BLE001  (events recorded without timing)
                    return "    n/a"
            print("  step %d: kernels %s .. %s   h2d done %s   d2h done %s" % (i, t(a), t(b), t(ei), t(eo)))


if __name__ == "__main__":
    main()
