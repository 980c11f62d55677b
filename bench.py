#!/usr/bin/env python3
"""bench.py — fwd+bwd rasterization throughput of the hot path (BASELINE.json configs[1]).

Workload (config.workload = "A"): 1 M synthetic Gaussians (SH degree 3), one 1920x1080 view per GPU, forward +
backward through the reference-facing plugin surface (diff_gaussian_rasterization._C over the C ABI).
A step = one forward + one backward of one view. With N GPUs every rank renders its own view of equal expected cost
(see make_inputs; no data-path collective: the path shards by view), so scaling is weak and value = N * W*H * K / t.
The same JSON line carries the other BASELINE configurations as sections (benchsections.py): `hier_step` (configs[2], 4 999 999
nodes, N = 1), `eval_path` (configs[3], 49 999 999 nodes, 1000-pose camera path sharded over the ranks), `dp_train` (configs[4],
gradient exchange over NCCL), `creator` (configs[0], GPU builder next to the reference's CPU creator + CPU cut tau sweep), `train`.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Prints ONE JSON line on rank 0 (see DESIGN.md "Measurement").
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "hierarchical-lod-gaussians_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

W, H, P, DEG, M = 1920, 1080, 1_000_000, 3, 16


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--P", type=int, default=P, help="number of Gaussians (default: the 1 M of config A)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sections", default="hier_step,eval_path,dp_train,creator",
                    help="comma list of the extra BASELINE configurations to run after the headline workload ('' = none)")
    ap.add_argument("--eval-frames", type=int, default=1000, help="camera-path length of the eval_path section")
    ap.add_argument("--eval-leaves", type=int, default=25_000_000, help="leaves of the eval_path hierarchy (N = 2L-1 nodes)")
    ap.add_argument("--hier-leaves", type=int, default=2_500_000, help="leaves of the hier_step hierarchy")
    ap.add_argument("--dp-mode", default="auto", choices=["auto", "factored", "allreduce"], help="gradient exchange of the dp_train section")
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU while the timed region runs (nvidia-smi, 100 ms period)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.proc = index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                f = [x.strip() for x in line.split(",")]
                if len(f) >= 7:
                    self.samples.append(f)
        except Exception:  # noqa: BLE001
            pass

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        sm = sorted(int(float(s[0])) for s in self.samples if s[0].replace(".", "").isdigit())
        mx = [int(float(s[1])) for s in self.samples if s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for s in self.samples for n, v in zip(names, s[3:7]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.samples)}


def make_inputs(rank, n_gauss):
    """Workload A on the host (numpy) for this rank. Weak scaling needs the same work on every GPU, and the cost of a view depends
    on where the camera stands in the scene, so rank r does not get another pose of rank 0's scene (measured: 13 % more instances
    for a camera 0.6 m to the side, which the max over ranks then reports as a scaling loss); it gets the same camera and a scene
    drawn from the same distribution with seed r, i.e. N views of equal expected cost. Rank 0 is the N = 1 workload."""
    import numpy as np
    from lodgs import synth
    cam = synth.make_camera(W=W, H=H, fx=1200.0, fy=1200.0)
    sc = synth.make_gaussians(n_gauss, seed=rank, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"])
    rng = np.random.default_rng(1234 + rank)
    target = rng.random((3, H, W), dtype=np.float32)
    return sc, cam, target


def main():
    args = parse()
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    from benchlib import run_ours, run_reference
    if args.impl == "reference":
        line = run_reference(args, rank, world, dev)
    else:
        line = run_ours(args, rank, world, dev)
    if rank == 0 and line is not None:
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
