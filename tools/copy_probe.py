#!/usr/bin/env python3
"""Host-link ceiling of the end-to-end step: copies only, no kernels.

Per rank, the byte counts of bench.py's end-to-end step on workload A (260 883 200 B each way), from / to pinned host memory:
  h2d / d2h alone, both directions at once on two streams, each as (a) the 6 + 6 separate tensors the pipeline moves and (b) ONE packed
  buffer each way. Run alone or under torchrun (one rank per GPU): every rank copies at the same time, the line reports the max over ranks,
  i.e. what the shared host complex delivers with N GPUs pulling on it.

  python tools/copy_probe.py                                                        (1 GPU)
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29544 tools/copy_probe.py
"""
import json
import os

import torch

# floats per Gaussian of the step's inputs (means, opacity, scales, rotations, SH 16x3, + the target image) and of its results
P, W, H = 1_000_000, 1920, 1080
IN_SIZES = [P * 3, P * 1, P * 3, P * 4, P * 48, 3 * H * W]
OUT_SIZES = [3 * H * W, P * 3, P * 48, P * 1, P * 3, P * 4]


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(dev)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    h_in = [torch.empty(n, dtype=torch.float32).pin_memory() for n in IN_SIZES]
    d_in = [torch.empty(n, dtype=torch.float32, device=dev) for n in IN_SIZES]
    h_out = [torch.empty(n, dtype=torch.float32).pin_memory() for n in OUT_SIZES]
    d_out = [torch.empty(n, dtype=torch.float32, device=dev) for n in OUT_SIZES]
    hp_in, dp_in = torch.empty(sum(IN_SIZES), dtype=torch.float32).pin_memory(), torch.empty(sum(IN_SIZES), dtype=torch.float32, device=dev)
    hp_out, dp_out = torch.empty(sum(OUT_SIZES), dtype=torch.float32).pin_memory(), torch.empty(sum(OUT_SIZES), dtype=torch.float32, device=dev)
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()

    def up(packed):
        with torch.cuda.stream(s_in):
            if packed:
                dp_in.copy_(hp_in, non_blocking=True)
            else:
                for d, h in zip(d_in, h_in):
                    d.copy_(h, non_blocking=True)

    def down(packed):
        with torch.cuda.stream(s_out):
            if packed:
                hp_out.copy_(dp_out, non_blocking=True)
            else:
                for h, d in zip(h_out, d_out):
                    h.copy_(d, non_blocking=True)

    def timed(fn, reps=20):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        cur = torch.cuda.current_stream()
        s_in.wait_stream(cur)
        s_out.wait_stream(cur)
        e0.record(cur)
        s_in.wait_event(e0)
        s_out.wait_event(e0)
        for _ in range(reps):
            fn()
        cur.wait_stream(s_in)
        cur.wait_stream(s_out)
        e1.record(cur)
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    b_in, b_out = 4 * sum(IN_SIZES), 4 * sum(OUT_SIZES)
    res = {"n_gpus": world, "h2d_bytes": b_in, "d2h_bytes": b_out, "cases": {}}
    for packed in (False, True):
        tag = "packed" if packed else "6+6 tensors"
        t_up = timed(lambda: up(packed))
        t_dn = timed(lambda: down(packed))
        t_bi = timed(lambda: (up(packed), down(packed)))
        res["cases"][tag] = {"h2d_ms": round(t_up, 4), "h2d_GBps_per_gpu": round(b_in / t_up / 1e6, 1),
                             "d2h_ms": round(t_dn, 4), "d2h_GBps_per_gpu": round(b_out / t_dn / 1e6, 1),
                             "both_ms": round(t_bi, 4), "both_GBps_per_gpu_per_direction": round(b_in / t_bi / 1e6, 1),
                             "both_GBps_aggregate_per_direction": round(world * b_in / t_bi / 1e6, 1)}
    if rank == 0:
        print(json.dumps(res))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
