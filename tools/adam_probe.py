#!/usr/bin/env python3
"""Per-tensor time and achieved bandwidth of lodgs_adam_step (28 B per updated scalar) for the six parameter shapes of workload A,
dense and with an 85 % row list, next to torch's fused elementwise add of the same size for scale."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hierarchical-lod-gaussians_b200"))
import torch  # noqa: E402

from lodgs.optim import Adam  # noqa: E402


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def main():
    P = 1_000_000
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(0)
    rel = (torch.rand(P, device=dev, generator=g) < 0.85).nonzero().reshape(-1)
    for name, shape in (("xyz", (P, 3)), ("f_dc", (P, 1, 3)), ("f_rest", (P, 15, 3)), ("shs", (P, 16, 3)), ("opacity", (P, 1)), ("scaling", (P, 3)),
                        ("rotation", (P, 4))):
        p = torch.randn(*shape, device=dev).requires_grad_(True)
        p.grad = torch.randn(*shape, device=dev)
        opt = Adam([p], lr=1e-6, eps=1e-15)
        w = p.numel() // P
        for label, r in (("dense", None), ("rows85", rel)):
            ms = timeit(lambda: opt.step(r))
            rows = P if r is None else r.numel()
            gb = rows * w * 28 / 1e9
            print("%-9s %-7s width %2d  %.4f ms  %.0f GB/s" % (name, label, w, ms, gb / (ms * 1e-3)))
    a, b = torch.randn(P * 48, device=dev), torch.randn(P * 48, device=dev)
    ms = timeit(lambda: a.add_(b))
    print("torch a.add_(b) 48M floats: %.4f ms  %.0f GB/s (12 B/elt)" % (ms, P * 48 * 12 / 1e9 / (ms * 1e-3)))


if __name__ == "__main__":
    main()
