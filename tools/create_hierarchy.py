#!/usr/bin/env python3
"""GaussianHierarchyCreator on the GPU: `python tools/create_hierarchy.py <plyfile> <source dir> <output prefix>` — the argument order of
the reference's executable (GH/mainHierarchyCreator.cpp:43-44; the source dir only fed its disabled appearance filter)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "hierarchical-lod-gaussians_b200"), ROOT]

if __name__ == "__main__":
    if len(sys.argv) < 4:
        raise SystemExit("usage: create_hierarchy.py <plyfile> <source dir> <output prefix>")
    import torch
    from lodgs import hier_build
    t = time.perf_counter()
    h = hier_build.create_dynamic_hierarchy(sys.argv[1], sys.argv[3])
    torch.cuda.synchronize()
    print("Kept: %d Gaussians, wrote %d nodes to %shierarchy.dhier in %.3f s" % ((h["pos"].size(0) + 1) // 2, h["pos"].size(0), sys.argv[3],
                                                                                 time.perf_counter() - t))
