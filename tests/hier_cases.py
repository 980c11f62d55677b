"""Seeded hierarchy-file contents shared by tests/golden/make_golden_hier.py and tests/test_hier_files_vs_reference.py."""
import numpy as np


def static_case(P=23, N=15, seed=5):
    """Random (not tree-consistent: the file code does not look at the values) .hier contents incl. half rounding edge cases."""
    r = np.random.default_rng(seed)
    h = {"pos": r.normal(0, 30, (P, 3)), "shs": r.normal(0, 1, (P, 48)), "alpha": r.random(P), "scales": r.normal(-3, 1, (P, 3)),
         "rot": r.normal(0, 1, (P, 4)), "boxes": r.normal(0, 50, (N, 8))}
    h = {k: v.astype(np.float32) for k, v in h.items()}
    if P and N:
        h["shs"][0, :6] = [65519.0, 1e-8, 6.1e-5, -70000.0, 2049.0, 2051.0]     # overflow edge, flush, subnormal, -inf, ties
        h["boxes"][0, :3] = [5.9604645e-8, 2.98e-8, -0.0]
    # HalfNode keeps depth / counts as int16 and the writer throws above 32000 (hierarchy_writer.cpp:97-98)
    nodes = np.stack([r.integers(0, 32000, N), r.integers(-1, 10 ** 6, N), r.integers(0, 10 ** 6, N), r.integers(0, 2, N), r.integers(0, 2, N),
                      r.integers(-1, 10 ** 6, N), r.integers(0, 3, N)], 1).astype(np.int32)
    h["nodes"] = nodes
    return h


def dynamic_case(deg, P=19, seed=9):
    r = np.random.default_rng(seed + deg)
    M3 = 3 * (deg + 1) ** 2
    h = {"pos": r.normal(0, 30, (P, 3)), "shs": r.normal(0, 1, (P, M3)), "alpha": r.random(P), "scales": r.normal(-3, 1, (P, 3)),
         "rot": r.normal(0, 1, (P, 4))}
    h = {k: v.astype(np.float32) for k, v in h.items()}
    h["nodes"] = r.integers(-1, 10 ** 6, (P, 6)).astype(np.int32)
    return h
