"""Seeded test cases shared by the golden generator, the CPU oracle tests and the GPU parity tests."""
import numpy as np

from lodgs import synth

CASES = {
    # name: (P, W, H, fx, seed, lod, zmax, scale_mean)
    "small": (600, 160, 96, 100.0, 3, False, 8.0, 0.05),
    "wide": (3000, 320, 176, 180.0, 5, False, 20.0, 0.04),
    "ragged": (900, 150, 83, 90.0, 7, False, 10.0, 0.06),     # image size not a multiple of 16
    "lod": (1200, 192, 112, 120.0, 11, True, 10.0, 0.05),
}


def make_case(name):
    P, W, H, fx, seed, lod, zmax, sm = CASES[name]
    cam = synth.make_camera(W=W, H=H, fx=fx, fy=fx)
    sc = synth.make_gaussians(P, seed=seed, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], zmax=zmax, scale_mean=sm)
    bg = np.array([0.1, 0.2, 0.3], np.float32)
    kw = {}
    if lod:
        rng = np.random.default_rng(seed + 100)
        K = P // 2
        idx = np.sort(rng.permutation(P)[:K]).astype(np.int32)
        par = rng.integers(0, P, K).astype(np.int32)
        par[rng.random(K) < 0.2] = -1
        kw = {"indices": idx, "parent_indices": par, "ts": rng.random(K).astype(np.float32),
              "kids": rng.integers(1, 5, K).astype(np.int32)}
    rng = np.random.default_rng(seed + 200)
    dL = (rng.random((3, H, W)).astype(np.float32) - 0.5) / (H * W)
    dLi = (rng.random((1, H, W)).astype(np.float32) - 0.5) / (H * W)
    return sc, cam, bg, kw, dL, dLi


def make_cut_case(L=2000, seed=0):
    rng = np.random.default_rng(seed)
    pts = (rng.uniform(-1, 1, (L, 3)) * np.array([40.0, 40.0, 4.0])).astype(np.float32)
    rad = np.exp(rng.normal(np.log(0.05), 0.5, L)).astype(np.float32)
    nodes, boxes, gpos, grad = synth.build_hierarchy(pts, rad)
    views = [np.array(v, np.float32) for v in ([0.0, 0.0, -30.0], [10.0, -5.0, 1.0], [80.0, 80.0, 20.0], [0.5, 0.25, 0.1])]
    return nodes, boxes, views


DYN_VIEWS = [([0.0, 0.0, -60.0], [0.0, 0.0, 1.0]), ([30.0, -20.0, 25.0], [-0.6, 0.4, -0.69]), ([5.0, 5.0, 1.0], [1.0, 0.0, 0.0]),
             ([-80.0, 10.0, -5.0], [0.99, -0.1, 0.05])]
DYN_TAUS = (0.0, 0.002, 0.01, 0.05, 0.5)


def make_dynamic_cut_case(L=2000, seed=21, skybox=4):
    """Dynamic-format hierarchy (nodes[N,6], positions, scales) + (viewpoint, zdir) pairs for the dynamic cut tests."""
    rng = np.random.default_rng(seed)
    pts = (rng.uniform(-1, 1, (L, 3)) * np.array([40.0, 40.0, 4.0])).astype(np.float32)
    rad = np.exp(rng.normal(np.log(0.05), 0.5, L)).astype(np.float32)
    nodes, boxes, gpos, grad = synth.build_hierarchy(pts, rad)
    dyn, pos, sca = synth.to_dynamic_hierarchy(nodes, boxes, gpos, grad, seed=seed, skybox=skybox)
    views = [(np.array(v, np.float32), (np.array(z, np.float32) / np.linalg.norm(z)).astype(np.float32)) for v, z in DYN_VIEWS]
    return dyn, pos, sca, views


SPT_CASES = {  # name -> (seed, tables, longest table, selected slots, share of zero ids, empty tables allowed)
    "small": (5, 40, 12, 17, 0.05, False),
    "one_slot": (6, 9, 30, 1, 0.0, False),
    "ragged": (7, 600, 300, 411, 0.02, True),
    "all_tables": (8, 250, 64, 250, 0.1, True),
    "large": (9, 20000, 128, 9000, 0.01, False),
}


def make_spt_case(name):
    """SPT tables as scene/gaussian_model.py:200-258 lays them out (runs sorted by max distance descending, the run's first entry
    carrying max = 1e12) + a selection (table ids, distances) like train_post.py:398's. The entry arrays carry one padding entry
    past the last run: the reference probes one entry past a fully selected run."""
    seed, n_tables, longest, n_sel, zero_share, allow_empty = SPT_CASES[name]
    rng = np.random.default_rng(seed)
    lens = rng.integers(0 if allow_empty else 1, longest + 1, n_tables)
    starts = np.zeros(n_tables + 1, np.int32)
    starts[1:] = np.cumsum(lens)
    total = int(starts[-1])
    spt_max = np.zeros(total + 1, np.float32)
    spt_min = np.zeros(total + 1, np.float32)
    for t in range(n_tables):
        n = int(lens[t])
        if n == 0:
            continue
        mx = np.sort(rng.uniform(0.5, 60.0, n).astype(np.float32))[::-1].copy()
        mx[0] = 1e12
        if n > 3:
            mx[2] = mx[1]  # ties
        spt_max[starts[t]:starts[t + 1]] = mx
        spt_min[starts[t]:starts[t + 1]] = np.minimum(mx, rng.uniform(0.0, 40.0, n).astype(np.float32))
    ids = rng.integers(1, 5_000_000, total + 1).astype(np.int32)
    ids[rng.random(total + 1) < zero_share] = 0
    sel = rng.permutation(n_tables)[:n_sel].astype(np.int32)
    dist = rng.uniform(0.2, 70.0, n_sel).astype(np.float32)
    if n_sel > 4:
        dist[1] = 1e13   # beyond every max distance, the root's included: nothing selected
        dist[2] = 0.0    # closer than everything: the whole run
    return ids, starts, spt_max, spt_min, sel, dist


def spt_cut_model(ids, starts, spt_max, spt_min, sel, dist):
    """The cut as csrc/cut_spt.cu computes it (per-slot windows), in numpy — checked against the literal restatement in oracle/."""
    S = len(sel)
    sizes = np.zeros(S, np.int64)
    for i in range(S):
        first = low = int(starts[sel[i]])
        high = int(starts[sel[i] + 1])
        pivot = (low + high) // 2
        while high - low > 1:
            if spt_max[pivot] > dist[i]:
                low = pivot
            else:
                high = pivot
            pivot = (low + high) // 2
        sizes[i] = high - first
    nonempty = np.nonzero(sizes > 0)[0]
    last = int(nonempty[-1]) if len(nonempty) else -1
    cut, counts = [], np.zeros(S, np.int32)
    for i in range(S):
        lo = 0 if i == 0 else 1
        hi = int(sizes[i]) - (1 if i == last else 0)
        if last < 0:
            hi = lo - 1
        elif sizes[i] <= 0:
            hi = 0 if i == 0 else lo - 1
        for o in range(lo, hi + 1):
            e = int(starts[sel[i]]) + o
            if 0 <= e < len(ids) and spt_min[e] < dist[i]:
                counts[i] += 1
                if ids[e] != 0:
                    cut.append(ids[e])
    prefix = np.zeros(S, np.int32)
    prefix[1:] = np.cumsum(counts)[:-1]
    return np.array(cut, np.int32), prefix


def make_morton_case(P=20000, seed=31, degenerate=False):
    """Positions + (min, max) as scene/gaussian_model.py:574 passes them (extent of the points); the extreme points sit exactly on
    the box faces (their scaled coordinate is 2^21, bit 21, which the code drops). `degenerate`: all z equal -> 0/0 on that axis."""
    rng = np.random.default_rng(seed)
    xyz = (rng.normal(0, 1, (P, 3)) * np.array([30.0, 12.0, 3.0])).astype(np.float32)
    if degenerate:
        xyz[:, 2] = 1.5
    return xyz, xyz.min(0), xyz.max(0)
