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
