"""TEST INFRASTRUCTURE ONLY: numpy front-end of oracle/_build/liboracle.so (the CPU restatement in cut_oracle.c /
raster_oracle.c). Imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg — never by the product."""
import ctypes as C
import os
import subprocess
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "liboracle.so")
_lib = None

_fp, _ip, _up, _u8, _u64 = (C.POINTER(C.c_float), C.POINTER(C.c_int32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint8),
                            C.POINTER(C.c_uint64))


class Scene(C.Structure):
    _fields_ = [("P", C.c_int), ("fullP", C.c_int), ("D", C.c_int), ("M", C.c_int), ("W", C.c_int), ("H", C.c_int),
                ("indices", _ip), ("parent_indices", _ip), ("ts", _fp), ("kids", _ip), ("means3D", _fp), ("shs", _fp),
                ("colors_precomp", _fp), ("opacities", _fp), ("scales", _fp), ("rotations", _fp), ("scale_modifier", C.c_float),
                ("viewmatrix", _fp), ("projmatrix", _fp), ("campos", _fp), ("tan_fovx", C.c_float), ("tan_fovy", C.c_float),
                ("background", _fp), ("do_depth", C.c_int)]


class State(C.Structure):
    _fields_ = [("radii", _ip), ("rects", _ip), ("tiles_touched", _up), ("point_offsets", _up), ("depths", _fp), ("means2D", _fp),
                ("conic_opacity", _fp), ("rgb", _fp), ("cov3D", _fp), ("clamped", _u8), ("R", C.c_int), ("keys_unsorted", _u64),
                ("vals_unsorted", _up), ("keys", _u64), ("point_list", _up), ("ranges", _up), ("final_T", _fp), ("n_contrib", _up),
                ("color", _fp), ("invdepth", _fp), ("seen", _ip)]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            subprocess.run(["make", "-C", HERE, "oracle"], check=True, stdout=subprocess.DEVNULL)
        _lib = C.CDLL(LIB)
        _lib.orc_expand_to_size.restype = C.c_int
        _lib.orc_expand_to_size.argtypes = [_ip, _fp, C.c_int, C.c_float, _fp, _ip, _ip, _ip]
        _lib.orc_get_interpolation_weights.argtypes = [_ip, C.c_int, C.c_float, _ip, _fp, _fp, _fp, _ip]
        _lib.orc_expand_to_size_dynamic.restype = C.c_int
        _lib.orc_expand_to_size_dynamic.argtypes = [_ip, _fp, _fp, C.c_int, C.c_float, _fp, _fp, _ip, _ip, _ip]
        _lib.orc_get_interpolation_weights_dynamic.argtypes = [_ip, C.c_int, C.c_float, _ip, _fp, _fp, _fp, _fp, _ip]
        _lib.orc_get_spt_cut.restype = C.c_int
        _lib.orc_get_spt_cut.argtypes = [C.c_int, _ip, C.c_int, _ip, _fp, _fp, _ip, _fp, _ip, _ip, _ip]
        _lib.orc_get_morton_codes.argtypes = [_fp, _fp, _fp, C.POINTER(C.c_int64), C.c_int]
        _lib.orc_expand_to_target.restype = C.c_long
        _lib.orc_expand_to_target.argtypes = [_ip, C.c_int, _ip]
        _lib.orc_forward.restype = C.c_int
        _lib.orc_forward.argtypes = [C.POINTER(Scene), C.POINTER(State)]
        _lib.orc_backward.argtypes = [C.POINTER(Scene), C.POINTER(State)] + [_fp] * 12
        _lib.orc_free_instances.argtypes = [C.POINTER(State)]
        _lib.orc_num_threads.restype = C.c_int
    return _lib


def _f(a):
    return None if a is None else a.ctypes.data_as(_fp)


def _i(a):
    return None if a is None else a.ctypes.data_as(_ip)


def _c(a, dt):
    return None if a is None else np.ascontiguousarray(a, dtype=dt)


def expand_to_size(nodes, boxes, tau, viewpoint):
    """-> (render_indices[K], parent_indices[K], node_ids[K])"""
    nodes, boxes, vp = _c(nodes, np.int32), _c(boxes, np.float32), _c(viewpoint, np.float32)
    N = nodes.shape[0]
    cap = int(np.maximum(nodes[:, 3] + nodes[:, 4], 1).sum())
    ri, pi, ni = (np.zeros(cap, np.int32) for _ in range(3))
    K = lib().orc_expand_to_size(_i(nodes), _f(boxes), N, float(tau), _f(vp), _i(ri), _i(pi), _i(ni))
    return ri[:K].copy(), pi[:K].copy(), ni[:K].copy()


def expand_to_size_omp(nodes, boxes, tau, viewpoint, bufs=None):
    """orc_expand_to_size_omp (all host cores) -> (render_indices[K], parent_indices[K], node_ids[K], threads used). `bufs` = reusable
    (ri, pi, ni, counts) arrays of length >= N (one Gaussian per node) for timing loops."""
    import ctypes as C
    nodes, boxes, vp = _c(nodes, np.int32), _c(boxes, np.float32), _c(viewpoint, np.float32)
    N = nodes.shape[0]
    if bufs is None:
        cap = int(np.maximum(nodes[:, 3] + nodes[:, 4], 1).sum())
        bufs = tuple(np.zeros(max(cap, N), np.int32) for _ in range(4))
    ri, pi, ni, counts = bufs
    used = C.c_int(0)
    fn = lib().orc_expand_to_size_omp
    fn.restype = C.c_int
    K = fn(_i(nodes), _f(boxes), N, C.c_float(float(tau)), _f(vp), _i(ri), _i(pi), _i(ni), _i(counts), C.byref(used))
    return ri[:K], pi[:K], ni[:K], used.value


def get_interpolation_weights(node_ids, tau, nodes, boxes, viewpoint):
    idx, nodes, boxes, vp = _c(node_ids, np.int32), _c(nodes, np.int32), _c(boxes, np.float32), _c(viewpoint, np.float32)
    K = idx.shape[0]
    ts, kids = np.zeros(K, np.float32), np.zeros(K, np.int32)
    lib().orc_get_interpolation_weights(_i(idx), K, float(tau), _i(nodes), _f(boxes), _f(vp), _f(ts), _i(kids))
    return ts, kids


def expand_to_size_dynamic(nodes, positions, scales, tau, viewpoint, zdir, parent_fill=-7):
    """-> (render_indices[K], parent_indices[K] (entries of roots keep parent_fill, as the reference leaves them), node_ids[K])"""
    nodes, pos, sca = _c(nodes, np.int32), _c(positions, np.float32), _c(scales, np.float32)
    vp, zd = _c(viewpoint, np.float32), _c(zdir, np.float32)
    N = nodes.shape[0]
    ri, ni = np.zeros(N, np.int32), np.zeros(N, np.int32)
    pi = np.full(N, parent_fill, np.int32)
    K = lib().orc_expand_to_size_dynamic(_i(nodes), _f(pos), _f(sca), N, float(tau), _f(vp), _f(zd), _i(ri), _i(pi), _i(ni))
    return ri[:K].copy(), pi[:K].copy(), ni[:K].copy()


def get_interpolation_weights_dynamic(node_ids, tau, nodes, positions, scales, viewpoint):
    idx, nodes, pos, sca, vp = (_c(node_ids, np.int32), _c(nodes, np.int32), _c(positions, np.float32), _c(scales, np.float32),
                                _c(viewpoint, np.float32))
    K = idx.shape[0]
    ts, kids = np.zeros(K, np.float32), np.zeros(K, np.int32)
    lib().orc_get_interpolation_weights_dynamic(_i(idx), K, float(tau), _i(nodes), _f(pos), _f(sca), _f(vp), _f(ts), _i(kids))
    return ts, kids


def get_spt_cut(gaussian_indices, SPT_starts, SPT_max, SPT_min, SPT_indices, SPT_distances):
    """Switching::getSPTCut restated on numpy arrays -> (cut i32[K], exclusive prefix of the per-slot counts i32[S])."""
    gi, st, si = _c(gaussian_indices, np.int32), _c(SPT_starts, np.int32), _c(SPT_indices, np.int32)
    mx, mn, sd = _c(SPT_max, np.float32), _c(SPT_min, np.float32), _c(SPT_distances, np.float32)
    S, n = si.size, min(gi.size, mx.size, mn.size)
    pos = np.zeros(1, np.int32)
    lib().orc_get_spt_cut(S, _i(gi), n, _i(st), _f(mx), _f(mn), _i(si), _f(sd), None, None, _i(pos))
    cut = np.zeros(max(int(pos[0]), 1), np.int32)
    counts = np.zeros(S, np.int32)
    K = lib().orc_get_spt_cut(S, _i(gi), n, _i(st), _f(mx), _f(mn), _i(si), _f(sd), _i(cut), _i(counts), _i(pos))
    return cut[:K].copy(), counts


def get_morton_codes(xyz, mn, mx):
    """MORTON::getMortonCode restated on numpy arrays -> i64[P]."""
    x, lo, hi = _c(xyz, np.float32), _c(mn, np.float32), _c(mx, np.float32)
    codes = np.zeros(x.shape[0], np.int64)
    lib().orc_get_morton_codes(_f(x), _f(lo), _f(hi), codes.ctypes.data_as(C.POINTER(C.c_int64)), x.shape[0])
    return codes


def expand_to_target(nodes, target):
    """Traversal::expandToTarget restated -> i32 Gaussian ids."""
    n = _c(nodes, np.int32)
    cnt = lib().orc_expand_to_target(_i(n), int(target), None)
    out = np.zeros(max(cnt, 1), np.int32)
    lib().orc_expand_to_target(_i(n), int(target), _i(out))
    return out[:cnt].copy()


def _scene(sc, cam, bg, indices, parent_indices, ts, kids, do_depth, colors, keep):
    a = {k: _c(sc[k], np.float32) for k in ("means3D", "opacities", "scales", "rotations")}
    a["shs"] = _c(sc["shs"], np.float32) if colors is None else None
    a["colors"] = _c(colors, np.float32)
    a["indices"], a["parent"], a["kids"] = _c(indices, np.int32), _c(parent_indices, np.int32), _c(kids, np.int32)
    a["ts"] = _c(ts, np.float32)
    a["view"], a["proj"], a["campos"] = _c(cam["viewmatrix"], np.float32), _c(cam["projmatrix"], np.float32), _c(cam["campos"], np.float32)
    a["bg"] = _c(bg, np.float32)
    keep.append(a)
    fullP = a["means3D"].shape[0]
    P = fullP if indices is None else a["indices"].shape[0]
    M = a["shs"].shape[1] if a["shs"] is not None else 0
    s = Scene(P, fullP, int(sc.get("sh_degree", 3)), M, cam["W"], cam["H"], _i(a["indices"]), _i(a["parent"]), _f(a["ts"]), _i(a["kids"]),
              _f(a["means3D"]), _f(a["shs"]), _f(a["colors"]), _f(a["opacities"]), _f(a["scales"]), _f(a["rotations"]), 1.0, _f(a["view"]),
              _f(a["proj"]), _f(a["campos"]), cam["tanfovx"], cam["tanfovy"], _f(a["bg"]), 1 if do_depth else 0)
    return s, P, fullP, M


def forward(sc, cam, bg, indices=None, parent_indices=None, ts=None, kids=None, do_depth=True, colors=None, want_state=False):
    """numpy in, dict of numpy out (same keys as the GPU test helpers)."""
    keep = []
    s, P, fullP, M = _scene(sc, cam, bg, indices, parent_indices, ts, kids, do_depth, colors, keep)
    W, H = cam["W"], cam["H"]
    ntiles = ((W + 15) // 16) * ((H + 15) // 16)
    o = {"radii": np.zeros(P, np.int32), "rects": np.zeros((P, 2), np.int32), "tiles_touched": np.zeros(P, np.uint32),
         "point_offsets": np.zeros(P, np.uint32), "depths": np.zeros(P, np.float32), "means2D": np.zeros((P, 2), np.float32),
         "conic_opacity": np.zeros((P, 4), np.float32), "rgb": np.zeros((P, 3), np.float32), "cov3D": np.zeros((P, 6), np.float32),
         "clamped": np.zeros((P, 3), np.uint8), "ranges": np.zeros((ntiles, 2), np.uint32), "final_T": np.zeros((H, W), np.float32),
         "n_contrib": np.zeros((H, W), np.uint32), "color": np.zeros((3, H, W), np.float32), "invdepth": np.zeros((1, H, W), np.float32),
         "seen": np.zeros(P, np.int32)}
    st = State()
    for k, tp in (("radii", _ip), ("rects", _ip), ("tiles_touched", _up), ("point_offsets", _up), ("depths", _fp), ("means2D", _fp),
                  ("conic_opacity", _fp), ("rgb", _fp), ("cov3D", _fp), ("clamped", _u8), ("ranges", _up), ("final_T", _fp),
                  ("n_contrib", _up), ("color", _fp), ("invdepth", _fp), ("seen", _ip)):
        setattr(st, k, o[k].ctypes.data_as(tp))
    R = lib().orc_forward(C.byref(s), C.byref(st))
    o["R"], o["P"] = R, P
    if R > 0:
        o["keys_unsorted"] = np.ctypeslib.as_array(st.keys_unsorted, (R,)).copy()
        o["vals_unsorted"] = np.ctypeslib.as_array(st.vals_unsorted, (R,)).copy()
        o["keys"] = np.ctypeslib.as_array(st.keys, (R,)).copy()
        o["point_list"] = np.ctypeslib.as_array(st.point_list, (R,)).copy()
    if want_state:
        o["_state"], o["_scene"], o["_keep"] = st, s, keep
    else:
        lib().orc_free_instances(C.byref(st))
    return o


def backward(fwd, dL_dpix, dL_dinv=None):
    """fwd = forward(..., want_state=True). -> dict of gradients (numpy)."""
    s, st = fwd["_scene"], fwd["_state"]
    fullP, M = s.fullP, s.M
    z = lambda *sh: np.zeros(sh, np.float32)  # noqa: E731
    g = {"dL_dmean2D": z(fullP, 3), "dL_dconic": z(fullP, 4), "dL_dopacity": z(fullP, 1), "dL_dcolor": z(fullP, 3),
         "dL_dinvdepth": z(fullP, 1) if dL_dinv is not None else None, "dL_dmean3D": z(fullP, 3), "dL_dcov3D": z(fullP, 6),
         "dL_dsh": z(fullP, M, 3), "dL_dscale": z(fullP, 3), "dL_drot": z(fullP, 4)}
    dp, di = _c(dL_dpix, np.float32), _c(dL_dinv, np.float32)
    lib().orc_backward(C.byref(s), C.byref(st), _f(dp), _f(di), _f(g["dL_dmean2D"]), _f(g["dL_dconic"]), _f(g["dL_dopacity"]),
                       _f(g["dL_dcolor"]), _f(g["dL_dinvdepth"]), _f(g["dL_dmean3D"]), _f(g["dL_dcov3D"]), _f(g["dL_dsh"]),
                       _f(g["dL_dscale"]), _f(g["dL_drot"]))
    return g


def free(fwd):
    if "_state" in fwd:
        lib().orc_free_instances(C.byref(fwd["_state"]))
        del fwd["_state"]


def bench_sample(W, H, n_gauss=1_000_000, budget_s=12.0):
    """cpu_baseline for bench.py: fwd+bwd of workload A on the host cores, bounded to ~budget_s by rendering a centred
    crop-free DOWNSCALED view (W/2 x H/2, same scene and camera field of view) repeated until the budget is spent — Mpix/s scale
    with the pixel count."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(HERE), "hierarchical-lod-gaussians_b200"))
    from lodgs import synth
    w, h = W // 2, H // 2
    cam = synth.make_camera(W=w, H=h, fx=1200.0 / 2, fy=1200.0 / 2)
    sc = synth.make_gaussians(n_gauss, seed=0, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"])
    bg = np.zeros(3, np.float32)
    target = np.random.default_rng(1234).random((3, h, w), dtype=np.float32)
    t0 = time.time()
    reps = 0
    while True:
        f = forward(sc, cam, bg, want_state=True)
        dL = (f["color"] - target) * (2.0 / f["color"].size)
        backward(f, dL, np.zeros((1, h, w), np.float32))
        free(f)
        reps += 1
        if time.time() - t0 > budget_s and reps >= 3:
            break
    dt = (time.time() - t0) / reps
    return {"value": round(w * h / 1e6 / dt, 3), "unit": "Mpix/s", "cores": int(lib().orc_num_threads()), "kind": "port",
            "sample": "oracle/raster_oracle.c (OpenMP) fwd+bwd, %d Gaussians, %dx%d view (1/4 of the 1080p pixels, same field of view), "
                      "%d repetitions, %.2f s per fwd+bwd, %.0f s of host time" % (n_gauss, w, h, reps, dt, reps * dt)}
