"""TEST INFRASTRUCTURE: ctypes loader for oracle/_ref/libref_lodgs.so — the UNMODIFIED reference CUDA kernels
(built by oracle/Makefile from /root/reference) behind oracle/ref_shim.cu. Used by the GPU parity tests, by
__graft_entry__.smoke() and by bench.py --impl reference only; never by the product path."""
import ctypes as C
import os

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libref_lodgs.so")  # ROOT = repo root
_p, _i, _f = C.c_void_p, C.c_int, C.c_float
_lib = None


def available():
    return os.path.exists(REF_LIB)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(REF_LIB)
        _lib.ref_forward.restype = _i
        _lib.ref_forward.argtypes = [_i, _i, _i, _p, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f, _f, _i,
                                     _p, _p, _p, _p, _i, _p]
        _lib.ref_backward.restype = None
        _lib.ref_backward.argtypes = [_i, _i, _i, _i, _p, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f, _f,
                                      _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i]
        _lib.ref_copy_d2d.restype = _i
        _lib.ref_copy_d2d.argtypes = [_p, _p, C.c_size_t]
        _lib.ref_geom_ptrs.argtypes = [_i, _p]
        _lib.ref_binning_ptrs.argtypes = [_i, _p]
        _lib.ref_image_ptrs.argtypes = [_i, _p]
        _lib.ref_expand_to_size.restype = _i
        _lib.ref_expand_to_size.argtypes = [_i, _f, _p, _p, _p, _f, _f, _f, _p, _p, _p]
        _lib.ref_get_ts_indexed.restype = None
        _lib.ref_get_ts_indexed.argtypes = [_i, _p, _f, _p, _p, _f, _f, _f, _f, _f, _f, _p, _p, _p]
        _lib.ref_expand_to_size_dynamic.restype = _i
        _lib.ref_expand_to_size_dynamic.argtypes = [_i, _f, _p, _p, _p, _p, _f, _f, _f, _p, _p, _p]
        _lib.ref_get_ts_indexed_dynamic.restype = None
        _lib.ref_get_ts_indexed_dynamic.argtypes = [_i, _p, _f, _p, _p, _p, _f, _f, _f, _f, _f, _f, _p, _p, _p]
        _lib.ref_get_morton_code.restype = None
        _lib.ref_get_morton_code.argtypes = [_p, _p, _p, _p, _i]
        _lib.ref_get_spt_cut_run.restype = _i
        _lib.ref_get_spt_cut_run.argtypes = [_i, _p, _p, _p, _p, _p, _p]
        _lib.ref_get_spt_cut_fetch.restype = _i
        _lib.ref_get_spt_cut_fetch.argtypes = [_i, _i, _p, _p]
    return _lib


def get_spt_cut(gaussian_indices, SPT_starts, SPT_max, SPT_min, SPT_indices, SPT_distances):
    """Switching::getSPTCut on CUDA tensors -> (cut i32[K], counts prefix i32[S]) as CUDA tensors."""
    S = SPT_indices.numel()
    torch.cuda.synchronize()
    K = lib().ref_get_spt_cut_run(S, gaussian_indices.data_ptr(), SPT_starts.data_ptr(), SPT_max.data_ptr(), SPT_min.data_ptr(),
                                  SPT_indices.data_ptr(), SPT_distances.data_ptr())
    cut = torch.empty(max(K, 0), dtype=torch.int32, device="cuda")
    counts = torch.empty(S, dtype=torch.int32, device="cuda")
    rc = lib().ref_get_spt_cut_fetch(max(K, 0), S, cut.data_ptr() if K > 0 else None, counts.data_ptr())
    if rc != 0:
        raise RuntimeError("ref_get_spt_cut_fetch: cuda error %d" % rc)
    return cut, counts


def ptr(t):
    return None if t is None or t.numel() == 0 else t.data_ptr()


def _from_dev(p, nbytes, dtype, shape):
    """Copy `nbytes` from raw device pointer p into a new CUDA tensor."""
    out = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    rc = lib().ref_copy_d2d(out.data_ptr(), p, nbytes)
    assert rc == 0, "ref_copy_d2d failed with CUDA error %d" % rc
    return out.view(dtype).reshape(shape)


def forward(sc, cam, bg, indices=None, parent_indices=None, ts=None, kids=None, do_depth=True, colors=None):
    """Run the reference forward on CUDA tensors in dict `sc` (means3D, shs, opacities, scales, rotations, sh_degree)."""
    L = lib()
    P = sc["means3D"].size(0) if indices is None else indices.size(0)
    W, H = cam["W"], cam["H"]
    dev = "cuda"
    out_color = torch.zeros(3, H, W, device=dev)
    out_inv = torch.zeros(1, H, W, device=dev) if do_depth else None
    radii = torch.zeros(P, dtype=torch.int32, device=dev)
    rects = torch.zeros(P, 2, dtype=torch.int32, device=dev)
    seen = torch.zeros(P, dtype=torch.int32, device=dev)
    shs = sc.get("shs") if colors is None else None
    M = shs.size(1) if shs is not None else 0
    torch.cuda.synchronize()
    R = L.ref_forward(P, sc["sh_degree"], M, ptr(bg), W, H, ptr(indices), ptr(parent_indices), ptr(ts), ptr(kids), ptr(sc["means3D"]),
                      ptr(shs), ptr(colors), ptr(sc["opacities"]), ptr(sc["scales"]), 1.0, ptr(sc["rotations"]), None,
                      ptr(cam["viewmatrix"]), ptr(cam["projmatrix"]), ptr(cam["campos"]), cam["tanfovx"], cam["tanfovy"], 0,
                      ptr(out_color), ptr(out_inv), ptr(radii), ptr(rects), 0, ptr(seen))
    torch.cuda.synchronize()
    res = {"R": R, "color": out_color, "invdepth": out_inv, "radii": radii, "rects": rects, "seen": seen, "P": P}
    g = (C.c_void_p * 10)()
    L.ref_geom_ptrs(P, g)
    res["depths"] = _from_dev(g[0], P * 4, torch.float32, (P,))
    res["means2D"] = _from_dev(g[4], P * 8, torch.float32, (P, 2))
    res["cov3D"] = _from_dev(g[5], P * 24, torch.float32, (P, 6))
    res["conic_opacity"] = _from_dev(g[6], P * 16, torch.float32, (P, 4))
    res["rgb"] = _from_dev(g[7], P * 12, torch.float32, (P, 3))
    res["tiles_touched"] = _from_dev(g[8], P * 4, torch.int32, (P,))
    res["point_offsets"] = _from_dev(g[9], P * 4, torch.int32, (P,))
    res["clamped"] = _from_dev(g[2] if parent_indices is not None else g[1], P * 3, torch.uint8, (P, 3))
    if R > 0:
        b = (C.c_void_p * 4)()
        L.ref_binning_ptrs(R, b)
        res["keys_unsorted"] = _from_dev(b[0], R * 8, torch.int64, (R,))
        res["vals_unsorted"] = _from_dev(b[1], R * 4, torch.int32, (R,))
        res["keys"] = _from_dev(b[2], R * 8, torch.int64, (R,))
        res["point_list"] = _from_dev(b[3], R * 4, torch.int32, (R,))
        im = (C.c_void_p * 3)()
        L.ref_image_ptrs(W * H, im)
        ntiles = ((W + 15) // 16) * ((H + 15) // 16)
        res["final_T"] = _from_dev(im[0], W * H * 4, torch.float32, (H, W))
        res["n_contrib"] = _from_dev(im[1], W * H * 4, torch.int32, (H, W))
        res["ranges"] = _from_dev(im[2], ntiles * 8, torch.int32, (ntiles, 2))
    return res


def backward(sc, cam, bg, fwd, dL_dpix, dL_dinv=None, indices=None, parent_indices=None, ts=None, kids=None, colors=None):
    L = lib()
    fullP = sc["means3D"].size(0)
    P = fwd["P"]
    W, H = cam["W"], cam["H"]
    shs = sc.get("shs") if colors is None else None
    M = shs.size(1) if shs is not None else 0
    z = lambda *s: torch.zeros(*s, device="cuda")  # noqa: E731
    g = {"dL_dmean2D": z(fullP, 3), "dL_dconic": z(fullP, 4), "dL_dopacity": z(fullP, 1), "dL_dcolor": z(fullP, 3),
         "dL_dinvdepth": z(fullP, 1) if dL_dinv is not None else None, "dL_dmean3D": z(fullP, 3), "dL_dcov3D": z(fullP, 6),
         "dL_dsh": z(fullP, M, 3), "dL_dscale": z(fullP, 3), "dL_drot": z(fullP, 4)}
    torch.cuda.synchronize()
    L.ref_backward(P, sc["sh_degree"], M, fwd["R"], ptr(bg), W, H, ptr(indices), ptr(parent_indices), ptr(ts), ptr(kids),
                   ptr(sc["means3D"]), ptr(shs), ptr(colors), ptr(sc["opacities"]), ptr(sc["scales"]), 1.0, ptr(sc["rotations"]), None,
                   ptr(cam["viewmatrix"]), ptr(cam["projmatrix"]), ptr(cam["campos"]), cam["tanfovx"], cam["tanfovy"], ptr(fwd["radii"]),
                   ptr(dL_dpix), ptr(dL_dinv), ptr(g["dL_dmean2D"]), ptr(g["dL_dconic"]), ptr(g["dL_dopacity"]), ptr(g["dL_dcolor"]),
                   ptr(g["dL_dinvdepth"]), ptr(g["dL_dmean3D"]), ptr(g["dL_dcov3D"]), ptr(g["dL_dsh"]), ptr(g["dL_dscale"]),
                   ptr(g["dL_drot"]), 0)
    torch.cuda.synchronize()
    return g
