"""TEST INFRASTRUCTURE: ctypes loader for oracle/_ref/libref_ar.so — the UNMODIFIED alt-rasterizer kernels of the reference
(submodules/alt-rasterizer, built by oracle/Makefile from /root/reference) behind oracle/ref_shim_ar.cu. Used by
tests/test_gpu_alt_rasterizer.py only; never by the product path."""
import ctypes as C
import os

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libref_ar.so")
_p, _i, _f, _u = C.c_void_p, C.c_int, C.c_float, C.c_uint
_lib = None


def available():
    return os.path.exists(REF_LIB)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(REF_LIB)
        _lib.refar_forward.restype = _i
        _lib.refar_forward.argtypes = [_i, _i, _i, _p, _i, _i, _p, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f, _f, _i, _p, _p, _i, _p, _i, _p]
        _lib.refar_backward.restype = None
        _lib.refar_backward.argtypes = [_i, _i, _i, _i, _i, _p, _i, _i, _p, _p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _p, _f, _f, _p, _p, _p,
                                        _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i]
        _lib.refar_adam.restype = None
        _lib.refar_adam.argtypes = [_p, _p, _p, _p, _p, _f, _f, _f, _f, _u, _u]
        _lib.refar_relocation.restype = None
        _lib.refar_relocation.argtypes = [_i, _p, _p, _p, _p, _i, _p, _p]
    return _lib


def ptr(t):
    return None if t is None or t.numel() == 0 else t.data_ptr()


def forward(sc, cam, bg, antialiasing):
    """sc: CUDA tensors means3D, dc [P,1,3], shs_rest [P,M,3], opacities, scales, rotations, sh_degree."""
    L = lib()
    P = sc["means3D"].size(0)
    W, H = cam["W"], cam["H"]
    out_color = torch.zeros(3, H, W, device="cuda")
    out_inv = torch.zeros(1, H, W, device="cuda")
    radii = torch.zeros(P, dtype=torch.int32, device="cuda")
    nb = C.c_int(0)
    torch.cuda.synchronize()
    R = L.refar_forward(P, sc["sh_degree"], sc["shs_rest"].size(1), ptr(bg), W, H, ptr(sc["means3D"]), ptr(sc["dc"]), ptr(sc["shs_rest"]),
                        None, ptr(sc["opacities"]), ptr(sc["scales"]), 1.0, ptr(sc["rotations"]), None, ptr(cam["viewmatrix"]),
                        ptr(cam["projmatrix"]), ptr(cam["campos"]), cam["tanfovx"], cam["tanfovy"], 0, ptr(out_color), ptr(out_inv),
                        int(antialiasing), ptr(radii), 0, C.addressof(nb))
    torch.cuda.synchronize()
    return {"R": R, "B": nb.value, "color": out_color, "invdepth": out_inv, "radii": radii, "P": P}


def backward(sc, cam, bg, fwd, dL_dpix, dL_dinv, antialiasing):
    L = lib()
    P = fwd["P"]
    W, H = cam["W"], cam["H"]
    M = sc["shs_rest"].size(1)
    z = lambda *s: torch.zeros(*s, device="cuda")  # noqa: E731
    g = {"dL_dmean2D": z(P, 3), "dL_dconic": z(P, 4), "dL_dopacity": z(P, 1), "dL_dcolor": z(P, 3), "dL_dinvdepth": z(P, 1),
         "dL_dmean3D": z(P, 3), "dL_dcov3D": z(P, 6), "dL_ddc": z(P, 1, 3), "dL_dsh": z(P, M, 3), "dL_dscale": z(P, 3), "dL_drot": z(P, 4)}
    torch.cuda.synchronize()
    L.refar_backward(P, sc["sh_degree"], M, fwd["R"], fwd["B"], ptr(bg), W, H, ptr(sc["means3D"]), ptr(sc["dc"]), ptr(sc["shs_rest"]), None,
                     ptr(sc["opacities"]), ptr(sc["scales"]), 1.0, ptr(sc["rotations"]), None, ptr(cam["viewmatrix"]),
                     ptr(cam["projmatrix"]), ptr(cam["campos"]), cam["tanfovx"], cam["tanfovy"], ptr(fwd["radii"]), ptr(dL_dpix),
                     ptr(dL_dinv), ptr(g["dL_dmean2D"]), ptr(g["dL_dconic"]), ptr(g["dL_dopacity"]), ptr(g["dL_dcolor"]),
                     ptr(g["dL_dinvdepth"]), ptr(g["dL_dmean3D"]), ptr(g["dL_dcov3D"]), ptr(g["dL_ddc"]), ptr(g["dL_dsh"]),
                     ptr(g["dL_dscale"]), ptr(g["dL_drot"]), int(antialiasing), 0)
    torch.cuda.synchronize()
    return g


def adam(param, grad, m, v, visible, lr, b1, b2, eps, N, M):
    torch.cuda.synchronize()
    lib().refar_adam(ptr(param), ptr(grad), ptr(m), ptr(v), ptr(visible), lr, b1, b2, eps, N, M)
    torch.cuda.synchronize()


def relocation(opacity_old, scale_old, N, binoms, n_max):
    P = opacity_old.numel()
    o = torch.zeros(P, device="cuda")
    s = torch.zeros(3 * P, device="cuda")
    torch.cuda.synchronize()
    lib().refar_relocation(P, ptr(opacity_old), ptr(scale_old), ptr(N), ptr(binoms), n_max, ptr(o), ptr(s))
    torch.cuda.synchronize()
    return o, s


# ---- the reference's own scratch state of its LAST forward (for the defect predicate below and tests/gpu_diag_ar.py)
def _copy_out(ptr_, n, dtype):
    t = torch.empty(n, dtype=dtype, device="cuda")
    L = lib()
    L.refar_copy.restype = _i
    rc = L.refar_copy(C.c_void_p(t.data_ptr()), C.c_void_p(ptr_), C.c_size_t(n * t.element_size()))
    assert rc == 0, rc
    return t


def internals(fwd, W, H):
    """Copies of the reference's ranges / point_list / n_contrib / max_contrib / bucket state / per-Gaussian 2-D state after `forward`."""
    out = (C.c_void_p * 16)()
    L = lib()
    L.refar_internals.restype = None
    L.refar_internals(C.c_int(fwd["P"]), C.c_int(W), C.c_int(H), C.c_int(fwd["R"]), C.c_int(fwd["B"]), out)
    nt = ((W + 15) // 16) * ((H + 15) // 16)
    P, R, B = fwd["P"], fwd["R"], fwd["B"]
    i32, f32 = torch.int32, torch.float32
    return {"ranges": _copy_out(out[0], nt * 2, i32).view(nt, 2), "point_list": _copy_out(out[1], R, i32),
            "n_contrib": _copy_out(out[2], W * H, i32), "max_contrib": _copy_out(out[3], nt, i32),
            "bucket_offsets": _copy_out(out[4], nt, i32), "bucket_to_tile": _copy_out(out[5], B, i32),
            "sampled_T": _copy_out(out[6], B * 256, f32).view(B, 256), "means2D": _copy_out(out[7], P * 2, f32).view(P, 2),
            "conic_opacity": _copy_out(out[8], P * 4, f32).view(P, 4), "rgb": _copy_out(out[9], P * 3, f32).view(P, 3),
            "final_T": _copy_out(out[10], W * H, f32)}


def skipped_instances(st, W, H):
    """The instances the reference's backward never visits, derived from ITS state.

    Cause (reference defect, not a convention): AR/cuda_rasterizer/forward.cu:427-431 reduces `last_contributor` with
    `cub::BlockReduce<uint32_t, BLOCK_SIZE>` — a specialisation for a ONE-dimensional block of 256 threads — inside renderCUDA, which is
    launched with a 16x16 block (forward.cu:454). CUB derives the linear thread id from its template block shape (cub/util_ptx.cuh
    RowMajorTid: threadIdx.y is ignored when BLOCK_DIM_Y == 1), so all eight warps take warp slot 0 and thread (0,0) combines the maximum
    over ITS warp — the tile's top two pixel rows — with seven never-written (uninitialised) shared-memory slots. `max_contrib[tile]` is
    therefore not the tile's last contributor, and AR/cuda_rasterizer/backward.cu:503 (`if (bucket_idx_in_tile * 32 >= max_contrib[tile_id])
    return;`) drops whole 32-splat buckets that pixels of the other fourteen rows did blend.

    -> (lost [n_list] bool per sorted-list entry, tile_of [n_list], pos [n_list], true_max [tiles])"""
    nt = st["ranges"].size(0)
    gx = (W + 15) // 16
    lens = (st["ranges"][:, 1] - st["ranges"][:, 0]).long()
    n_list = int(lens.sum())
    tile_of = torch.repeat_interleave(torch.arange(nt, device="cuda"), lens)
    # the tile lists are contiguous from entry 0 in tile order (entries culled per tile are padded to the END of the sorted list)
    start = st["ranges"][:, 0].long()
    pos = torch.arange(n_list, device="cuda") - (start[tile_of] - int(start[lens > 0][0]) if n_list else 0)
    nc = torch.zeros(((H + 15) // 16) * 16, gx * 16, dtype=torch.int32, device="cuda")
    nc[:H, :W] = st["n_contrib"].view(H, W)
    true_max = nc.view((H + 15) // 16, 16, gx, 16).permute(0, 2, 1, 3).reshape(nt, 256).max(1).values.long()
    bucket_start = (pos // 32) * 32
    mc = st["max_contrib"].long() & 0xffffffff   # uint32 in the reference; the uninitialised shared-memory slots often make it huge
    lost = (bucket_start >= mc[tile_of]) & (pos < true_max[tile_of])
    return lost, tile_of, pos, true_max


def color_grad_of_instances(st, W, H, dL, which):
    """dL_dcolor contribution [P,3] of the sorted-list entries selected by `which` (bool per entry), by replaying the reference's own lists
    in torch (alpha, running T, `pos < n_contrib` as in AR forward.cu:364-418). Small scenes only."""
    lost, tile_of, pos, _ = skipped_instances(st, W, H)
    P = st["means2D"].size(0)
    out = torch.zeros(P, 3, device="cuda", dtype=torch.float64)
    gx = (W + 15) // 16
    dLf = dL.reshape(3, -1).double()
    k = torch.arange(256, device="cuda")
    first = int(st["ranges"][(st["ranges"][:, 1] > st["ranges"][:, 0])][0, 0]) if which.numel() else 0
    for t in torch.unique(tile_of[which]).tolist():
        lo, hi = (int(x) for x in st["ranges"][t])
        ids = st["point_list"][lo:hi].long()
        ty, tx = divmod(t, gx)
        px, py = tx * 16 + k % 16, ty * 16 + k // 16
        valid = (px < W) & (py < H)
        pid = (py * W + px).clamp(max=W * H - 1)
        xy, co = st["means2D"][ids], st["conic_opacity"][ids]
        dx, dy = xy[:, 0:1] - px.float()[None], xy[:, 1:2] - py.float()[None]
        power = -0.5 * (co[:, 0:1] * dx * dx + co[:, 2:3] * dy * dy) - co[:, 1:2] * dx * dy
        alpha = torch.clamp_max(co[:, 3:4] * torch.exp(power), 0.99)
        ok = (power <= 0) & (alpha >= 1.0 / 255.0)
        p_ = torch.arange(hi - lo, device="cuda")[:, None]
        a = torch.where(ok, alpha, torch.zeros_like(alpha))
        Tb = torch.cumprod(torch.cat([torch.ones(1, 256, device="cuda"), 1 - a[:-1]], 0), 0)
        w = torch.where(ok & (p_ < st["n_contrib"][pid][None]) & valid[None], a * Tb, torch.zeros_like(a)).double()
        sel = which[lo - first:hi - first]
        contrib = w @ dLf[:, pid].t()                              # [n, 3]
        out.index_add_(0, ids[sel], contrib[sel])
    return out
