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
