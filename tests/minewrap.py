"""Test helper: run the product path (diff_gaussian_rasterization._C over liblodgs.so) and expose the same
dictionary of intermediates as refshim.forward(), by viewing the opaque buffers through lodgs_*_layout."""
import ctypes as C

import torch

import lodgs
from diff_gaussian_rasterization import _C

E = lambda dev="cuda", dt=torch.float32: torch.empty(0, device=dev, dtype=dt)  # noqa: E731


def to_cuda(d):
    out = {}
    for k, v in d.items():
        out[k] = torch.from_numpy(v).cuda().contiguous() if hasattr(v, "dtype") and not torch.is_tensor(v) else v
    return out


def _view(buf, off, nbytes, dtype, shape):
    return buf[off:off + nbytes].view(dtype).reshape(shape)


def forward(sc, cam, bg, indices=None, parent_indices=None, ts=None, kids=None, do_depth=True, colors=None):
    ei = E(dt=torch.int32)
    args = (bg, indices if indices is not None else ei, parent_indices if parent_indices is not None else ei,
            ts if ts is not None else E(), kids if kids is not None else ei, sc["means3D"],
            colors if colors is not None else E(), sc["opacities"], sc["scales"], sc["rotations"], 1.0, E(), cam["viewmatrix"],
            cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], cam["H"], cam["W"], sc["shs"] if colors is None else E(),
            sc["sh_degree"], cam["campos"], False, False, do_depth)
    R, color, radii, geom, binning, img, invdepth, seen = _C.rasterize_gaussians(*args)
    torch.cuda.synchronize()
    P = radii.numel()
    W, H = cam["W"], cam["H"]
    res = {"R": R, "color": color, "invdepth": invdepth if do_depth else None, "radii": radii, "seen": seen, "P": P,
           "geom": geom, "binning": binning, "img": img}
    o = (C.c_size_t * 9)()
    lodgs.lib.lodgs_geom_layout(P, o)
    # per-Gaussian blend records (include/lodgs.h): rows of culled Gaussians are never written -> masked to zero here
    vis = (radii > 0)
    recs = torch.where(vis[:, None], _view(geom, o[1], P * 48, torch.float32, (P, 12)), torch.zeros((), device=geom.device))
    res["recs"] = recs
    res["depths"] = torch.where(vis, _view(geom, o[0], P * 4, torch.float32, (P,)), torch.zeros((), device=geom.device))
    res["means2D"] = recs[:, 0:2].contiguous()
    res["conic_opacity"] = recs[:, 2:6].contiguous()
    res["rgb"] = recs[:, 6:9].contiguous()
    res["cov3D"] = _view(geom, o[4], P * 24, torch.float32, (P, 6))
    res["clamped_bits"] = _view(geom, o[5], P, torch.uint8, (P,))
    res["tiles_touched"] = _view(geom, o[6], P * 4, torch.int32, (P,))
    res["rects"] = _view(geom, o[8], P * 8, torch.int32, (P, 2))
    if R > 0:
        b = (C.c_size_t * 5)()
        lodgs.lib.lodgs_binning_layout(R, b)
        res["tiles_unsorted"] = _view(binning, b[0], R * 4, torch.int32, (R,))   # depth-ordered emission
        res["vals_unsorted"] = _view(binning, b[1], R * 4, torch.int32, (R,))
        res["tiles_sorted"] = _view(binning, b[2], R * 4, torch.int32, (R,))
        res["point_list"] = _view(binning, b[3], R * 4, torch.int32, (R,))
        keys = torch.empty(R, dtype=torch.int64, device=geom.device)
        lodgs.check(lodgs.lib.lodgs_sorted_keys(P, R, geom.data_ptr(), binning.data_ptr(), keys.data_ptr(), lodgs.current_stream()))
        torch.cuda.synchronize()
        res["keys"] = keys
        im = (C.c_size_t * 3)()
        lodgs.lib.lodgs_image_layout(W, H, im)
        ntiles = ((W + 15) // 16) * ((H + 15) // 16)
        res["final_T"] = _view(img, im[0], W * H * 4, torch.float32, (H, W))
        res["n_contrib"] = _view(img, im[1], W * H * 4, torch.int32, (H, W))
        res["ranges"] = _view(img, im[2], ntiles * 8, torch.int32, (ntiles, 2))
        o_order = (im[2] + ntiles * 8 + 127) // 128 * 128                           # csrc/state.cuh: arrays are carved 128-byte aligned
        res["tile_order"] = _view(img, o_order, (ntiles + 1) * 4, torch.int32, (ntiles + 1,))   # [ntiles] = number of LONG tiles
    return res


def backward(sc, cam, bg, fwd, dL_dpix, dL_dinv=None, indices=None, parent_indices=None, ts=None, kids=None, colors=None):
    ei = E(dt=torch.int32)
    args = (bg, indices if indices is not None else ei, parent_indices if parent_indices is not None else ei,
            ts if ts is not None else E(), kids if kids is not None else ei, sc["means3D"], fwd["radii"],
            colors if colors is not None else E(), sc["opacities"], sc["scales"], sc["rotations"], 1.0, E(), cam["viewmatrix"],
            cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], dL_dpix, dL_dinv if dL_dinv is not None else E(),
            sc["shs"] if colors is None else E(), sc["sh_degree"], cam["campos"], fwd["geom"], fwd["R"], fwd["binning"], fwd["img"],
            False)
    m2, col, op, m3, cov, sh, s, r = _C.rasterize_gaussians_backward(*args)
    torch.cuda.synchronize()
    return {"dL_dmean2D": m2, "dL_dcolor": col, "dL_dopacity": op, "dL_dmean3D": m3, "dL_dcov3D": cov, "dL_dsh": sh,
            "dL_dscale": s, "dL_drot": r}
