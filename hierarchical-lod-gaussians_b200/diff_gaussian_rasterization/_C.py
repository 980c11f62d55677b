"""diff_gaussian_rasterization._C — same functions, argument order and return tuples as the reference's pybind
module (HR/ext.cpp:15-19, HR/rasterize_points.cu:36-245), implemented as a thin ctypes shim over liblodgs.so.
`mark_visible`, which the reference's Python calls but never binds, is provided too."""
import ctypes as C

import torch

import lodgs

import lodgs as _lodgs
_lodgs.load()   # the plugin cannot work without the native library: fail at import, loudly
from lodgs import check, current_stream, lib, ptr

NUM_CHANNELS = 3

# Persistent high-water scratch arena per device for the binning buffer would save an allocation per call; the
# reference hands fresh torch byte tensors back to Python (rasterize_points.cu:85-92) and so do we: they are what
# autograd saves for backward.

LOD_IN_RECORDS = True   # False: stage 2 patches t and 1/kids into the records (lod_patch_kernel); kept for the C ABI's older call sequence


def _byte_buffer(nbytes, device):
    # torch allocations are 512-byte aligned, which satisfies the library's 128-byte requirement
    return torch.empty(int(nbytes), dtype=torch.uint8, device=device)


def _f(t):
    return t.contiguous() if t is not None else t


def rasterize_gaussians(background, indices, parent_indices, ts, kids, means3D, colors, opacity, scales, rotations,
                        scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height,
                        image_width, sh, degree, campos, prefiltered, debug, do_depth):
    return forward_impl(background, indices, parent_indices, ts, kids, means3D, colors, opacity, scales, rotations, scale_modifier,
                        cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                        prefiltered, debug, do_depth, 0)


def forward_impl(background, indices, parent_indices, ts, kids, means3D, colors, opacity, scales, rotations, scale_modifier,
                 cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                 prefiltered, debug, do_depth, _flags):
    """Shared by this module (_flags = 0) and alt_gaussian_rasterization._C (LODGS_RASTER_* flags)."""
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    dev = means3D.device
    if dev.type != "cuda":
        raise RuntimeError("lodgs rasterizer needs CUDA tensors (there is no CPU path)")
    P = means3D.size(0) if indices.numel() == 0 else indices.size(0)
    H, W = int(image_height), int(image_width)
    # the blend writes every pixel and the preprocess every radius, so these are not pre-filled; they are zeroed below only
    # when nothing is rendered (the reference returns its zero-filled tensors in that case, rasterizer_impl.cu:332-333)
    out_color = torch.empty((NUM_CHANNELS, H, W), dtype=torch.float32, device=dev)
    out_invdepth = torch.empty((1 if do_depth else 0, H, W), dtype=torch.float32, device=dev)
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    seen = torch.zeros((P,), dtype=torch.int32, device=dev)
    geom = torch.empty(0, dtype=torch.uint8, device=dev)
    binning = torch.empty(0, dtype=torch.uint8, device=dev)
    img = torch.empty(0, dtype=torch.uint8, device=dev)
    rendered = 0
    if P != 0:
        M = sh.size(1) if sh.numel() != 0 else 0
        t = [_f(x.to(dev)) if x.numel() and x.device != dev else _f(x) for x in
             (background, indices, parent_indices, ts, kids, means3D, colors, opacity, scales, rotations, cov3D_precomp,
              viewmatrix, projmatrix, sh, campos)]
        (background, indices, parent_indices, ts, kids, means3D, colors, opacity, scales, rotations, cov3D_precomp,
         viewmatrix, projmatrix, sh, campos) = t
        stream = current_stream()
        geom = _byte_buffer(lib.lodgs_geom_bytes(P), dev)
        img = _byte_buffer(lib.lodgs_image_bytes(W, H), dev)
        n = C.c_int32(0)
        # with both ts and kids the preprocess kernel stores t and 1/kids into the blend records itself; stage 2 then gets kids = None
        in_records = LOD_IN_RECORDS and ts.numel() != 0 and kids.numel() != 0
        check(lib.lodgs_forward_preprocess_lod(P, int(degree), M, W, H, ptr(indices), ptr(parent_indices), ptr(ts),
                                               ptr(kids) if in_records else None, ptr(means3D),
                                               ptr(sh), ptr(colors), ptr(opacity), ptr(scales), float(scale_modifier),
                                               ptr(rotations), ptr(cov3D_precomp), ptr(viewmatrix), ptr(projmatrix), ptr(campos),
                                               float(tan_fovx), float(tan_fovy), ptr(radii), geom.data_ptr(), geom.numel(),
                                               C.addressof(n), int(_flags), stream))
        rendered = int(n.value)
        if rendered > 0:
            binning = _byte_buffer(lib.lodgs_binning_bytes(rendered), dev)
            check(lib.lodgs_forward_render(P, W, H, rendered, ptr(background), ptr(ts), None if in_records else ptr(kids), ptr(colors), ptr(radii),
                                           geom.data_ptr(), binning.data_ptr(), binning.numel(), img.data_ptr(), img.numel(),
                                           ptr(out_color), ptr(out_invdepth) if do_depth else None, ptr(seen), stream))
        if debug:
            torch.cuda.synchronize()
    if rendered == 0:
        out_color.zero_()
        out_invdepth.zero_()
        if P == 0:
            radii.zero_()
    return rendered, out_color, radii, geom, binning, img, out_invdepth, seen


def rasterize_gaussians_backward(background, indices, parent_indices, ts, kids, means3D, radii, colors, opacities, scales,
                                 rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                                 dL_dout_color, dL_dout_invdepth, sh, degree, campos, geomBuffer, R, binningBuffer,
                                 imageBuffer, debug):
    return backward_impl(background, indices, parent_indices, ts, kids, means3D, radii, colors, opacities, scales, rotations,
                         scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, dL_dout_invdepth, sh,
                         degree, campos, geomBuffer, R, binningBuffer, imageBuffer, debug, 0)


def rasterize_gaussians_backward_rgb(background, indices, parent_indices, ts, kids, means3D, radii, colors, opacities, scales,
                                     rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                                     dL_dout_color, dL_dout_invdepth, sh, degree, campos, geomBuffer, R, binningBuffer,
                                     imageBuffer, debug):
    """rasterize_gaussians_backward without the SH gradient: the sixth result is None and a ninth, dL_drgb [fullP,3] (colour gradient
    after the clamp mask), is appended — the input of lodgs.optim.sh_adam_from_rgb (not in the reference: lodgs_backward_rgb)."""
    return backward_impl(background, indices, parent_indices, ts, kids, means3D, radii, colors, opacities, scales, rotations,
                         scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, dL_dout_invdepth, sh,
                         degree, campos, geomBuffer, R, binningBuffer, imageBuffer, debug, 0, want_rgb=True)


def backward_impl(background, indices, parent_indices, ts, kids, means3D, radii, colors, opacities, scales, rotations, scale_modifier,
                  cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, dL_dout_invdepth, sh, degree, campos,
                  geomBuffer, R, binningBuffer, imageBuffer, debug, _flags, want_rgb=False):
    dev = means3D.device
    fullP = means3D.size(0)
    P = fullP if indices.numel() == 0 else indices.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if sh.numel() != 0 else 0
    has_depth = dL_dout_invdepth is not None and dL_dout_invdepth.numel() != 0
    # The five tensors the blend accumulates into with atomics are carved from ONE zero-filled allocation (one fill
    # kernel instead of five); the reference allocates each with torch::zeros (rasterize_points.cu:178-187).
    acc = torch.zeros((fullP * (3 + NUM_CHANNELS + 4 + 1 + (1 if has_depth else 0)),), dtype=torch.float32, device=dev)
    o = 0

    def carve(*shape):
        nonlocal o
        n = 1
        for d in shape:
            n *= d
        v = acc[o:o + n].view(*shape)
        o += n
        return v

    dL_dconic = carve(fullP, 2, 2)  # first: keeps the 16-byte rows aligned
    dL_dmeans2D, dL_dcolors, dL_dopacity = carve(fullP, 3), carve(fullP, NUM_CHANNELS), carve(fullP, 1)
    dL_dinvdepths = carve(fullP, 1) if has_depth else torch.zeros((0, 1), dtype=torch.float32, device=dev)
    # Tensors the library stores every row of need no zero-fill.
    own_geom = P != 0 and bool(lib.lodgs_backward_overwrites_geometry(int(indices.numel() != 0), int(parent_indices.numel() != 0)))
    alloc = (lambda *shape: torch.empty(shape, dtype=torch.float32, device=dev)) if own_geom else \
        (lambda *shape: torch.zeros(shape, dtype=torch.float32, device=dev))
    z = lambda *shape: torch.zeros(shape, dtype=torch.float32, device=dev)  # noqa: E731
    dL_dmeans3D, dL_dcov3D = alloc(fullP, 3), alloc(fullP, 6)
    has_scales = scales.numel() != 0
    dL_dscales = alloc(fullP, 3) if has_scales else z(fullP, 3)
    dL_drotations = alloc(fullP, 4) if has_scales else z(fullP, 4)
    dL_drgb = None
    if want_rgb:
        dL_dsh = None
        dL_drgb = alloc(fullP, NUM_CHANNELS)
    else:
        dL_dsh = torch.empty((fullP, M, 3), dtype=torch.float32, device=dev)
        if not (P != 0 and sh.numel() and sh.device == dev and sh.is_contiguous() and
                lib.lodgs_backward_overwrites_dsh(M, int(indices.numel() != 0), int(parent_indices.numel() != 0), ptr(sh), ptr(dL_dsh))):
            dL_dsh.zero_()
    if P != 0:
        t = [_f(x.to(dev)) if x.numel() and x.device != dev else _f(x) for x in
             (background, indices, parent_indices, ts, kids, means3D, colors, opacities, scales, rotations, cov3D_precomp,
              viewmatrix, projmatrix, sh, campos)]
        (background, indices, parent_indices, ts, kids, means3D, colors, opacities, scales, rotations, cov3D_precomp,
         viewmatrix, projmatrix, sh, campos) = t
        dL_dout_color = dL_dout_color.contiguous()
        if has_depth:
            dL_dout_invdepth = dL_dout_invdepth.contiguous()
        check(lib.lodgs_backward_rgb(P, fullP, int(degree), M, int(R), W, H, ptr(background), ptr(indices), ptr(parent_indices),
                                 ptr(ts), ptr(kids), ptr(means3D), ptr(sh), ptr(colors), ptr(opacities), ptr(scales),
                                 float(scale_modifier), ptr(rotations), ptr(cov3D_precomp), ptr(viewmatrix), ptr(projmatrix),
                                 ptr(campos), float(tan_fovx), float(tan_fovy), ptr(radii), ptr(geomBuffer), ptr(binningBuffer),
                                 ptr(imageBuffer), ptr(dL_dout_color), ptr(dL_dout_invdepth) if has_depth else None,
                                 ptr(dL_dmeans2D), ptr(dL_dconic), ptr(dL_dopacity), ptr(dL_dcolors),
                                 ptr(dL_dinvdepths) if has_depth else None, ptr(dL_dmeans3D), ptr(dL_dcov3D), ptr(dL_dsh),
                                 ptr(dL_dscales), ptr(dL_drotations), ptr(dL_drgb), int(_flags), current_stream()))
        if debug:
            torch.cuda.synchronize()
    if want_rgb:
        return dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, None, dL_dscales, dL_drotations, dL_drgb
    return dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations


def mark_visible(means3D, viewmatrix, projmatrix):
    P = means3D.size(0)
    present = torch.zeros((P,), dtype=torch.bool, device=means3D.device)
    if P != 0:
        check(lib.lodgs_mark_visible(P, ptr(means3D.contiguous()), ptr(viewmatrix.contiguous()), ptr(projmatrix.contiguous()),
                                     present.data_ptr(), current_stream()))
    return present


def compute_relocation(opacity_old, scale_old, N, binoms, n_max):
    """HR/rasterize_points.cu ComputeRelocationCUDA: returns (final_opacity [P], final_scale [3P]) like the reference."""
    P = opacity_old.size(0)
    dev = opacity_old.device
    if dev.type != "cuda":
        raise RuntimeError("lodgs compute_relocation needs CUDA tensors (there is no CPU path)")
    final_opacity = torch.zeros((P,), dtype=torch.float32, device=dev)
    final_scale = torch.zeros((3 * P,), dtype=torch.float32, device=dev)
    if P != 0:
        check(lib.lodgs_compute_relocation(P, ptr(opacity_old.contiguous().float()), ptr(scale_old.contiguous().float()),
                                           ptr(N.to(dev).int().contiguous()), ptr(binoms.to(dev).contiguous().float()), int(n_max),
                                           ptr(final_opacity), ptr(final_scale), current_stream()))
    return final_opacity, final_scale
