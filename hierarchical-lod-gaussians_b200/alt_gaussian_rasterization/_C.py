"""alt_gaussian_rasterization._C — the reference's second rasterizer plugin (submodules/alt-rasterizer, "AR": config
kbuffer.json) served by the same liblodgs.so kernels: same functions, argument order and return tuples as AR/ext.cpp and
AR/rasterize_points.cu:44-306.

What differs from diff_gaussian_rasterization and how it is mapped:
* tile rectangle from the square of the largest 3-sigma radius, opacity compensation only with `antialiasing`
  -> LODGS_RASTER_SQUARE_RECT / LODGS_RASTER_NO_ANTIALIAS of lodgs_forward_preprocess_ex / lodgs_backward_ex;
* SH coefficients split into `dc [P,1,3]` and `sh [P,M,3]` -> concatenated to one [P,M+1,3] block for the kernels and the
  gradient split again (one extra copy each way; the kernels want the coefficients of a Gaussian contiguous);
* per-Gaussian backward over 32-splat buckets with a `sampleBuffer` of per-bucket pixel states (AR/cuda_rasterizer/
  rasterizer_impl.cu:288-295, backward.cu:451-656): the tile-reduced backward of csrc/blend_bwd.cu computes the same
  gradients without that state, so `sampleBuffer` is empty; `num_buckets` (which only sizes that state) is still the reference's
  value: the reference drops (Gaussian, tile) instances that cannot reach alpha 1/255 anywhere inside the tile before counting
  (rasterizer_impl.cu:143-187) — pairs the blend kernels here skip per pixel with the same result — and csrc/ar_buckets.cu
  evaluates that tile test on the sorted instance list just to report the same count;
* two quirks of AR's backward are reproduced on request of this module (LODGS_RASTER_BACKGROUND_TWICE,
  LODGS_RASTER_CLAMPED_ALPHA_GRAD, see include/lodgs.h): the background term of dL/dalpha counted twice, and no gradient
  cut for alphas clamped at 0.99;
* no LOD inputs, no `seen`, inverse depth always rendered.
"""
import torch

import lodgs

import lodgs as _lodgs
_lodgs.load()   # the plugin cannot work without the native library: fail at import, loudly
from diff_gaussian_rasterization import _C as _hr
from lodgs.optim import adam_update

NUM_CHANNELS = 3


def _flags(antialiasing):
    return (lodgs.RASTER_SQUARE_RECT | lodgs.RASTER_BACKGROUND_TWICE | lodgs.RASTER_CLAMPED_ALPHA_GRAD |
            (0 if antialiasing else lodgs.RASTER_NO_ANTIALIAS))


def _cat_sh(dc, sh):
    if dc.numel() == 0 and sh.numel() == 0:
        return sh
    if sh.numel() == 0:
        return dc.contiguous()
    return torch.cat((dc, sh.to(dc.device)), dim=1).contiguous()


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix,
                        tan_fovx, tan_fovy, image_height, image_width, dc, sh, degree, campos, prefiltered, antialiasing, debug):
    dev = means3D.device
    ei = torch.empty(0, dtype=torch.int32, device=dev)
    ef = torch.empty(0, dtype=torch.float32, device=dev)
    shs = _cat_sh(dc, sh)
    rendered, color, radii, geom, binning, img, invdepth, _seen = _hr.forward_impl(
        background, ei, ei, ef, ei, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix,
        tan_fovx, tan_fovy, image_height, image_width, shs, degree, campos, prefiltered, debug, True, _flags(antialiasing))
    num_buckets = 0
    if rendered > 0:
        # the reference's value: per-tile lists after ITS tile-level opacity cull, ceil(length / 32) summed (csrc/ar_buckets.cu)
        import ctypes as C
        ntiles = ((int(image_width) + 15) // 16) * ((int(image_height) + 15) // 16)
        ws = torch.empty(ntiles * 4 + 512, dtype=torch.uint8, device=dev)
        out2 = (C.c_int32 * 2)()
        lodgs.check(lodgs.lib.lodgs_ar_bucket_count(radii.numel(), int(image_width), int(image_height), int(rendered), geom.data_ptr(),
                                                    binning.data_ptr(), ws.data_ptr(), ws.numel(), out2, lodgs.current_stream()))
        num_buckets = int(out2[0])
    sample = torch.empty(0, dtype=torch.uint8, device=dev)
    return rendered, num_buckets, color, invdepth, radii, geom, binning, img, sample


def rasterize_gaussians_backward(background, means3D, radii, colors, opacities, scales, rotations, scale_modifier, cov3D_precomp,
                                 viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, dc, sh, dL_dout_invdepth, degree, campos,
                                 geomBuffer, R, binningBuffer, imageBuffer, B, sampleBuffer, antialiasing, debug):
    dev = means3D.device
    ei = torch.empty(0, dtype=torch.int32, device=dev)
    ef = torch.empty(0, dtype=torch.float32, device=dev)
    shs = _cat_sh(dc, sh)
    g2, gcol, gop, g3, gcov, gsh, gs, gr = _hr.backward_impl(
        background, ei, ei, ef, ei, means3D, radii, colors, opacities, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
        projmatrix, tan_fovx, tan_fovy, dL_dout_color, dL_dout_invdepth, shs, degree, campos, geomBuffer, R, binningBuffer, imageBuffer,
        debug, _flags(antialiasing))
    P = means3D.size(0)
    if shs.numel() == 0:
        gdc = torch.zeros((P, 1, 3), dtype=torch.float32, device=dev)
        grest = torch.zeros((P, 0, 3), dtype=torch.float32, device=dev)
    else:
        gdc = gsh[:, :1].contiguous()
        grest = gsh[:, 1:].contiguous()
    return g2, gcol, gop, g3, gcov, gdc, grest, gs, gr


def mark_visible(means3D, viewmatrix, projmatrix):
    return _hr.mark_visible(means3D, viewmatrix, projmatrix)


def compute_relocation(opacity_old, scale_old, N, binoms, n_max):
    return _hr.compute_relocation(opacity_old, scale_old, N, binoms, n_max)


def adamUpdate(param, param_grad, exp_avg, exp_avg_sq, visible, lr, b1, b2, eps, N, M):
    adam_update(param, param_grad, exp_avg, exp_avg_sq, visible, lr, b1, b2, eps, N, M)
