"""Render façade with the reference's signatures (gaussian_renderer/__init__.py): `render` (:35-161, training entry point of
train_single.py:97), `render_post` (:241-402, hierarchy cuts, render_hierarchy.py:82), `render_coarse` (:405-516), `render_on_disk`
(:163-236), `render_vanilla` (:621-712, the alt-rasterizer surface train_post.py:510 drives), `occlusion_cull` (:24-33) and a
`network_gui` placeholder module so that `from gaussian_renderer import ..., network_gui` of the reference scripts resolves.

`render_post` keeps the reference's behaviour for both values of `interp_python`: the Gaussians are gathered (and with
interp_python=True lerped with their parents) in torch, and render / parent indices and weights are CLEARED before the rasterizer call
(GR:348-350), so the kernel neither interpolates nor applies the LOD alpha. The in-kernel LOD path that BASELINE.json's north star
names (indices, parents, t and sibling counts handed to the rasterizer) is an explicit extra: `interp_kernel=True`.
`pc` / `gaussians` is any object exposing the reference GaussianModel getters (scene/gaussian_model.py).
`render_stp` is not served: its rasterizer (stp_gaussian_rasterization) is not part of the reference tree (GR:17, commented out)."""
import math

import torch

import alt_gaussian_rasterization
from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

from . import network_gui  # noqa: F401


def _empty(dtype):
    return torch.empty(0, dtype=dtype, device="cuda")


_SH_C = (0.28209479177387814, 0.4886025119029199,
         (1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396),
         (-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658, 1.445305721320277,
          -0.5900435899266435))


def _sh_basis(deg, d):
    """Real SH basis values [P, (deg+1)^2] at unit directions d [P,3] (the polynomials of utils/sh_utils.py:57-112)."""
    x, y, z = d[:, 0], d[:, 1], d[:, 2]
    cols = [torch.full_like(x, _SH_C[0])]
    if deg > 0:
        cols += [-_SH_C[1] * y, _SH_C[1] * z, -_SH_C[1] * x]
    if deg > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        c = _SH_C[2]
        cols += [c[0] * xy, c[1] * yz, c[2] * (2.0 * zz - xx - yy), c[3] * xz, c[4] * (xx - yy)]
    if deg > 2:
        c = _SH_C[3]
        cols += [c[0] * y * (3 * xx - yy), c[1] * xy * z, c[2] * y * (4 * zz - xx - yy), c[3] * z * (2 * zz - 3 * xx - 3 * yy),
                 c[4] * x * (4 * zz - xx - yy), c[5] * z * (xx - yy), c[6] * x * (xx - 3 * yy)]
    return torch.stack(cols, 1)


def _colors_from_shs_python(pc, viewpoint_camera):
    """pipe.convert_SHs_python (GR:114-119): SH -> RGB in torch, + 0.5, clamped at 0."""
    feats = pc.get_features                                   # [P, M, 3]
    d = pc.get_xyz - viewpoint_camera.camera_center.reshape(1, 3)
    d = d / d.norm(dim=1, keepdim=True)
    basis = _sh_basis(pc.active_sh_degree, d)                 # [P, m]
    rgb = (basis.unsqueeze(2) * feats[:, :basis.size(1)]).sum(1)
    return torch.clamp_min(rgb + 0.5, 0.0)


def _shs_or_colors(pc, viewpoint_camera, pipe, override_color):
    if override_color is not None:
        return None, override_color
    if getattr(pipe, "convert_SHs_python", False):
        return None, _colors_from_shs_python(pc, viewpoint_camera)
    return pc.get_features, None


def _scales_rot_or_cov(pc, pipe, scaling_modifier):
    if getattr(pipe, "compute_cov3D_python", False):
        return None, None, pc.get_covariance(scaling_modifier)
    return pc.get_scaling, pc.get_rotation, None


def _screenspace(ref):
    pts = torch.zeros_like(ref, dtype=ref.dtype, requires_grad=True, device="cuda") + 0
    try:
        pts.retain_grad()
    except Exception:  # noqa: BLE001
        pass
    return pts


def _settings(viewpoint_camera, bg_color, scaling_modifier, sh_degree, debug, do_depth, render_indices=None, parent_indices=None,
              ts=None, kids=None):
    return GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
        tanfovx=math.tan(viewpoint_camera.FoVx * 0.5), tanfovy=math.tan(viewpoint_camera.FoVy * 0.5), bg=bg_color,
        scale_modifier=scaling_modifier, viewmatrix=viewpoint_camera.world_view_transform,
        projmatrix=viewpoint_camera.full_proj_transform, sh_degree=sh_degree, campos=viewpoint_camera.camera_center,
        prefiltered=False, debug=debug, do_depth=do_depth,
        render_indices=_empty(torch.int32) if render_indices is None else render_indices,
        parent_indices=_empty(torch.int32) if parent_indices is None else parent_indices,
        interpolation_weights=_empty(torch.float32) if ts is None else ts,
        num_node_kids=_empty(torch.int32) if kids is None else kids)


def _visibility(pc, radii, indices):
    subfilter = radii > 0
    if indices is None:
        return subfilter, subfilter
    vis_filter = torch.zeros(pc._xyz.size(0), dtype=bool, device="cuda")
    w = vis_filter[indices]
    w[subfilter] = True
    vis_filter[indices] = w
    return vis_filter, subfilter


def render(viewpoint_camera, pc, pipe, bg_color: torch.Tensor, scaling_modifier=1.0, override_color=None, indices=None,
           use_trained_exp=False):
    """Render the scene (GR:35-161). Background tensor (bg_color) must be on GPU!"""
    screenspace_points = _screenspace(pc.get_xyz)
    rasterizer = GaussianRasterizer(raster_settings=_settings(viewpoint_camera, bg_color, scaling_modifier, pc.active_sh_degree,
                                                              pipe.debug, True))
    means3D, means2D, opacity = pc.get_xyz, screenspace_points, pc.get_opacity
    scales, rotations, cov3D_precomp = _scales_rot_or_cov(pc, pipe, scaling_modifier)
    shs, colors_precomp = _shs_or_colors(pc, viewpoint_camera, pipe, override_color)
    if indices is not None:
        means3D = means3D[indices].contiguous()
        means2D = means2D[indices].contiguous()
        shs = shs[indices].contiguous()          # like the reference (GR:123): the index path assumes SHs and scale / rotation
        opacity = opacity[indices].contiguous()
        scales = scales[indices].contiguous()
        rotations = rotations[indices].contiguous()
    rendered_image, radii, depth_image = rasterizer(means3D=means3D, means2D=means2D, shs=shs, colors_precomp=colors_precomp,
                                                    opacities=opacity, scales=scales, rotations=rotations, cov3D_precomp=cov3D_precomp)
    if use_trained_exp:
        exposure = pc.get_exposure_from_name(viewpoint_camera.image_name)
        rendered_image = torch.matmul(rendered_image.permute(1, 2, 0), exposure[:3, :3]).permute(2, 0, 1) + exposure[:3, 3, None, None]
    rendered_image = rendered_image.clamp(0, 1)
    vis_filter, subfilter = _visibility(pc, radii, indices)
    return {"render": rendered_image, "depth": depth_image, "viewspace_points": screenspace_points,
            "visibility_filter": vis_filter.nonzero().flatten().long(), "radii": radii[subfilter]}


def render_coarse(viewpoint_camera, pc, pipe, bg_color: torch.Tensor, scaling_modifier=1.0, zfar=0.0, override_color=None, indices=None):
    """Render the scene for the coarse optimisation (GR:405-516): `render` without exposure and depth; debug is forced on and the
    visibility filter is returned as a boolean mask. Background tensor (bg_color) must be on GPU!"""
    screenspace_points = _screenspace(pc.get_xyz)
    rasterizer = GaussianRasterizer(raster_settings=_settings(viewpoint_camera, bg_color, scaling_modifier, pc.active_sh_degree, True, False))
    means3D, means2D, opacity = pc.get_xyz, screenspace_points, pc.get_opacity
    scales, rotations, cov3D_precomp = _scales_rot_or_cov(pc, pipe, scaling_modifier)
    shs, colors_precomp = _shs_or_colors(pc, viewpoint_camera, pipe, override_color)
    if indices is not None:
        means3D = means3D[indices].contiguous()
        means2D = means2D[indices].contiguous()
        shs = shs[indices].contiguous()
        opacity = opacity[indices].contiguous()
        scales = scales[indices].contiguous()
        rotations = rotations[indices].contiguous()
    rendered_image, radii, _ = rasterizer(means3D=means3D, means2D=means2D, shs=shs, colors_precomp=colors_precomp, opacities=opacity,
                                          scales=scales, rotations=rotations, cov3D_precomp=cov3D_precomp)
    vis_filter, subfilter = _visibility(pc, radii, indices)
    return {"render": rendered_image, "viewspace_points": screenspace_points, "visibility_filter": vis_filter, "radii": radii[subfilter]}


def render_on_disk(viewpoint_camera, means3D, opacity, scales, rotations, shs, pipe, bg_color: torch.Tensor, scaling_modifier=1.0,
                   override_color=None, sh_degree=3):
    """GR:163-236: renders explicitly passed (already activated) parameters and returns the rasterizer's `seen` flags, which
    occlusion_cull consumes. The vendored Python wrapper returns (color, radii, invdepths) while this caller unpacks (image, seen, depth)
    (SURVEY §2.3-1); here `seen` comes from the native 8-tuple via GaussianRasterizer.last_seen."""
    screenspace_points = torch.zeros_like(means3D, dtype=torch.float32, requires_grad=False, device="cuda") + 0
    means2D = torch.nn.Parameter(screenspace_points)
    pipe.debug = True
    rasterizer = GaussianRasterizer(raster_settings=_settings(viewpoint_camera, bg_color, scaling_modifier, int(sh_degree), pipe.debug, False))
    rendered_image, _radii, _depth = rasterizer(means3D=means3D, means2D=means2D, shs=shs, colors_precomp=None, opacities=opacity,
                                                scales=scales, rotations=rotations, cov3D_precomp=None)
    return {"render": rendered_image.clamp(0, 1), "viewspace_points": screenspace_points, "seen": rasterizer.last_seen}


def occlusion_cull(indices, gaussians, camera, pipe, background, opacity_multiplier=1, scale_multiplier=1):
    """GR:24-33: which of the Gaussians `indices` reach at least one pixel (bool, CPU)."""
    means3D = gaussians._xyz[indices].cuda().contiguous()
    opacity = torch.clamp(gaussians.opacity_activation(gaussians._opacity[indices].cuda().contiguous()) * opacity_multiplier, 0, 1)
    scales = gaussians.scaling_activation(gaussians._scaling[indices].cuda().contiguous()) * scale_multiplier
    rotations = gaussians.rotation_activation(gaussians._rotation[indices].cuda().contiguous())
    shs = torch.cat((gaussians._features_dc[indices].cuda().contiguous(), gaussians._features_rest[indices].cuda().contiguous()), dim=1).contiguous()
    return render_on_disk(camera, means3D, opacity, scales, rotations, shs, pipe, background)["seen"].to(torch.bool).cpu()


def render_vanilla(viewpoint_camera, means3D, opacity, scales, rotations, dc, shs, pipe, bg_color: torch.Tensor, sh_degree=3,
                   scaling_modifier=1.0, override_color=None, use_trained_exp=False, anti_aliasing=True):
    """GR:621-712: the alt-rasterizer surface on explicitly passed parameters (train_post.py:510, train_coarse.py:103). The vendored
    function builds the result dict and falls off its end without returning it; its callers index the result, so it is returned here."""
    raster_settings = alt_gaussian_rasterization.GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
        tanfovx=math.tan(viewpoint_camera.FoVx * 0.5), tanfovy=math.tan(viewpoint_camera.FoVy * 0.5), bg=bg_color, scale_modifier=1.0,
        viewmatrix=viewpoint_camera.world_view_transform, projmatrix=viewpoint_camera.full_proj_transform, sh_degree=sh_degree,
        campos=viewpoint_camera.camera_center, prefiltered=False, debug=False, antialiasing=anti_aliasing)
    rasterizer = alt_gaussian_rasterization.GaussianRasterizer(raster_settings=raster_settings)
    screenspace_points = _screenspace(means3D)
    if override_color is not None:
        shs = None
    rendered_image, radii, _ = rasterizer(means3D=means3D, means2D=screenspace_points, dc=dc, shs=shs, colors_precomp=override_color,
                                          opacities=opacity, scales=scales, rotations=rotations, cov3D_precomp=None)
    return {"render": rendered_image.clamp(0, 1), "viewspace_points": screenspace_points, "radii": radii}


def render_stp(*args, **kwargs):
    raise NotImplementedError("render_stp needs stp_gaussian_rasterization, which the reference does not vendor (its import is commented "
                              "out, gaussian_renderer/__init__.py:17)")


def render_post(viewpoint_camera, gaussians, pipe, bg_color: torch.Tensor, scaling_modifier=1.0, override_color=None,
                render_indices=torch.Tensor([]).int(), parent_indices=torch.Tensor([]).int(),
                interpolation_weights=torch.Tensor([]).float(), num_node_siblings=torch.Tensor([]).int(), interp_python=True,
                on_disk=True, use_trained_exp=False, iteration=0, interp_kernel=False):
    """Render the scene from a hierarchy cut (GR:241-402). Background tensor (bg_color) must be on GPU!

    interp_python=True : parent interpolation in torch (GR:306-337), skybox points prepended.
    interp_python=False: plain gather by render_indices (GR:338-345).
    In both cases the index / parent / weight tensors are cleared afterwards (GR:346-348) and only num_node_kids reaches the kernel, as
    in the reference. interp_kernel=True (not in the reference's signature) hands indices, parents, weights and sibling counts to the
    rasterizer instead: interpolation inside preprocess, LOD alpha in the blend, parent redirect in the backward."""
    screenspace_points = _screenspace(gaussians.get_xyz)
    means3D, means2D, opacity = gaussians.get_xyz, screenspace_points, gaussians.get_opacity
    scales, rotations, cov3D_precomp = _scales_rot_or_cov(gaussians, pipe, scaling_modifier)
    shs, colors_precomp = _shs_or_colors(gaussians, viewpoint_camera, pipe, override_color)
    skybox = int(getattr(gaussians, "skybox_points", 0))
    n = render_indices.size(0)
    ri = pi = ts = None
    kids = num_node_siblings
    if n != 0 and interp_kernel:
        ri = render_indices[:n].int().contiguous()
        pi = parent_indices[:n].int().contiguous()
        ts = interpolation_weights[:n].float().contiguous()
        kids = num_node_siblings[:n].int().contiguous()
    elif n != 0:
        render_inds = render_indices.long()
        if interp_python:
            interps = interpolation_weights[:n].unsqueeze(1)
            interps_inv = (1 - interpolation_weights[:n]).unsqueeze(1)
            parent_inds = parent_indices[:n].long()
            means3D_base = (interps * means3D[render_inds] + interps_inv * means3D[parent_inds]).contiguous()
            scales_base = (interps * scales[render_inds] + interps_inv * scales[parent_inds]).contiguous()
            if override_color is None:
                shs_base = (interps.unsqueeze(2) * shs[render_inds] + interps_inv.unsqueeze(2) * shs[parent_inds]).contiguous()
            parents, rots = rotations[parent_inds], rotations[render_inds]
            dots = torch.bmm(rots.unsqueeze(1), parents.unsqueeze(2)).flatten()
            parents[dots < 0] *= -1
            rotations_base = ((interps * rots) + interps_inv * parents).contiguous()
            opacity_base = (interps * opacity[render_inds] + interps_inv * opacity[parent_inds]).contiguous()
            skybox_inds = torch.arange(0, skybox, device="cuda").long()
            means3D = torch.cat((means3D[skybox_inds], means3D_base)).contiguous()
            if override_color is None:
                shs = torch.cat((shs[skybox_inds], shs_base)).contiguous()
            opacity = torch.cat((opacity[skybox_inds], opacity_base)).contiguous()
            rotations = torch.cat((rotations[skybox_inds], rotations_base)).contiguous()
            means2D = means2D[:(skybox + n)].contiguous()
            scales = torch.cat((scales[skybox_inds], scales_base)).contiguous()
            # skybox points are not interpolated; the reference slices with [:-skybox_points] (GR:333-334), i.e. it keeps the first
            # len - skybox entries (and nothing at all for skybox_points == 0: -0 == 0)
            kids = torch.cat((torch.ones(skybox, device="cuda", dtype=torch.int32),
                              num_node_siblings[:num_node_siblings.size(0) - skybox if skybox else 0].to(device="cuda", dtype=torch.int32)))
        else:
            means3D = means3D[render_inds].contiguous()
            means2D = means2D[render_inds].contiguous()
            if override_color is None:
                shs = shs[render_inds].contiguous()
            opacity = opacity[render_inds].contiguous()
            scales = scales[render_inds].contiguous()
            rotations = rotations[render_inds].contiguous()
    pipe.debug = True
    rasterizer = GaussianRasterizer(raster_settings=_settings(viewpoint_camera, bg_color, scaling_modifier, gaussians.active_sh_degree,
                                                              pipe.debug, False, ri, pi, ts, kids.int() if kids is not None else None))
    rendered_image, radii, _ = rasterizer(means3D=means3D, means2D=means2D, shs=shs, colors_precomp=colors_precomp, opacities=opacity,
                                          scales=scales, rotations=rotations, cov3D_precomp=cov3D_precomp)
    rendered_image = rendered_image.clamp(0, 1)
    radii = radii[skybox:]
    vis_filter = radii > 0
    return {"render": rendered_image, "viewspace_points": screenspace_points, "visibility_filter": vis_filter,
            "radii": radii[vis_filter]}
