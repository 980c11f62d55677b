"""Render façade with the reference's signatures (gaussian_renderer/__init__.py:35-161 `render`, :241-398 `render_post`).

`render()` is the training-time entry point (train_single.py:97): it builds GaussianRasterizationSettings, optionally
gathers by `indices`, rasterizes, applies the exposure affine and clamps. `render_post()` is the hierarchy entry point
(render_hierarchy.py:82): with interp_python=True it does the parent interpolation in torch like the reference; with
interp_python=False the per-Gaussian render/parent indices, interpolation weights and sibling counts go to the rasterizer,
which interpolates inside preprocess and blends with the LOD alpha (the path BASELINE.json's north star names).
`pc` / `gaussians` is any object exposing the reference GaussianModel getters (scene/gaussian_model.py)."""
import math

import torch

from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer


def _empty(dtype):
    return torch.empty(0, dtype=dtype, device="cuda")


def _shs_or_colors(pc, viewpoint_camera, pipe, override_color):
    if override_color is not None:
        return None, override_color
    if getattr(pipe, "convert_SHs_python", False):
        raise NotImplementedError("convert_SHs_python: SH evaluation happens inside the preprocess kernel; pass override_color "
                                  "for precomputed colours")
    return pc.get_features, None


def render(viewpoint_camera, pc, pipe, bg_color: torch.Tensor, scaling_modifier=1.0, override_color=None, indices=None,
           use_trained_exp=False):
    """Render the scene. Background tensor (bg_color) must be on GPU!"""
    screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True, device="cuda") + 0
    try:
        screenspace_points.retain_grad()
    except Exception:  # noqa: BLE001
        pass
    tanfovx = math.tan(viewpoint_camera.FoVx * 0.5)
    tanfovy = math.tan(viewpoint_camera.FoVy * 0.5)
    raster_settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width), tanfovx=tanfovx,
        tanfovy=tanfovy, bg=bg_color, scale_modifier=scaling_modifier, viewmatrix=viewpoint_camera.world_view_transform,
        projmatrix=viewpoint_camera.full_proj_transform, sh_degree=pc.active_sh_degree, campos=viewpoint_camera.camera_center,
        prefiltered=False, debug=pipe.debug, do_depth=True, render_indices=_empty(torch.int32), parent_indices=_empty(torch.int32),
        interpolation_weights=_empty(torch.float32), num_node_kids=_empty(torch.int32))
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)
    means3D, means2D, opacity = pc.get_xyz, screenspace_points, pc.get_opacity
    scales = rotations = cov3D_precomp = None
    if getattr(pipe, "compute_cov3D_python", False):
        cov3D_precomp = pc.get_covariance(scaling_modifier)
    else:
        scales, rotations = pc.get_scaling, pc.get_rotation
    shs, colors_precomp = _shs_or_colors(pc, viewpoint_camera, pipe, override_color)
    if indices is not None:
        means3D = means3D[indices].contiguous()
        means2D = means2D[indices].contiguous()
        if shs is not None:
            shs = shs[indices].contiguous()
        opacity = opacity[indices].contiguous()
        if scales is not None:
            scales = scales[indices].contiguous()
            rotations = rotations[indices].contiguous()
    rendered_image, radii, depth_image = rasterizer(means3D=means3D, means2D=means2D, shs=shs, colors_precomp=colors_precomp,
                                                    opacities=opacity, scales=scales, rotations=rotations, cov3D_precomp=cov3D_precomp)
    if use_trained_exp:
        exposure = pc.get_exposure_from_name(viewpoint_camera.image_name)
        rendered_image = torch.matmul(rendered_image.permute(1, 2, 0), exposure[:3, :3]).permute(2, 0, 1) + exposure[:3, 3, None, None]
    rendered_image = rendered_image.clamp(0, 1)
    subfilter = radii > 0
    if indices is not None:
        vis_filter = torch.zeros(pc._xyz.size(0), dtype=bool, device="cuda")
        w = vis_filter[indices]
        w[subfilter] = True
        vis_filter[indices] = w
    else:
        vis_filter = subfilter
    return {"render": rendered_image, "depth": depth_image, "viewspace_points": screenspace_points,
            "visibility_filter": vis_filter.nonzero().flatten().long(), "radii": radii[subfilter]}


def render_post(viewpoint_camera, gaussians, pipe, bg_color: torch.Tensor, scaling_modifier=1.0, override_color=None,
                render_indices=torch.Tensor([]).int(), parent_indices=torch.Tensor([]).int(),
                interpolation_weights=torch.Tensor([]).float(), num_node_siblings=torch.Tensor([]).int(), interp_python=True,
                on_disk=True, use_trained_exp=False, iteration=0):
    """Render the scene from a hierarchy cut. Background tensor (bg_color) must be on GPU!"""
    tanfovx = math.tan(viewpoint_camera.FoVx * 0.5)
    tanfovy = math.tan(viewpoint_camera.FoVy * 0.5)
    screenspace_points = torch.zeros_like(gaussians.get_xyz, dtype=gaussians.get_xyz.dtype, requires_grad=True, device="cuda") + 0
    try:
        screenspace_points.retain_grad()
    except Exception:  # noqa: BLE001
        pass
    means3D, means2D, opacity = gaussians.get_xyz, screenspace_points, gaussians.get_opacity
    scales, rotations = gaussians.get_scaling, gaussians.get_rotation
    shs, colors_precomp = _shs_or_colors(gaussians, viewpoint_camera, pipe, override_color)
    skybox = int(getattr(gaussians, "skybox_points", 0))
    n = render_indices.size(0)
    if n != 0 and interp_python:
        # parent interpolation in torch, as the reference does (GR:304-337); the kernel then only applies the LOD alpha
        ri, pi = render_indices.long(), parent_indices[:n].long()
        t = interpolation_weights[:n].unsqueeze(1)
        ti = 1 - t
        means3D_b = (t * means3D[ri] + ti * means3D[pi]).contiguous()
        scales_b = (t * scales[ri] + ti * scales[pi]).contiguous()
        if shs is not None:
            shs_b = (t.unsqueeze(2) * shs[ri] + ti.unsqueeze(2) * shs[pi]).contiguous()
        parents, rots = rotations[pi], rotations[ri]
        dots = torch.bmm(rots.unsqueeze(1), parents.unsqueeze(2)).flatten()
        parents[dots < 0] *= -1
        rot_b = (t * rots + ti * parents).contiguous()
        op_b = (t * opacity[ri] + ti * opacity[pi]).contiguous()
        sky = torch.arange(0, skybox, device="cuda").long()
        means3D = torch.cat((means3D[sky], means3D_b)).contiguous()
        if shs is not None:
            shs = torch.cat((shs[sky], shs_b)).contiguous()
        opacity = torch.cat((opacity[sky], op_b)).contiguous()
        rotations = torch.cat((rotations[sky], rot_b)).contiguous()
        scales = torch.cat((scales[sky], scales_b)).contiguous()
        means2D = means2D[:(skybox + n)].contiguous()
        ones_f = torch.ones(skybox, device="cuda", dtype=torch.float32)
        ones_i = torch.ones(skybox, device="cuda", dtype=torch.int32)
        interpolation_weights = torch.cat((ones_f, interpolation_weights[:n].detach().float()))
        num_node_siblings = torch.cat((ones_i, num_node_siblings[:n].int()))
        render_indices = parent_indices = _empty(torch.int32)
        # as the reference (GR:348-350): indices and weights are cleared, only num_node_kids reaches the kernel, whose
        # LOD alpha blend is therefore off (ts == nullptr, forward.cu:505)
        lod_ts, lod_kids = _empty(torch.float32), num_node_siblings.contiguous()
    elif n != 0:
        render_indices = render_indices[:n].int().contiguous()
        parent_indices = parent_indices[:n].int().contiguous()
        lod_ts = interpolation_weights[:n].float().contiguous()
        lod_kids = num_node_siblings[:n].int().contiguous()
    else:
        render_indices = parent_indices = _empty(torch.int32)
        lod_ts, lod_kids = _empty(torch.float32), _empty(torch.int32)
    raster_settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width), tanfovx=tanfovx,
        tanfovy=tanfovy, bg=bg_color, scale_modifier=scaling_modifier, viewmatrix=viewpoint_camera.world_view_transform,
        projmatrix=viewpoint_camera.full_proj_transform, sh_degree=gaussians.active_sh_degree,
        campos=viewpoint_camera.camera_center, prefiltered=False, debug=getattr(pipe, "debug", False),
        render_indices=render_indices, parent_indices=parent_indices, interpolation_weights=lod_ts, num_node_kids=lod_kids,
        do_depth=False)
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)
    rendered_image, radii, _ = rasterizer(means3D=means3D, means2D=means2D, shs=shs, colors_precomp=colors_precomp, opacities=opacity,
                                          scales=scales, rotations=rotations, cov3D_precomp=None)
    rendered_image = rendered_image.clamp(0, 1)
    radii = radii[skybox:]
    vis_filter = radii > 0
    return {"render": rendered_image, "viewspace_points": screenspace_points, "visibility_filter": vis_filter,
            "radii": radii[vis_filter]}
