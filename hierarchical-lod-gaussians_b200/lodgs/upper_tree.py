"""Host side of `lodgs_upper_tree_cut` (csrc/cut_upper.cu): the per-frame cut of the upper tree that precedes the SPT cut, with the
argument meaning of the reference's Python (scene/gaussian_model.py:55-103,354-394; train_post.py:329-356; hierarchy_viewer.py:262-345)."""
import ctypes as C

import torch

from . import check, current_stream, lib, ptr

_ws = {}


def extract_frustum_planes(view_proj_matrix):
    """GaussianModel.extract_frustum_planes (scene/gaussian_model.py:55-78): left / right / bottom / top planes of a (4,4) full projection
    transform, normalised; (4,4) tensor of (nx, ny, nz, d)."""
    m = view_proj_matrix.T
    planes = torch.stack([m[3] + m[0], m[3] - m[0], m[3] + m[1], m[3] - m[1]])
    planes = planes / torch.norm(planes[:, :3], dim=1, keepdim=True)
    return planes


def coarse_cut(upper_tree_nodes, upper_tree_xyz, min_distance_squared, camera_position, bounds=None, planes=None, distance_multiplier=1.0,
               root_node=0):
    """One launch for `cut_hierarchy_on_condition(upper_tree_nodes, LOD_detail_cut, return_upper_tree=False, root_node=..., leave_out_of_cut_
    condition=cull)` and the lists train_post.py derives from its result. `bounds` None = no frustum culling (Use_Frustum_Culling off).
    -> dict: coarse_cut (node ids), SPT_indices, SPT_node_indices, SPT_distances, upper_tree_nodes_to_render — CUDA tensors, the
    reference's element order."""
    dev = upper_tree_nodes.device
    if dev.type != "cuda":
        raise RuntimeError("coarse_cut needs CUDA tensors (there is no CPU path)")
    nodes = upper_tree_nodes.to(torch.int32).contiguous()
    N = nodes.size(0)
    xyz = upper_tree_xyz.to(torch.float32).contiguous()
    md = min_distance_squared.to(torch.float32).contiguous()
    cam = camera_position.to(device=dev, dtype=torch.float32).contiguous()
    b = pl = None
    n_planes = 0
    if bounds is not None:
        b = bounds.to(torch.float32).contiguous()
        pl = planes.to(device=dev, dtype=torch.float32).contiguous()
        n_planes = pl.size(0)
    need = int(lib.lodgs_upper_cut_workspace_bytes(N))
    ws = _ws.get(dev.index)
    if ws is None or ws.numel() < need:
        ws = torch.empty(need, dtype=torch.uint8, device=dev)
        _ws[dev.index] = ws
    cap = max(N, 1)
    cut, spt_i, spt_n, direct = (torch.empty(cap, dtype=torch.int32, device=dev) for _ in range(4))
    spt_d = torch.empty(cap, dtype=torch.float32, device=dev)
    counts = (C.c_int32 * 3)()
    check(lib.lodgs_upper_tree_cut(ptr(nodes), N, int(root_node), ptr(xyz), ptr(b), ptr(pl), n_planes, ptr(md), ptr(cam), float(distance_multiplier),
                                   cut.data_ptr(), cap, spt_i.data_ptr(), spt_n.data_ptr(), spt_d.data_ptr(), direct.data_ptr(), ws.data_ptr(),
                                   ws.numel(), counts, current_stream()))
    k, s, d = counts[0], counts[1], counts[2]
    return {"coarse_cut": cut[:k], "SPT_indices": spt_i[:s], "SPT_node_indices": spt_n[:s], "SPT_distances": spt_d[:s],
            "upper_tree_nodes_to_render": direct[:d]}
