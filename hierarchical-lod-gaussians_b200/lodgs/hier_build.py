"""GPU construction of the Gaussian hierarchy: the host side of `lodgs_hier_build` (csrc/hier_build.cu), mirroring what the reference's
CPU creator does around its three stages (GH/mainHierarchyCreator.cpp:41-184): PLY records -> activated Gaussians (GH/loader.cpp:78-126),
validity filter (mainHierarchyCreator.cpp:87-149), kd-tree + merge + alignment on the device, `.dhier` output (GH/writer.cpp:266-288).
No CPU fallback: tensors must be CUDA."""
import numpy as np
import torch

from . import check, current_stream, lib


def build_hierarchy(pos, scales, rotations, opacities, shs, dynamic=True, static=True):
    """pos[L,3], scales[L,3] (activated), rotations[L,4] (w,x,y,z, normalised), opacities[L] (activated), shs[L,F] float32 CUDA tensors ->
    dict with one Gaussian per node in pre-order (`pos`, `rotations`, `scales`, `opacities`, `shs`, N = 2L-1 rows) and the node tables
    `dyn_nodes`[N,6] (HierarchyNode) and / or `nodes`[N,7] + `boxes`[N,2,4] (static format)."""
    if pos.device.type != "cuda":
        raise RuntimeError("build_hierarchy needs CUDA tensors (there is no CPU path)")
    dev = pos.device
    f = lambda t: t.detach().to(device=dev, dtype=torch.float32).contiguous()  # noqa: E731
    pos, scales, rotations, opacities = f(pos), f(scales), f(rotations), f(opacities).reshape(-1)
    L = pos.size(0)
    shs = f(shs).reshape(L, -1)
    F = shs.size(1)
    N = 2 * L - 1
    out = {"pos": torch.empty(N, 3, device=dev), "rotations": torch.empty(N, 4, device=dev), "scales": torch.empty(N, 3, device=dev),
           "opacities": torch.empty(N, device=dev), "shs": torch.empty(N, F, device=dev)}
    if dynamic:
        out["dyn_nodes"] = torch.empty(N, 6, dtype=torch.int32, device=dev)
    if static:
        out["nodes"] = torch.empty(N, 7, dtype=torch.int32, device=dev)
        out["boxes"] = torch.empty(N, 2, 4, device=dev)
    ws = torch.empty(lib.lodgs_hier_build_workspace_bytes(L), dtype=torch.uint8, device=dev)
    p = lambda k: out[k].data_ptr() if k in out else None  # noqa: E731
    check(lib.lodgs_hier_build(L, F, pos.data_ptr(), scales.data_ptr(), rotations.data_ptr(), opacities.data_ptr(),
                               shs.data_ptr() if F else None, p("pos"), p("rotations"), p("scales"), p("opacities"), p("shs") if F else None,
                               p("dyn_nodes"), p("nodes"), p("boxes"), ws.data_ptr(), ws.numel(), current_stream()))
    return out


def read_less_rich_ply(path):
    """The binary PLY the creator reads with loadPly<LessRichPoint> (GH/loader.cpp:78-126; record = GH/common.h:39-47, 26 f32) ->
    numpy [count, 26]."""
    with open(path, "rb") as fh:
        fh.readline()
        fh.readline()
        count = int(fh.readline().split()[2])
        while fh.readline().strip() != b"end_header":
            pass
        return np.frombuffer(fh.read(count * 26 * 4), dtype="<f4").reshape(count, 26).copy()


def activate_and_filter(records, device, skyboxpoints=0):
    """records[count,26] (position 3, normal 3, shs 12, opacity logit 1, log-scale 3, rotation 4) -> (pos, scales, rotations, opacities,
    shs[.,48]) on `device` of the Gaussians the creator keeps: loader activation (GH/loader.cpp:105-124: sigmoid, exp, normalise; a
    LessRichPoint only carries its three DC coefficients over) and the validity filter of mainHierarchyCreator.cpp:87-149, evaluated
    on the host by `lodgs_creator_activate` with the creator's libm calls, so the build starts from the creator's bits."""
    r = np.ascontiguousarray(records[skyboxpoints:], dtype=np.float32)
    n = r.shape[0]
    pos, scales, rot = np.empty((n, 3), np.float32), np.empty((n, 3), np.float32), np.empty((n, 4), np.float32)
    opac, shs, keep = np.empty(n, np.float32), np.empty((n, 48), np.float32), np.empty(n, np.uint8)
    P = lambda a: a.ctypes.data  # noqa: E731
    check(lib.lodgs_creator_activate(P(r), n, P(pos), P(scales), P(rot), P(opac), P(shs), P(keep)))
    k = keep.astype(bool)
    return tuple(torch.from_numpy(a[k]).to(device) for a in (pos, scales, rot, opac, shs))


def log_scales_like_writer(scales):
    """log(scale) as GH/writer.cpp:119,130 stores it (libm logf on the host) -> CPU tensor."""
    s = np.ascontiguousarray(scales.detach().cpu().numpy(), dtype=np.float32)
    out = np.empty_like(s)
    check(lib.lodgs_logf_host(s.ctypes.data, s.size, out.ctypes.data))
    return torch.from_numpy(out)


def create_dynamic_hierarchy(ply_path, out_prefix, device="cuda"):
    """GaussianHierarchyCreator <ply> <source dir> <output prefix> (GH/mainHierarchyCreator.cpp) on the GPU: writes
    <out_prefix>hierarchy.dhier with the reference's layout (log-scales, activated opacities, SH degree 3) and returns the build."""
    from gaussian_hierarchy import _C as gh
    h = build_hierarchy(*activate_and_filter(read_less_rich_ply(ply_path), device), dynamic=True, static=False)
    gh.write_dynamic_hierarchy(out_prefix + "hierarchy.dhier", h["pos"], h["shs"], h["opacities"], log_scales_like_writer(h["scales"]),
                               h["rotations"], h["dyn_nodes"], 3)
    return h
