"""gaussian_hierarchy._C — expand_to_size / get_interpolation_weights / load_hierarchy / write_hierarchy with the
reference's signatures and tensor layouts (GH/ext.cpp:15-27, GH/torch/torch_interface.cpp), bound to liblodgs.so.

Device conventions of the reference are kept: expand_to_size takes `viewpoint` as a CUDA tensor (read
inside the kernel) and `viewdir` as a CPU tensor; get_interpolation_weights reads `viewpoint` on the host. Either
device is accepted and copied as needed."""
import ctypes as C
import os

import torch

from lodgs import LodgsError, check, current_stream, lib, ptr

import lodgs as _lodgs
_lodgs.load()   # the plugin cannot work without the native library: fail at import, loudly

_ws_cache = {}          # (device index) -> workspace tensor (grown on demand; the reference mallocs per call)
_traversable = {}       # _hier_key -> bool, from lodgs_cut_prepare
_last_count = {}        # _hier_key -> K of the previous cut of this hierarchy (the next frame's is close to it)
TRAVERSAL_MIN_NODES = 16_000_000   # below this the single-pass flat kernel is faster than ~25 grid-synchronised levels
TRAVERSAL_MAX_CUT_FRACTION = 0.10  # traversal time grows with the nodes it visits (about 2 K), the flat pass reads all N at
#                                    ~3 TB/s: measured cross-over on 33.5 M nodes is K ~ 0.12 N (profiles/r1_cut_flat_ab_s15.json)


def _hier_key(nodes, boxes):
    """Identity of a hierarchy for the two caches: both tensors' storage AND their version counters, so an in-place edit of nodes or
    boxes, or a new hierarchy that the caching allocator placed at a recycled address, never inherits a stale 'monotone' verdict (a new
    tensor at an old address starts at version 0, so the entry is additionally dropped when its tensors die: see _remember)."""
    return (nodes.data_ptr(), boxes.data_ptr(), nodes.size(0), nodes._version, boxes._version)


def _remember(cache, key, value, nodes, boxes):
    import weakref
    cache[key] = value
    for t in (nodes, boxes):   # drop the entry with the tensors it describes (address reuse after free)
        try:
            weakref.finalize(t, cache.pop, key, None)
        except TypeError:
            pass


def _workspace(N, dev):
    need = int(lib.lodgs_cut_workspace_bytes(N))
    ws = _ws_cache.get(dev.index)
    if ws is None or ws.numel() < need:
        ws = torch.empty(need, dtype=torch.uint8, device=dev)
        _ws_cache[dev.index] = ws
    return ws


def prepare_hierarchy(nodes, boxes):
    """Validate once that the traversal cut selects exactly the flat rule's set (monotone sizes, single root)."""
    N = nodes.size(0)
    ws = _workspace(N, nodes.device)
    ok = C.c_int32(0)
    check(lib.lodgs_cut_prepare(ptr(nodes), ptr(boxes), N, ws.data_ptr(), ws.numel(), C.addressof(ok), current_stream()))
    _remember(_traversable, _hier_key(nodes, boxes), bool(ok.value), nodes, boxes)
    return bool(ok.value)


def expand_to_size(nodes, boxes, size, viewpoint, viewdir, render_indices, parent_indices, nodes_for_render_indices,
                   mode=None):
    """-> number of cut entries K; writes render_indices[:K], parent_indices[:K], nodes_for_render_indices[:K]."""
    if nodes.device.type != "cuda":
        raise RuntimeError("expand_to_size needs CUDA tensors (there is no CPU path)")
    dev = nodes.device
    nodes, boxes = nodes.contiguous(), boxes.contiguous()
    N = nodes.size(0)
    vp = viewpoint.to(device=dev, dtype=torch.float32).contiguous()
    key = _hier_key(nodes, boxes)
    if mode is None and os.environ.get("LODGS_CUT_MODE"):        # A/B: 0 = flat pass, 1 = level-synchronous traversal
        mode = int(os.environ["LODGS_CUT_MODE"])
    if mode is None:
        # traversal only pays on a big hierarchy cut shallowly; the previous frame's K is the predictor (first call: flat)
        mode = 1 if (N >= TRAVERSAL_MIN_NODES and _traversable.get(key, False)
                     and _last_count.get(key, N) < TRAVERSAL_MAX_CUT_FRACTION * N) else 0
    if mode == 1 and not _traversable.get(key, False):
        if not prepare_hierarchy(nodes, boxes):
            mode = 0
    ws = _workspace(N, dev)
    capacity = min(render_indices.numel(), parent_indices.numel() if parent_indices.numel() else 1 << 62,
                   nodes_for_render_indices.numel() if nodes_for_render_indices.numel() else 1 << 62)
    count = C.c_int32(0)
    check(lib.lodgs_expand_to_size(ptr(nodes), ptr(boxes), N, float(size), ptr(vp), ptr(render_indices), ptr(parent_indices),
                                   ptr(nodes_for_render_indices), int(capacity), ws.data_ptr(), ws.numel(), int(mode),
                                   C.addressof(count), current_stream()))
    if key in _last_count:
        _last_count[key] = int(count.value)
    else:
        _remember(_last_count, key, int(count.value), nodes, boxes)
    return int(count.value)


def get_interpolation_weights(indices, size, nodes, boxes, viewpoint, viewdir, ts, num_kids):
    K = indices.size(0)
    if K == 0:
        return
    v = viewpoint.detach().to("cpu", torch.float32)
    check(lib.lodgs_get_interpolation_weights(ptr(indices.contiguous()), K, float(size), ptr(nodes.contiguous()),
                                              ptr(boxes.contiguous()), float(v[0]), float(v[1]), float(v[2]), ptr(ts),
                                              ptr(num_kids), current_stream()))


def expand_to_size_dynamic(nodes, positions, scales, size, viewpoint, viewdir, render_indices, parent_indices,
                           nodes_for_render_indices):
    """GH/torch/torch_interface.cpp:171-194 (ExpandToSizeDynamic): cut on the dynamic hierarchy format, nodes [N,6] i32
    {depth, parent, child_count, first_child, next_sibling, max_side_length}, positions/scales [N,3]. `viewpoint` is read on
    the device, `viewdir` on the host, as in the reference. -> K; a cut entry is the node id."""
    if nodes.device.type != "cuda":
        raise RuntimeError("expand_to_size_dynamic needs CUDA tensors (there is no CPU path)")
    dev = nodes.device
    nodes, positions, scales = nodes.contiguous(), positions.contiguous(), scales.contiguous()
    N = nodes.size(0)
    vp = viewpoint.to(device=dev, dtype=torch.float32).contiguous()
    z = viewdir.detach().to("cpu", torch.float32)
    ws = _workspace(N, dev)
    capacity = min(render_indices.numel(), parent_indices.numel() if parent_indices.numel() else 1 << 62,
                   nodes_for_render_indices.numel() if nodes_for_render_indices.numel() else 1 << 62)
    count = C.c_int32(0)
    check(lib.lodgs_expand_to_size_dynamic(ptr(nodes), ptr(positions), ptr(scales), N, float(size), ptr(vp), float(z[0]), float(z[1]),
                                           float(z[2]), ptr(render_indices), ptr(parent_indices), ptr(nodes_for_render_indices),
                                           int(capacity), ws.data_ptr(), ws.numel(), C.addressof(count), current_stream()))
    return int(count.value)


def get_interpolation_weights_dynamic(indices, size, nodes, positions, scales, viewpoint, viewdir, ts, num_kids):
    """GH/torch/torch_interface.cpp:222-244 (GetTsIndexedDynamic); `viewpoint` is read on the host."""
    K = indices.size(0)
    if K == 0:
        return
    v = viewpoint.detach().to("cpu", torch.float32)
    check(lib.lodgs_get_interpolation_weights_dynamic(ptr(indices.contiguous()), K, float(size), ptr(nodes.contiguous()),
                                                      ptr(positions.contiguous()), ptr(scales.contiguous()), float(v[0]), float(v[1]),
                                                      float(v[2]), ptr(ts), ptr(num_kids), current_stream()))


def get_spt_cut_cuda(number_of_SPTs, gaussian_indices, SPT_starts, SPT_max, SPT_min, SPT_indices, SPT_distances):
    """GH/torch/torch_interface.cpp:287-316 (GetSPTCut): -> (cut ids i32[K], exclusive prefix sum of the per-slot counts i32[S]),
    both on the device. Slot i selects table SPT_indices[i] at distance SPT_distances[i]."""
    if gaussian_indices.device.type != "cuda":
        raise RuntimeError("get_spt_cut_cuda needs CUDA tensors (there is no CPU path)")
    dev = gaussian_indices.device
    S = int(number_of_SPTs)
    counts = torch.empty(S, dtype=torch.int32, device=dev)
    if S == 0:
        return torch.empty(0, dtype=torch.int32, device=dev), counts
    gi, st = gaussian_indices.contiguous(), SPT_starts.contiguous()
    mx, mn = SPT_max.contiguous(), SPT_min.contiguous()
    si, sd = SPT_indices.to(torch.int32).contiguous(), SPT_distances.to(torch.float32).contiguous()
    for t, dt, name in ((gi, torch.int32, "gaussian_indices"), (st, torch.int32, "SPT_starts"), (mx, torch.float32, "SPT_max"),
                        (mn, torch.float32, "SPT_min")):
        if t.dtype != dt:
            raise RuntimeError("get_spt_cut_cuda: %s must be %s" % (name, dt))
    n_entries = min(gi.numel(), mx.numel(), mn.numel())
    ws = torch.empty(lib.lodgs_spt_cut_workspace_bytes(S), dtype=torch.uint8, device=dev)
    count = C.c_int32(0)
    check(lib.lodgs_spt_cut_count(S, ptr(gi), n_entries, ptr(st), ptr(mx), ptr(mn), ptr(si), ptr(sd), ptr(counts), ws.data_ptr(), ws.numel(),
                                  C.addressof(count), current_stream()))
    cut = torch.empty(int(count.value), dtype=torch.int32, device=dev)
    check(lib.lodgs_spt_cut_write(S, ptr(gi), n_entries, ptr(st), ptr(mn), ptr(si), ptr(sd), ws.data_ptr(), ptr(cut), cut.numel(),
                                  current_stream()))
    return cut, counts


def get_morton_indices(xyz, min, max, code):  # noqa: A002  (the reference's argument names)
    """GH/torch/torch_interface.cpp GetMortonCode: writes the 63-bit Morton code of every row of `xyz` into `code` (i64[P], CUDA)."""
    if xyz.device.type != "cuda" or code.device.type != "cuda":
        raise RuntimeError("get_morton_indices needs CUDA tensors (there is no CPU path)")
    if code.dtype != torch.int64 or not code.is_contiguous() or code.numel() < xyz.size(0):
        raise RuntimeError("get_morton_indices: code must be a contiguous int64 tensor with one entry per point")
    dev = xyz.device
    x = xyz.detach().to(torch.float32).contiguous()
    mn = min.detach().to(device=dev, dtype=torch.float32).contiguous()
    mx = max.detach().to(device=dev, dtype=torch.float32).contiguous()
    check(lib.lodgs_get_morton_codes(ptr(x), ptr(mn), ptr(mx), code.data_ptr(), x.size(0), current_stream()))


def load_hierarchy(filename):
    """-> (pos[P,3], shs[P,16,3], alpha[P,1], scales[P,3], rot[P,4], nodes[N,7] i32, boxes[N,2,4]) CPU tensors."""
    P, N = C.c_int32(0), C.c_int32(0)
    fn = str(filename).encode()
    try:
        check(lib.lodgs_hier_probe(fn, C.addressof(P), C.addressof(N)))
    except LodgsError as e:
        raise RuntimeError(str(e).split(": ", 1)[-1])
    P, N = P.value, N.value
    pos, shs = torch.empty(P, 3), torch.empty(P, 16, 3)
    alpha, scales, rot = torch.empty(P, 1), torch.empty(P, 3), torch.empty(P, 4)
    nodes, boxes = torch.empty(N, 7, dtype=torch.int32), torch.empty(N, 2, 4)
    check(lib.lodgs_hier_load(fn, pos.data_ptr(), shs.data_ptr(), alpha.data_ptr(), scales.data_ptr(), rot.data_ptr(),
                              nodes.data_ptr(), boxes.data_ptr()))
    return pos, shs, alpha, scales, rot, nodes, boxes


def write_hierarchy(filename, pos, shs, opacities, log_scales, rotations, nodes, boxes, compressed=False):
    t = [x.detach().to("cpu").contiguous() for x in (pos, shs, opacities, log_scales, rotations)]
    nodes_c, boxes_c = nodes.detach().to("cpu", torch.int32).contiguous(), boxes.detach().to("cpu", torch.float32).contiguous()
    check(lib.lodgs_hier_write(str(filename).encode(), t[0].size(0), nodes_c.size(0), t[0].data_ptr(), t[1].data_ptr(),
                               t[2].data_ptr(), t[3].data_ptr(), t[4].data_ptr(), nodes_c.data_ptr(), boxes_c.data_ptr(),
                               1 if compressed else 0))


def load_dynamic_hierarchy(filename):
    """GH/torch/torch_interface.cpp:43-75 (LoadDynamicHierarchy) -> (pos[P,3], shs[P,(deg+1)^2,3], alpha[P,1], scales[P,3], rot[P,4],
    nodes[P,6] i32) CPU tensors; one node per Gaussian, as the reference loads them."""
    P, deg, N = C.c_int32(0), C.c_int32(0), C.c_int32(0)
    fn = str(filename).encode()
    try:
        check(lib.lodgs_dhier_probe(fn, C.addressof(P), C.addressof(deg), C.addressof(N)))
    except LodgsError as e:
        raise RuntimeError(str(e).split(": ", 1)[-1])
    P, deg = P.value, deg.value
    pos, shs = torch.empty(P, 3), torch.empty(P, (deg + 1) ** 2, 3)
    alpha, scales, rot = torch.empty(P, 1), torch.empty(P, 3), torch.empty(P, 4)
    nodes = torch.empty(P, 6, dtype=torch.int32)
    check(lib.lodgs_dhier_load(fn, pos.data_ptr(), shs.data_ptr(), alpha.data_ptr(), scales.data_ptr(), rot.data_ptr(), nodes.data_ptr()))
    return pos, shs, alpha, scales, rot, nodes


def write_dynamic_hierarchy(filename, pos, shs, opacities, log_scales, rotations, nodes, SH_degree):
    """GH/torch/torch_interface.cpp:107-134 (WriteDynamicHierarchy): tensors on any device; shs holds 3*(SH_degree+1)^2 floats per row."""
    t = [x.detach().to("cpu", torch.float32).contiguous() for x in (pos, shs, opacities, log_scales, rotations)]
    nodes_c = nodes.detach().to("cpu", torch.int32).contiguous()
    P, deg = t[0].size(0), int(SH_degree)
    if not 0 <= deg <= 3 or t[1].numel() != P * 3 * (deg + 1) ** 2:
        raise RuntimeError("write_dynamic_hierarchy: shs must hold 3*(SH_degree+1)^2 floats per Gaussian")
    check(lib.lodgs_dhier_write(str(filename).encode(), P, nodes_c.size(0), deg, t[0].data_ptr(), t[1].data_ptr(), t[2].data_ptr(),
                                t[3].data_ptr(), t[4].data_ptr(), nodes_c.data_ptr()))


def expand_to_target(nodes, target):
    """GH/torch/torch_interface.cpp:136-142 (ExpandToTarget): depth-target cut on the host -> i32 CPU tensor of Gaussian ids."""
    n = nodes.detach().to("cpu", torch.int32).contiguous()
    count = C.c_int64(0)
    check(lib.lodgs_expand_to_target(n.data_ptr(), n.size(0), int(target), None, 0, C.addressof(count)))
    out = torch.empty(count.value, dtype=torch.int32)
    check(lib.lodgs_expand_to_target(n.data_ptr(), n.size(0), int(target), out.data_ptr(), out.numel(), C.addressof(count)))
    return out
