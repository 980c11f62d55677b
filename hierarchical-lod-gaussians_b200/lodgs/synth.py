"""Seeded synthetic inputs: Gaussians in a view frustum, the 1080p camera, kd-tree hierarchies.
Everything is generated on the CPU with numpy so every implementation sees identical bits."""
import math

import numpy as np


def projection_matrix(znear, zfar, tanfovx, tanfovy):
    """utils/graphics_utils.py:51-77 (getProjectionMatrix), row-major 4x4."""
    top, right = tanfovy * znear, tanfovx * znear
    bottom, left = -top, -right
    P = np.zeros((4, 4), dtype=np.float64)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def make_camera(W=1920, H=1080, fx=1200.0, fy=1200.0, R=None, T=None, znear=0.01, zfar=100.0):
    """Camera tensors as scene/cameras.py:102-107 builds them: viewmatrix = W2C^T, projmatrix = viewmatrix @ P^T,
    campos = inv(viewmatrix)[3,:3]. R is camera-to-world rotation (as stored by the reference), T world-to-camera."""
    tanfovx, tanfovy = W / (2.0 * fx), H / (2.0 * fy)
    R = np.eye(3) if R is None else np.asarray(R, dtype=np.float64)
    T = np.zeros(3) if T is None else np.asarray(T, dtype=np.float64)
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = R.T
    Rt[:3, 3] = T
    Rt[3, 3] = 1.0
    view = Rt.T.astype(np.float32)  # world_view_transform
    proj = projection_matrix(znear, zfar, tanfovx, tanfovy).T.astype(np.float32)
    full = (view.astype(np.float32) @ proj.astype(np.float32)).astype(np.float32)
    campos = np.linalg.inv(view.astype(np.float64))[3, :3].astype(np.float32)
    return {"W": W, "H": H, "tanfovx": float(tanfovx), "tanfovy": float(tanfovy), "viewmatrix": view, "projmatrix": full,
            "campos": campos}


def look_at_camera(eye, target, up=(0.0, -1.0, 0.0), **kw):
    eye, target, up = (np.asarray(v, dtype=np.float64) for v in (eye, target, up))
    z = target - eye
    z /= np.linalg.norm(z)
    x = np.cross(-up, z)  # image y points down
    x /= np.linalg.norm(x)
    y = np.cross(z, x)
    R = np.stack([x, y, z], axis=1)  # camera-to-world, columns = camera axes in world
    T = -R.T @ eye
    return make_camera(R=R, T=T, **kw)


def make_gaussians(P, seed=0, tanfovx=0.8, tanfovy=0.45, zmin=1.0, zmax=40.0, scale_mean=0.02, scale_sigma=0.6, sh_degree=3):
    """Scene A: activated parameters as gaussian_renderer.render() would pass them."""
    rng = np.random.default_rng(seed)
    M = 16
    z = rng.uniform(zmin, zmax, P)
    x = rng.uniform(-1, 1, P) * 0.9 * z * tanfovx * 1.2
    y = rng.uniform(-1, 1, P) * 0.9 * z * tanfovy * 1.2
    means = np.stack([x, y, z], 1).astype(np.float32)
    scales = np.exp(rng.normal(math.log(scale_mean), scale_sigma, (P, 3))).astype(np.float32)
    q = rng.normal(0, 1, (P, 4))
    rots = (q / np.linalg.norm(q, axis=1, keepdims=True)).astype(np.float32)
    opac = (1.0 / (1.0 + np.exp(-rng.normal(0, 2, (P, 1))))).astype(np.float32)
    shs = np.zeros((P, M, 3), dtype=np.float32)
    shs[:, 0, :] = rng.normal(0.5, 0.5, (P, 3)) / 0.2821
    shs[:, 1:, :] = rng.normal(0, 0.1, (P, M - 1, 3))
    return {"means3D": means, "scales": scales, "rotations": rots, "opacities": opac, "shs": shs, "sh_degree": sh_degree}


def build_hierarchy(points, radii, seed=0):
    """Binary kd-tree over `points` [L,3] (median split on the longest axis of the node's AABB, boxes = union of
    +-3*radius cubes: GH/PointbasedKdTreeGenerator.cpp:16-68), flattened with upstream populateRec numbering
    (GH/writer.cpp:21-96: children of a node occupy a contiguous id block, blocks appended in DFS pre-order).
    One Gaussian per node: leaves own Gaussians [0,L) in leaf order of creation; interior node i owns a merged
    Gaussian (count_merged=1). Returns nodes[N,7] i32, boxes[N,2,4] f32, gauss_pos[N,3], gauss_rad[N].
    Pure Python recursion: for test sizes only (see lodgs_synth_* for the large benchmarks)."""
    import sys
    sys.setrecursionlimit(100000)
    L = points.shape[0]
    pts = np.asarray(points, dtype=np.float32)
    rad = np.asarray(radii, dtype=np.float32)

    # explicit tree first
    class T:  # noqa: D401
        __slots__ = ("lo", "hi", "children", "bmin", "bmax", "leaf")

    def gen(idx):
        t = T()
        t.bmin = (pts[idx] - 3.0 * rad[idx, None]).min(0)
        t.bmax = (pts[idx] + 3.0 * rad[idx, None]).max(0)
        t.children = []
        t.leaf = -1
        if len(idx) == 1:
            t.leaf = int(idx[0])
            return t
        axis = int(np.argmax(t.bmax - t.bmin))
        order = idx[np.argsort(pts[idx, axis], kind="stable")]
        mid = len(order) // 2
        t.children = [gen(order[:mid]), gen(order[mid:])]
        return t

    root = gen(np.arange(L))
    nodes, boxes, gpos, grad = [], [], [], []

    def gaussian_of(t):
        if t.leaf >= 0:
            return pts[t.leaf], rad[t.leaf]
        c = 0.5 * (t.bmin + t.bmax)
        return c.astype(np.float32), np.float32(0.25 * float((t.bmax - t.bmin).max()))

    def new_node(t, depth, parent):
        nid = len(nodes)
        p, r = gaussian_of(t)
        gid = len(gpos)
        gpos.append(p)
        grad.append(r)
        is_leaf = t.leaf >= 0
        w = 0.0 if is_leaf else float((t.bmax - t.bmin).max())
        nodes.append([depth, parent, gid, 1 if is_leaf else 0, 0 if is_leaf else 1, -1, 0])
        boxes.append([[t.bmin[0], t.bmin[1], t.bmin[2], w], [t.bmax[0], t.bmax[1], t.bmax[2], w]])
        return nid

    def populate(t, nid, depth):
        if not t.children:
            return
        start = len(nodes)
        nodes[nid][5] = start
        nodes[nid][6] = len(t.children)
        ids = [new_node(c, depth + 1, nid) for c in t.children]
        for c, cid in zip(t.children, ids):
            populate(c, cid, depth + 1)

    rid = new_node(root, 0, -1)
    populate(root, rid, 0)
    return (np.asarray(nodes, dtype=np.int32), np.asarray(boxes, dtype=np.float32), np.asarray(gpos, dtype=np.float32),
            np.asarray(grad, dtype=np.float32))


def tau_to_threshold(tau_px, tanfovx, width):
    """render_hierarchy.py:56."""
    return (2.0 * (tau_px + 0.5)) * tanfovx / (0.5 * width)


def build_hierarchy_torch(k, device, seed=0, extent=(200.0, 200.0, 10.0), radius_mean=0.1):
    """Large synthetic hierarchy for the cut benchmarks, built with torch ops on `device` in
    seconds: a perfect binary tree over L = 2^k leaves (N = 2L-1 nodes, one Gaussian per node, Gaussian id = node id).
    Leaves are the points in Morton order (so siblings are spatial neighbours, like a kd-tree's), boxes are unions of the
    children's boxes bottom-up (leaf box = point +- 3*radius, GH/PointbasedKdTreeGenerator.cpp:19-33), box.w = largest box
    side for interior nodes and 0 for leaves (GH/ClusterMerger.cpp:165-168). Node ids are level order: the two children of
    node i are 2i+1, 2i+2 (a contiguous block, as upstream populateRec produces). Returns nodes[N,7] i32, boxes[N,2,4] f32."""
    import torch
    g = torch.Generator(device="cpu").manual_seed(seed)
    L = 1 << k
    N = 2 * L - 1
    ext = torch.tensor(extent, dtype=torch.float32)
    pts = ((torch.rand(L, 3, generator=g) * 2 - 1) * ext).to(device)
    rad = torch.exp(torch.randn(L, generator=g) * 0.5 + float(np.log(radius_mean))).to(device)
    # Morton order over a 2^10 grid per axis
    q = ((pts / ext.to(device) * 0.5 + 0.5).clamp(0, 1 - 1e-7) * 1024).long()
    code = torch.zeros(L, dtype=torch.int64, device=device)
    for b in range(10):
        for a in range(3):
            code |= ((q[:, a] >> b) & 1) << (3 * b + a)
    order = torch.argsort(code)
    pts, rad = pts[order], rad[order]
    boxes = torch.zeros(N, 2, 4, dtype=torch.float32, device=device)
    lo, hi = pts - 3.0 * rad[:, None], pts + 3.0 * rad[:, None]
    first = L - 1
    boxes[first:, 0, :3], boxes[first:, 1, :3] = lo, hi
    for d in range(k - 1, -1, -1):
        n = 1 << d
        lo = lo.view(n, 2, 3).min(1).values
        hi = hi.view(n, 2, 3).max(1).values
        s = n - 1
        boxes[s:s + n, 0, :3], boxes[s:s + n, 1, :3] = lo, hi
        w = (hi - lo).max(1).values
        boxes[s:s + n, 0, 3], boxes[s:s + n, 1, 3] = w, w
    ids = torch.arange(N, dtype=torch.int64, device=device)
    depth = torch.floor(torch.log2((ids + 1).double())).long()
    leaf = ids >= first
    nodes = torch.stack([depth, torch.where(ids == 0, torch.full_like(ids, -1), (ids - 1) // 2), ids, leaf.long(), (~leaf).long(),
                         torch.where(leaf, torch.full_like(ids, -1), 2 * ids + 1), torch.where(leaf, torch.zeros_like(ids),
                                                                                              torch.full_like(ids, 2))], 1).int()
    return nodes.contiguous(), boxes.contiguous()


def to_dynamic_hierarchy(nodes, boxes, gauss_pos, gauss_rad, seed=0, skybox=0):
    """Re-express a static hierarchy (build_hierarchy) in the DYNAMIC format of this fork (GH/types.h:60-79 `HierarchyNode`):
    nodes[N,6] i32 {depth, parent, child_count, first_child, next_sibling, max_side_length}, positions[N,3], scales[N,3]
    (one Gaussian per node: position = the node's Gaussian, anisotropic scales around its radius). `skybox` extra nodes with
    depth -1 are appended (they are never selected by the cut, GH/runtime_switching.cu:556-560)."""
    rng = np.random.default_rng(seed)
    N = nodes.shape[0]
    dyn = np.zeros((N + skybox, 6), np.int32)
    dyn[:N, 0] = nodes[:, 0]
    dyn[:N, 1] = nodes[:, 1]
    dyn[:N, 2] = nodes[:, 6]
    dyn[:N, 3] = np.where(nodes[:, 6] > 0, nodes[:, 5], -1)
    nxt = np.full(N, -1, np.int32)
    for i in range(N):
        c0, cc = nodes[i, 5], nodes[i, 6]
        for k in range(cc - 1):
            nxt[c0 + k] = c0 + k + 1
    dyn[:N, 4] = nxt
    side = (boxes[:, 1, :3] - boxes[:, 0, :3]).max(1)
    dyn[:N, 5] = np.minimum(side, 2 ** 30).astype(np.int32)
    pos = np.zeros((N + skybox, 3), np.float32)
    sca = np.zeros((N + skybox, 3), np.float32)
    gid = nodes[:, 2]
    pos[:N] = gauss_pos[gid]
    sca[:N] = gauss_rad[gid, None] * rng.uniform(0.4, 1.0, (N, 3)).astype(np.float32)
    if skybox:
        dyn[N:, 0] = -1
        dyn[N:, 1] = -1
        dyn[N:, 3] = -1
        dyn[N:, 4] = -1
        pos[N:] = rng.uniform(-500, 500, (skybox, 3)).astype(np.float32)
        sca[N:] = 50.0
    return dyn, pos, sca
