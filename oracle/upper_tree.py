"""TEST INFRASTRUCTURE ONLY — the upper-tree cut as a checker.

`reference_functions()` compiles the reference's OWN source of GaussianModel.extract_frustum_planes / frustum_cull_spheres /
cut_hierarchy_on_condition (scene/gaussian_model.py:55-103,354-394) out of the file's syntax tree — the module itself cannot be imported
(plyfile, simple_knn, ... are absent) — and returns them as plain functions; works wherever /root/reference exists. `coarse_cut` restates
the call train_post.py:329-356 makes with them plus the list split that follows, on CPU tensors; tests/test_upper_tree_pin.py checks the
restatement against the compiled reference functions and against tests/golden/upper_tree.npz (their outputs, tests/golden/make_golden_upper.py)."""
import ast
import os

import numpy as np
import torch

REF_PATH = "/root/reference/scene/gaussian_model.py"
CHILD_COUNT, FIRST_CHILD, NEXT_SIBLING, GAUSSIAN = 2, 3, 4, 5


def reference_functions():
    if not os.path.exists(REF_PATH):
        return None
    tree = ast.parse(open(REF_PATH).read())
    want = {"extract_frustum_planes", "frustum_cull_spheres", "cut_hierarchy_on_condition"}
    body = [n for n in tree.body if isinstance(n, ast.Assign) and all(isinstance(t, ast.Name) and t.id.startswith("hierarchy_node_") for t in n.targets)]
    for cls in (n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "GaussianModel"):
        body += [n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name in want]
    ns = {"torch": torch}
    exec(compile(ast.Module(body=body, type_ignores=[]), REF_PATH, "exec"), ns)
    return {k: ns[k] for k in want}


def sum3(t):
    """Row sum of a [n,3] tensor as torch evaluates it where the reference runs (CUDA): (a0 + a2) + a1 (measured on a B200 with torch 2.11,
    profiles/r2_torch_sum3_probe.txt). On CUDA tensors the literal torch.sum is used; on CPU tensors, whose reduction adds (a0 + a1) + a2,
    the CUDA association is written out."""
    return t.sum(dim=-1) if t.is_cuda else (t[:, 0] + t[:, 2]) + t[:, 1]


def cut_hierarchy_on_condition(hierarchy, condition, root_node, leave_out):
    """scene/gaussian_model.py:354-394 with return_upper_tree=False."""
    stack = torch.tensor([root_node], dtype=torch.int32, device=hierarchy.device)
    cut = torch.empty(0, dtype=torch.int32, device=hierarchy.device)
    while len(stack) > 0:
        stack = stack[leave_out(stack)]
        cut = torch.cat((cut, stack[hierarchy[stack.long(), CHILD_COUNT] == 0]))
        stack = stack[hierarchy[stack.long(), CHILD_COUNT] > 0]
        m = condition(stack)
        cut = torch.cat((cut, stack[~m]))
        stack = stack[m]
        first = hierarchy[stack.long(), FIRST_CHILD]
        second = hierarchy[first.long(), NEXT_SIBLING]
        stack = torch.cat((first, second))
    return cut


def frustum_cull_spheres(points, radii, planes):
    """scene/gaussian_model.py:80-103."""
    visible = torch.ones(len(points), dtype=torch.bool, device=points.device)
    for normal, distance in zip(planes[:, :3], planes[:, 3]):
        sd = sum3(points * normal) + distance
        visible[sd + radii < 0] = False
    return visible


def coarse_cut(nodes, xyz, min_dist_sq, campos, bounds, planes, mult=1.0, root=0, fns=None):
    """train_post.py:329-356 on CPU tensors; `fns` = reference_functions() to run the reference's own code instead of the restatement."""
    cull_fn = fns["frustum_cull_spheres"] if fns else (lambda self, p, r, pl: frustum_cull_spheres(p, r, pl))
    if bounds is not None:
        cull = lambda idx: cull_fn(None, xyz[idx.long()], bounds[idx.long()], planes)  # noqa: E731
    else:
        cull = lambda idx: torch.ones(len(idx), dtype=torch.bool, device=idx.device)  # noqa: E731
    lod = lambda idx: min_dist_sq[idx.long()] > sum3((campos - xyz[idx.long()]).square()) * mult  # noqa: E731
    if fns:
        cut = fns["cut_hierarchy_on_condition"](None, nodes, lod, return_upper_tree=False, root_node=root, leave_out_of_cut_condition=cull)
    else:
        cut = cut_hierarchy_on_condition(nodes, lod, root, cull)
    cl = cut.long()
    leaf_nodes = cut[nodes[cl, CHILD_COUNT] == 0]
    ll = leaf_nodes.long()
    spt_indices = nodes[ll][nodes[ll, FIRST_CHILD] >= 0, FIRST_CHILD]
    direct = nodes[ll][nodes[ll, FIRST_CHILD] <= 0, GAUSSIAN]
    spt_nodes = leaf_nodes[nodes[ll, FIRST_CHILD] >= 0]
    d2 = sum3((xyz[spt_nodes.long()] - campos).pow(2))
    # the reference evaluates this on CUDA, where torch's sqrt is the IEEE-rounded sqrtf; torch's vectorised CPU sqrt is not (about 0.6 % of
    # values come out 1 ulp off), so on CPU tensors the root is taken with numpy's
    root_ = d2.sqrt() if d2.is_cuda else torch.from_numpy(np.sqrt(d2.numpy()))
    spt_dist = root_ * mult
    return {"coarse_cut": cut, "SPT_indices": spt_indices, "SPT_node_indices": spt_nodes, "SPT_distances": spt_dist,
            "upper_tree_nodes_to_render": direct}


def make_upper_tree(n_leaves, seed, spt_share=0.7, zero_share=0.05):
    """Seeded upper tree in the layout scene/gaussian_model.py:262-304 builds: pre-order binary tree, leaves carry an SPT id (>= 0), -1, or 0
    in first_child; min_distance_squared from a per-node size; bounding-sphere radii. -> dict of numpy arrays + cameras."""
    rng = np.random.default_rng(seed)
    N = 2 * n_leaves - 1
    nodes = np.zeros((N, 6), np.int32)
    xyz = np.zeros((N, 3), np.float32)
    size = np.zeros(N, np.float32)
    pts = (rng.uniform(-1, 1, (n_leaves, 3)) * np.array([60.0, 60.0, 6.0])).astype(np.float32)
    counter = [0]
    spt = [0]

    def rec(idx, depth, parent):
        i = counter[0]
        counter[0] += 1
        nodes[i, 0], nodes[i, 1] = depth, parent
        xyz[i] = pts[idx].mean(0)
        size[i] = float(np.linalg.norm(pts[idx].max(0) - pts[idx].min(0))) + 0.05
        if len(idx) == 1:
            u = rng.random()
            nodes[i, 2] = 0
            if u < zero_share:
                nodes[i, 3] = 0
            elif u < zero_share + spt_share:
                spt[0] += 1
                nodes[i, 3] = spt[0]
            else:
                nodes[i, 3] = -1
            nodes[i, 5] = 1000 + i
            return i
        ax = int(np.argmax(pts[idx].max(0) - pts[idx].min(0)))
        o = idx[np.argsort(pts[idx, ax], kind="stable")]
        h = len(o) // 2
        nodes[i, 2] = 2
        nodes[i, 5] = 1000 + i
        a = rec(o[:h], depth + 1, i)
        b = rec(o[h:], depth + 1, i)
        nodes[i, 3] = a
        nodes[a, 4] = b
        nodes[b, 4] = 0
        return i
    import sys
    sys.setrecursionlimit(100000)
    rec(np.arange(n_leaves), 0, -1)
    gran = 0.02
    md = ((size / gran) ** 2).astype(np.float32)
    md[0] = 1e12
    bounds = (0.5 * size + rng.uniform(0, 0.5, N)).astype(np.float32)
    return {"nodes": nodes, "xyz": xyz, "min_dist_sq": md, "bounds": bounds}


def make_view(seed):
    """-> (campos [3], full projection transform [4,4]) of a camera looking into the slab."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "hierarchical-lod-gaussians_b200"))
    from lodgs import synth
    rng = np.random.default_rng(seed)
    eye = rng.uniform(-1, 1, 3) * np.array([80.0, 80.0, 5.0]) + np.array([0.0, 0.0, 8.0])
    cam = synth.look_at_camera(eye=tuple(eye), target=tuple(rng.uniform(-20, 20, 3) * np.array([1, 1, 0.1])), up=(0.0, 0.0, 1.0), W=640, H=360, fx=400.0,
                               fy=400.0)
    return cam["campos"].astype(np.float32), cam["projmatrix"].astype(np.float32)
