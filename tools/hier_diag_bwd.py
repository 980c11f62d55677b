import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "hierarchical-lod-gaussians_b200"), ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch
from gaussian_hierarchy import _C as gh
from lodgs import synth
from oracle import refshim
import bench_hier_step as B
import minewrap
dev = torch.device("cuda", 0)
nodes, boxes, sc = B.make_scene(18, dev)
sc["sh_degree"] = 3
N = nodes.size(0)
cam = minewrap.to_cuda(synth.look_at_camera(eye=(-150.0, 20.0, 18.0), target=(0.0, 0.0, 0.0), up=(0.0, 0.0, 1.0), W=1920, H=1080, fx=1200.0, fy=1200.0))
bg = torch.zeros(3, device=dev)
ri, pi, ni = (torch.empty(N, dtype=torch.int32, device=dev) for _ in range(3))
ts, kids = torch.empty(N, device=dev), torch.empty(N, dtype=torch.int32, device=dev)
tau = synth.tau_to_threshold(3.0, cam["tanfovx"], 1920)
K = gh.expand_to_size(nodes, boxes, tau, cam["campos"], torch.zeros(3), ri, pi, ni)
gh.get_interpolation_weights(ni[:K], tau, nodes, boxes, cam["campos"].cpu(), torch.zeros(3), ts, kids)
kw = {"indices": ri[:K].clone(), "parent_indices": pi[:K].clone(), "ts": ts[:K].clone(), "kids": kids[:K].clone()}
print("K", K, "ts min/max", float(ts[:K].min()), float(ts[:K].max()), "kids min/max", int(kids[:K].min()), int(kids[:K].max()), "parents<0", int((pi[:K] < 0).sum()))
target = torch.rand(3, 1080, 1920, device=dev)
m = minewrap.forward(sc, cam, bg, **kw)
r = refshim.forward(sc, cam, bg, **kw)
dL = (m["color"] - target) * (2.0 / target.numel())
dLi = torch.zeros(1, 1080, 1920, device=dev)
gm = minewrap.backward(sc, cam, bg, m, dL, dLi, **kw)
gr = refshim.backward(sc, cam, bg, r, dL, dLi, **kw)
for k in gm:
    a, b = gm[k], gr[k]
    print("%-14s mine norm %.4e nnz %d | ref norm %.4e nnz %d | rel %.3e" % (k, float(a.double().norm()), int((a != 0).sum()), float(b.double().norm()),
                                                                          int((b != 0).sum()), float((a.double() - b.double()).norm() / max(float(b.double().norm()), 1e-300))))
