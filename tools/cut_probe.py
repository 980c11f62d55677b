import sys, os, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hierarchical-lod-gaussians_b200")); sys.path.insert(0, ROOT)
import torch, numpy as np
import lodgs
from lodgs import synth, lib, check, ptr, current_stream
from gaussian_hierarchy import _C as gh
dev = torch.device("cuda")
for k in (21, 24):
    nodes, boxes = synth.build_hierarchy_torch(k, dev)
    N = nodes.size(0)
    gh.prepare_hierarchy(nodes, boxes)
    ws = gh._workspace(N, dev)
    ri, pi, ni = (torch.empty(N, dtype=torch.int32, device=dev) for _ in range(3))
    ts, kids = torch.empty(N, device=dev), torch.empty(N, dtype=torch.int32, device=dev)
    vp = torch.tensor([120.0, 0.0, 30.0], device=dev)
    for tau_px in (3.0, 15.0):
        tau = synth.tau_to_threshold(tau_px, 0.8, 1920)
        for mode in (0, 1):
            def cut(sync):
                cnt = C.c_int32(0)
                check(lib.lodgs_expand_to_size(ptr(nodes), ptr(boxes), N, float(tau), ptr(vp), ptr(ri), ptr(pi), ptr(ni), N, ws.data_ptr(),
                                               ws.numel(), mode, C.addressof(cnt) if sync else None, current_stream()))
                return cnt.value
            K = cut(True)
            for sync in (False, True):
                for _ in range(3): cut(sync)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(20): cut(sync)
                e1.record(); torch.cuda.synchronize()
                print("k=%d tau=%g mode=%d sync=%d K=%d  %.4f ms  (%.0f GB/s flat-equivalent)" % (k, tau_px, mode, sync, K, e0.elapsed_time(e1) / 20, (60.0 * N + 12.0 * K) / (e0.elapsed_time(e1) / 20 * 1e-3) / 1e9))
        idx = ni[:K]
        for _ in range(3):
            check(lib.lodgs_get_interpolation_weights(ptr(idx), K, float(tau), ptr(nodes), ptr(boxes), 120.0, 0.0, 30.0, ptr(ts), ptr(kids), current_stream()))
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            check(lib.lodgs_get_interpolation_weights(ptr(idx), K, float(tau), ptr(nodes), ptr(boxes), 120.0, 0.0, 30.0, ptr(ts), ptr(kids), current_stream()))
        e1.record(); torch.cuda.synchronize()
        print("k=%d weights K=%d %.4f ms (%.0f GB/s of 108 B/entry)" % (k, K, e0.elapsed_time(e1) / 20, 108.0 * K / (e0.elapsed_time(e1) / 20 * 1e-3) / 1e9))
