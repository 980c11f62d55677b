"""TEST INFRASTRUCTURE ONLY — ctypes binding of oracle/_ref/libref_hier.so: the reference's UNMODIFIED hierarchy file reader / writer
(GH/hierarchy_loader.cpp:26-189, GH/hierarchy_writer.cpp:27-168) and depth-target traversal (GH/traversal.cpp:15-39), compiled against
oracle/eigen_stub by oracle/Makefile; plus a runner for oracle/_ref/GaussianHierarchyCreator (GH/mainHierarchyCreator.cpp)."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_ref", "libref_hier.so")
CREATOR = os.path.join(HERE, "_ref", "GaussianHierarchyCreator")
_lib = None


def available():
    return os.path.exists(LIB)


def creator_available():
    return os.path.exists(CREATOR)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(LIB)
        _lib.ref_expand_to_target.restype = C.c_long
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def load(path):
    """HierarchyLoader::load -> dict of numpy arrays (pos[P,3] shs[P,48] alpha[P] scales[P,3] rot[P,4] nodes[N,7] boxes[N,8])."""
    P, N = C.c_int(0), C.c_int(0)
    if lib().ref_hier_load(path.encode(), C.byref(P), C.byref(N)) != 0:
        raise RuntimeError("reference loader threw")
    P, N = P.value, N.value
    o = {"pos": np.empty((P, 3), np.float32), "shs": np.empty((P, 48), np.float32), "alpha": np.empty(P, np.float32),
         "scales": np.empty((P, 3), np.float32), "rot": np.empty((P, 4), np.float32), "nodes": np.empty((N, 7), np.int32),
         "boxes": np.empty((N, 8), np.float32)}
    lib().ref_hier_fetch(*[_p(o[k]) for k in ("pos", "shs", "alpha", "scales", "rot", "nodes", "boxes")])
    return o


def write(path, h, compressed):
    a = {k: np.ascontiguousarray(v) for k, v in h.items()}
    if lib().ref_hier_write(path.encode(), a["pos"].shape[0], a["nodes"].shape[0], _p(a["pos"]), _p(a["shs"]), _p(a["alpha"]), _p(a["scales"]),
                            _p(a["rot"]), _p(a["nodes"]), _p(a["boxes"]), int(bool(compressed))) != 0:
        raise RuntimeError("reference writer threw")


def load_dynamic(path):
    P, deg = C.c_int(0), C.c_int(0)
    if lib().ref_dhier_load(path.encode(), C.byref(P), C.byref(deg)) != 0:
        raise RuntimeError("reference loader threw")
    P, deg = P.value, deg.value
    M3 = 3 * (deg + 1) ** 2
    o = {"pos": np.empty((P, 3), np.float32), "shs": np.empty((P, M3), np.float32), "alpha": np.empty(P, np.float32),
         "scales": np.empty((P, 3), np.float32), "rot": np.empty((P, 4), np.float32), "nodes": np.empty((P, 6), np.int32), "sh_degree": deg}
    # the reference keeps one 48-float slot per Gaussian and fills the file's packed block into the front of that ARRAY
    # (hierarchy_loader.cpp:160-175); the torch binding then views the first P*M3 floats (torch_interface.cpp:62-64)
    lib().ref_dhier_fetch(_p(o["pos"]), _p(o["shs"]), M3, _p(o["alpha"]), _p(o["scales"]), _p(o["rot"]), _p(o["nodes"]))
    return o


def write_dynamic(path, h, sh_degree):
    a = {k: np.ascontiguousarray(v) for k, v in h.items() if isinstance(v, np.ndarray)}
    if lib().ref_dhier_write(path.encode(), a["pos"].shape[0], a["nodes"].shape[0], _p(a["pos"]), _p(a["shs"]), _p(a["alpha"]), _p(a["scales"]),
                             _p(a["rot"]), _p(a["nodes"]), int(sh_degree)) != 0:
        raise RuntimeError("reference writer threw")


def expand_to_target(nodes, target):
    nodes = np.ascontiguousarray(nodes, np.int32)
    n = lib().ref_expand_to_target(_p(nodes), int(target), None, 0)
    out = np.empty(n, np.int32)
    lib().ref_expand_to_target(_p(nodes), int(target), _p(out), n)
    return out


def run_creator(ply_path, out_prefix, timeout=3600):
    """GaussianHierarchyCreator <ply> <source dir (unused: the appearance filter is commented out, mainHierarchyCreator.cpp:165-177)>
    <output prefix> -> writes <prefix>hierarchy.dhier (+ .gdf). Returns wall seconds."""
    import time
    t = time.perf_counter()
    r = subprocess.run([CREATOR, ply_path, os.path.dirname(ply_path) or ".", out_prefix], stdout=subprocess.DEVNULL, stderr=subprocess.PIPE,
                       timeout=timeout)
    if r.returncode != 0:
        raise RuntimeError("reference creator failed: %s" % r.stderr.decode()[-300:])
    return time.perf_counter() - t
