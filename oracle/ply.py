"""TEST INFRASTRUCTURE ONLY — writes the binary PLY the reference's CPU creator reads (GH/loader.cpp:78-126 `loadPly<LessRichPoint>`;
record = GH/common.h:39-47: position 3, normal 3, shs 12, opacity 1 (logit), scale 3 (log), rotation 4 = 26 f32). The bundled
input.ply of BASELINE.json configs[0] is not in the reference tree (.MISSING_LARGE_BLOBS), so a seeded synthetic one stands in."""
import numpy as np


def synth_points(n, seed=0, extent=(40.0, 40.0, 4.0), scale_mean=0.05, distinct=True):
    """Seeded Gaussians. `distinct`: no two points share a coordinate on any axis (duplicates are nudged to the next float): the kd-tree
    split then selects a unique SET at every node — std::nth_element leaves the order of equal keys unspecified, so with ties the
    reference creator's own result is one of several valid ones."""
    rng = np.random.default_rng(seed)
    pos = (rng.random((n, 3), dtype=np.float32) - 0.5) * np.asarray(extent, np.float32)
    if distinct:
        for a in range(3):
            order = np.argsort(pos[:, a], kind="stable")
            s = pos[order, a].copy()
            while True:
                dup = np.nonzero(s[1:] <= s[:-1])[0]
                if dup.size == 0:
                    break
                s[dup + 1] = np.nextafter(s[dup], np.float32(np.inf))
            pos[order, a] = s
    shs = np.concatenate([rng.normal(0.5, 0.5, (n, 3)) / 0.2821, rng.normal(0, 0.1, (n, 9))], 1).astype(np.float32)
    opacity = rng.normal(0.0, 2.0, (n, 1)).astype(np.float32)
    log_scale = rng.normal(np.log(scale_mean), 0.6, (n, 3)).astype(np.float32)
    rot = rng.normal(0, 1, (n, 4)).astype(np.float32)
    return pos, shs, opacity, log_scale, rot


def write_less_rich_ply(path, pos, shs, opacity, log_scale, rot):
    n = pos.shape[0]
    rec = np.concatenate([pos, np.zeros((n, 3), np.float32), shs, opacity.reshape(n, 1), log_scale, rot], 1).astype("<f4")
    assert rec.shape[1] == 26
    names = ["x", "y", "z", "nx", "ny", "nz"] + ["f_dc_%d" % i for i in range(3)] + ["f_rest_%d" % i for i in range(9)] + \
            ["opacity", "scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"]
    hdr = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % n + "".join("property float %s\n" % s for s in names) + "end_header\n"
    with open(path, "wb") as fh:
        fh.write(hdr.encode())
        fh.write(rec.tobytes())
