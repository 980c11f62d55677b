"""GPU: the factored SH gradient path (lodgs_backward_rgb + lodgs_sh_adam_from_rgb, csrc/sh_adam.cu) against the materialised one
(dL_dsh written by the backward, lodgs_adam_step on it — itself pinned to scene/OurAdam.py and AR's adamUpdateCUDA by the other tests):
identical geometry gradients, dL_dsh == basis (x) dL_drgb, and the same parameters / moments after the update, for one view and for the
sum over several views (the multi-view training exchange), index list / visibility mask / all rows, both Adam conventions."""
import numpy as np
import pytest
import torch

import cases
import lodgs
from diff_gaussian_rasterization import _C
from lodgs import synth
from lodgs.optim import sh_adam_from_rgb

pytestmark = pytest.mark.gpu


def _view(sc, cam, bg, seed):
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    c = {k: (t(v) if isinstance(v, np.ndarray) else v) for k, v in cam.items()}
    ei, ef = torch.empty(0, dtype=torch.int32, device="cuda"), torch.empty(0, device="cuda")
    fa = (bg, ei, ei, ef, ei, sc["means3D"], ef, sc["opacities"], sc["scales"], sc["rotations"], 1.0, ef, c["viewmatrix"], c["projmatrix"],
          c["tanfovx"], c["tanfovy"], c["H"], c["W"], sc["shs"], 3, c["campos"], False, False, True)
    R, color, radii, geom, binning, img, invdepth, seen = _C.rasterize_gaussians(*fa)
    g = torch.Generator(device="cuda").manual_seed(seed)
    dL = (torch.rand(3, c["H"], c["W"], device="cuda", generator=g) - 0.5) / color.numel()
    dLi = (torch.rand(1, c["H"], c["W"], device="cuda", generator=g) - 0.5) / color.numel()
    ba = (bg, ei, ei, ef, ei, sc["means3D"], radii, ef, sc["opacities"], sc["scales"], sc["rotations"], 1.0, ef, c["viewmatrix"], c["projmatrix"],
          c["tanfovx"], c["tanfovy"], dL, dLi, sc["shs"], 3, c["campos"], geom, R, binning, img, False)
    return c, radii, _C.rasterize_gaussians_backward(*ba), _C.rasterize_gaussians_backward_rgb(*ba)


def _scene(name="wide"):
    scn, cam, bg, _, _, _ = cases.make_case(name)
    sc = {k: torch.from_numpy(v).cuda() for k, v in scn.items() if isinstance(v, np.ndarray)}
    W, H, fx = cam["W"], cam["H"], cam["W"] / (2 * cam["tanfovx"])
    cams = [cam] + [synth.make_camera(W=W, H=H, fx=fx, fy=fx, T=(dx, dy, 0.0)) for dx, dy in ((0.4, 0.0), (-0.3, 0.2))]
    return sc, cams, torch.from_numpy(bg).cuda()


def test_backward_rgb_equals_backward_and_factorises_dsh():
    sc, cams, bg = _scene()
    c, radii, full, rgb = _view(sc, cams[0], bg, 1)
    for i in (0, 1, 2, 3, 4, 6, 7):                                   # everything but the SH gradient: same kernels (atomic order differs)
        assert float((full[i] - rgb[i]).norm()) <= 1e-5 * float(full[i].norm()) + 1e-30, i
    assert rgb[5] is None and rgb[8].shape == (sc["means3D"].size(0), 3)
    vis = radii > 0
    assert float(rgb[8][~vis].abs().max()) == 0.0
    # dL_dsh[k] = basis_k(dir) * dL_drgb: coefficient 0 carries the constant basis value
    # (two runs of the blend backward: the order of its float atomics differs, and a colour gradient is a sum of terms of both signs, so
    # the comparison is relative to each Gaussian's own gradient scale, not element by element)
    def close(a, b):
        scale = torch.maximum(a.abs().amax(dim=1, keepdim=True), b.abs().amax(dim=1, keepdim=True))
        return bool(((a - b).abs() <= 1e-3 * scale + 1e-30).all()) and float((a - b).norm()) <= 1e-5 * float(a.norm())
    assert close(full[5][:, 0], 0.28209479177387814 * rgb[8])
    d = torch.nn.functional.normalize(sc["means3D"] - c["campos"], dim=1)
    assert close(full[5][:, 2], 0.4886025119029199 * d[:, 2:3] * rgb[8])


@pytest.mark.parametrize("bias,rows", [(True, "list"), (True, "all"), (False, "mask"), (False, "all")])
@pytest.mark.parametrize("n_views", [1, 3])
def test_sh_adam_from_rgb_matches_adam_on_summed_dsh(bias, rows, n_views):
    sc, cams, bg = _scene()
    P = sc["means3D"].size(0)
    views = [_view(sc, cams[v], bg, 10 + v) for v in range(n_views)]
    dsh = sum(v[2][5] for v in views)                                  # what an all-reduce of dL_dsh delivers
    drgb = torch.stack([v[3][8] for v in views])
    campos = torch.stack([v[0]["campos"] for v in views])
    seen = torch.stack([v[1] > 0 for v in views]).any(0)
    g = torch.Generator(device="cuda").manual_seed(5)
    m0, v0 = torch.randn(P, 16, 3, device="cuda", generator=g) * 1e-4, torch.rand(P, 16, 3, device="cuda", generator=g) * 1e-8
    lr_dc, lr_rest, eps, step = 2.5e-3, 1.25e-4, 1e-15, 3
    relevant = seen.nonzero().reshape(-1) if rows == "list" else None
    visible = seen if rows == "mask" else None
    # materialised path: separate contiguous dc / rest tensors, lodgs_adam_step on each
    want = {}
    for lo, hi, lr in ((0, 1, lr_dc), (1, 16, lr_rest)):
        p, m, v, gr = (x[:, lo:hi].contiguous() for x in (sc["shs"].clone(), m0, v0, dsh))
        lodgs.check(lodgs.lib.lodgs_adam_step(p.data_ptr(), gr.data_ptr(), m.data_ptr(), v.data_ptr(), relevant.data_ptr() if relevant is not None else None,
                                              relevant.numel() if relevant is not None else 0, visible.data_ptr() if visible is not None else None, P,
                                              p.numel() // P, lr, 0.9, 0.999, eps, step, 1 if bias else 0, lodgs.current_stream()))
        want[lo] = (p, m, v)
    # factored path on ONE fused [P,16,3] block
    shs, m, v = sc["shs"].clone(), m0.clone(), v0.clone()
    sh_adam_from_rgb(drgb, campos, sc["means3D"], shs[:, :1], m[:, :1], v[:, :1], shs[:, 1:], m[:, 1:], v[:, 1:], 3, lr_dc, lr_rest, eps=eps,
                     step=step, relevant=relevant, visible=visible, bias_correction=bias)
    torch.cuda.synchronize()
    for lo, hi in ((0, 1), (1, 16)):
        wp, wm, wv = want[lo]
        assert torch.allclose(m[:, lo:hi], wm, rtol=1e-5, atol=2e-7 * float(wm.abs().max()))
        assert torch.allclose(v[:, lo:hi], wv, rtol=1e-5, atol=2e-7 * float(wv.abs().max()))
        assert torch.allclose(shs[:, lo:hi], wp, rtol=2e-6, atol=2e-5 * (lr_dc if lo == 0 else lr_rest))
    if rows != "all":                                                   # untouched rows keep their bits
        assert torch.equal(shs[~seen], sc["shs"][~seen]) and torch.equal(m[~seen], m0[~seen])
    assert not torch.equal(shs[seen], sc["shs"][seen])


def test_multi_view_trainer_modes_agree():
    """lodgs.dp.MultiViewTrainer at world size 1: the factored exchange path and the all-reduce path take the same training steps."""
    from lodgs import dp
    scn, cam, bg, _, _, _ = cases.make_case("small")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    c = {k: (t(v) if isinstance(v, np.ndarray) else v) for k, v in cam.items()}
    shs = t(scn["shs"])
    par = {"means3D": t(scn["means3D"]), "dc": shs[:, :1].contiguous(), "shs_rest": shs[:, 1:].contiguous(), "opacities": t(scn["opacities"]),
           "scales": t(scn["scales"]), "rotations": t(scn["rotations"])}
    lrs = {"means3D": 1.6e-4, "dc": 2.5e-3, "shs_rest": 1.25e-4, "opacities": 5e-2, "scales": 5e-3, "rotations": 1e-3}
    target = torch.rand(3, c["H"], c["W"], device="cuda", generator=torch.Generator(device="cuda").manual_seed(3))
    res = {}
    for mode in ("factored", "allreduce"):
        tr = dp.MultiViewTrainer(par, lrs, c, t(bg), c["W"], c["H"], 3, 0, 1, mode=mode)
        for _ in range(3):
            tr.step(target)
        torch.cuda.synchronize()
        res[mode] = {k: v.clone() for k, v in tr.par.items()}
    for k in res["factored"]:
        a, b = res["factored"][k], res["allreduce"][k]
        assert torch.allclose(a, b, rtol=1e-5, atol=1e-5 * lrs[k]), (k, float((a - b).abs().max()))
        assert not torch.equal(a, par[k])
