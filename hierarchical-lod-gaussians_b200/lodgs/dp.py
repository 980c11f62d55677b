"""Multi-view data-parallel training of one chunk (BASELINE.json configs[4]): one process per GPU, every rank renders its own
view of the replicated scene through the alt-rasterizer conventions (dc + SH rest, antialiasing; AR/alt_gaussian_rasterization), the
parameter gradients are exchanged and every rank applies the same sparse Adam step (AR SparseGaussianAdam semantics: rows with a
non-zero summed opacity gradient, train_single.py:171-172; no bias correction), so the replicas stay identical.

Two exchanges:
* "allreduce": NCCL sum all-reduce of all six gradient tensors, 59 floats = 236 B per Gaussian (what SURVEY §8e describes).
* "factored" (default): the SH gradient is an outer product basis(mean - campos_v) (x) dL_drgb_v, so only its 3-float colour factor travels:
  one all-reduce over the 11 geometry floats (means 3, scales 3, rotations 4, opacity 1 — one flat buffer the backward writes into
  directly) and one all-gather of dL_drgb [P,3]; `lodgs_sh_adam_from_rgb` then rebuilds sum_v dL_dsh_v per row in registers, views in
  rank order on every rank, and applies the update. 44 + 12 B per Gaussian leave a rank instead of 236, dL_dsh is never written, and
  the SH update reads 24 instead of 28 B per scalar.
No CPU fallback; the exchange runs on NCCL (NVLink / NVSwitch) when torch.distributed is initialised, and is skipped at world size 1."""
import ctypes as C

import torch
import torch.distributed as dist

from . import RASTER_BACKGROUND_TWICE, RASTER_CLAMPED_ALPHA_GRAD, RASTER_SQUARE_RECT, check, current_stream, lib, ptr
from .optim import sh_adam_from_rgb

AR_FLAGS = RASTER_SQUARE_RECT | RASTER_BACKGROUND_TWICE | RASTER_CLAMPED_ALPHA_GRAD     # antialiasing on: no NO_ANTIALIAS


def exchange_gradients(flat, drgb=None, drgb_all=None, dsh=None):
    """The one exchange step of multi-view training, on any torch.distributed backend (NCCL on the box, gloo in the CPU tests).
    flat: the geometry gradients of all parameters in one buffer -> summed in place by ONE all-reduce. Then either the factored SH gradient
    (drgb [P,3] of this rank's view gathered into drgb_all [world,P,3], rank order) or the materialised one (dsh summed in place).
    -> (handle of the flat all-reduce, handle of the SH part, bytes this rank handed to the collectives)"""
    w_flat = dist.all_reduce(flat, async_op=True)
    nbytes = flat.numel() * flat.element_size()
    if dsh is None:
        w_sh = dist.all_gather_into_tensor(drgb_all.view(-1, drgb.size(-1)), drgb.view(-1, drgb.size(-1)), async_op=True)   # concatenation form: every backend takes it
        nbytes += drgb.numel() * drgb.element_size()
    else:
        w_sh = dist.all_reduce(dsh, async_op=True)
        nbytes += dsh.numel() * dsh.element_size()
    return w_flat, w_sh, nbytes


class MultiViewTrainer:
    def __init__(self, par, lrs, cam, bg, W, H, degree, rank, world, mode="auto", eps=1e-15):
        """par: dict of CUDA tensors means3D [P,3], dc [P,1,3], shs_rest [P,K,3], opacities [P,1], scales [P,3], rotations [P,4] (copied);
        lrs: learning rate per key; cam: dict with viewmatrix / projmatrix / campos / tanfovx / tanfovy of THIS rank's view."""
        self.mode = "factored" if mode == "auto" else mode
        if self.mode not in ("factored", "allreduce"):
            raise ValueError("mode must be 'factored' or 'allreduce'")
        self.rank, self.world, self.cam, self.bg, self.W, self.H, self.degree, self.lrs, self.eps = rank, world, cam, bg, W, H, degree, lrs, eps
        dev = par["means3D"].device
        P = par["means3D"].size(0)
        self.P, self.dev = P, dev
        self.M = 1 + par["shs_rest"].size(1)
        self.shs = torch.cat((par["dc"], par["shs_rest"]), dim=1).contiguous()          # one [P,M,3] block: what the kernels read
        self.par = {"means3D": par["means3D"].clone(), "dc": self.shs[:, :1], "shs_rest": self.shs[:, 1:], "opacities": par["opacities"].clone(),
                    "scales": par["scales"].clone(), "rotations": par["rotations"].clone()}
        z = torch.zeros_like
        self.m = {k: z(self.par[k]) for k in ("means3D", "opacities", "scales", "rotations")}
        self.v = {k: z(self.par[k]) for k in ("means3D", "opacities", "scales", "rotations")}
        self.m_shs, self.v_shs = z(self.shs), z(self.shs)
        # gradient buffers, allocated once: the 11 geometry floats in one flat block (a single collective), the blend's accumulators in another
        self.flat = torch.empty(11 * P, device=dev)
        self.g = {"means3D": self.flat[:3 * P].view(P, 3), "scales": self.flat[3 * P:6 * P].view(P, 3), "rotations": self.flat[6 * P:10 * P].view(P, 4),
                  "opacities": self.flat[10 * P:].view(P, 1)}
        self.acc = torch.empty(11 * P, device=dev)                                       # conic 4, mean2D 3, colour 3, inverse depth 1
        self.dcov = torch.empty(P, 6, device=dev)
        self.drgb = torch.empty(P, 3, device=dev)
        self.dsh = torch.empty(P, self.M, 3, device=dev) if self.mode == "allreduce" else None
        self.zero_inv = torch.zeros(1, H, W, device=dev)
        self.ei, self.ef = torch.empty(0, dtype=torch.int32, device=dev), torch.empty(0, device=dev)
        self.campos_all = cam["campos"].reshape(1, 3).contiguous()
        self.drgb_all = self.drgb.view(1, P, 3)
        if world > 1:
            self.campos_all = torch.empty(world, 3, device=dev)
            dist.all_gather_into_tensor(self.campos_all, cam["campos"].reshape(1, 3).contiguous())
            self.drgb_all = torch.empty(world, P, 3, device=dev)

    def _adam(self, key, vis):
        p = self.par[key]
        check(lib.lodgs_adam_step(p.data_ptr(), self.g[key].data_ptr(), self.m[key].data_ptr(), self.v[key].data_ptr(), None, 0, vis.data_ptr(),
                                  self.P, p.numel() // self.P, float(self.lrs[key]), 0.9, 0.999, float(self.eps), 1, 0, current_stream()))

    def step(self, target, timed=False):
        """One training step on this rank's view against `target` [3,H,W] (MSE). -> (events or None, bytes this rank handed to the collectives)"""
        from diff_gaussian_rasterization import _C as hr
        c, P, W, H = self.cam, self.P, self.W, self.H
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)] if timed else None
        if timed:
            ev[0].record()
        R, color, radii, geom, binning, img, invdepth, _seen = hr.forward_impl(
            self.bg, self.ei, self.ei, self.ef, self.ei, self.par["means3D"], self.ef, self.par["opacities"], self.par["scales"],
            self.par["rotations"], 1.0, self.ef, c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, self.shs, self.degree,
            c["campos"], False, False, True, AR_FLAGS)
        dL = (color - target) * (2.0 / color.numel())
        self.acc.zero_()
        self.g["opacities"].zero_()                       # accumulated by the blend backward before the preprocess backward finishes it
        a = self.acc
        dconic, dmean2D, dcolor, dinv = a[:4 * P], a[4 * P:7 * P], a[7 * P:10 * P], a[10 * P:]
        factored = self.mode == "factored"
        check(lib.lodgs_backward_rgb(P, P, self.degree, self.M, int(R), W, H, ptr(self.bg), None, None, None, None, ptr(self.par["means3D"]),
                                     ptr(self.shs), None, ptr(self.par["opacities"]), ptr(self.par["scales"]), 1.0, ptr(self.par["rotations"]), None,
                                     ptr(c["viewmatrix"]), ptr(c["projmatrix"]), ptr(c["campos"]), float(c["tanfovx"]), float(c["tanfovy"]),
                                     ptr(radii), ptr(geom), ptr(binning), ptr(img), ptr(dL), ptr(self.zero_inv), ptr(dmean2D), ptr(dconic),
                                     ptr(self.g["opacities"]), ptr(dcolor), ptr(dinv), ptr(self.g["means3D"]), ptr(self.dcov),
                                     None if factored else ptr(self.dsh), ptr(self.g["scales"]), ptr(self.g["rotations"]),
                                     ptr(self.drgb) if factored else None, AR_FLAGS, current_stream()))
        if timed:
            ev[1].record()
        nbytes = 0
        w_flat = w_rgb = w_sh = None
        if self.world > 1:
            if factored:
                w_flat, w_rgb, nbytes = exchange_gradients(self.flat, self.drgb, self.drgb_all)
            else:
                w_flat, w_sh, nbytes = exchange_gradients(self.flat, dsh=self.dsh)
            w_flat.wait()
        if timed:
            ev[2].record()
        vis = (self.g["opacities"] != 0).reshape(-1)
        for k in ("opacities", "scales", "rotations"):
            self._adam(k, vis)
        if factored:
            if w_rgb is not None:
                w_rgb.wait()
            sh_adam_from_rgb(self.drgb_all, self.campos_all, self.par["means3D"], self.par["dc"], self.m_shs[:, :1], self.v_shs[:, :1],
                             self.par["shs_rest"], self.m_shs[:, 1:], self.v_shs[:, 1:], self.degree, self.lrs["dc"], self.lrs["shs_rest"],
                             eps=self.eps, visible=vis, bias_correction=False)
        else:
            if w_sh is not None:
                w_sh.wait()
            # dc and rest carry different learning rates: two strided halves of the fused block, updated from the summed dL_dsh by the
            # same kernel (one view whose colour factor is the row itself is not expressible) -> plain element-wise Adam per half
            for lo, hi, key in ((0, 1, "dc"), (1, self.M, "shs_rest")):
                pv, gv = self.shs[:, lo:hi].contiguous(), self.dsh[:, lo:hi].contiguous()
                mv, vv = self.m_shs[:, lo:hi].contiguous(), self.v_shs[:, lo:hi].contiguous()
                check(lib.lodgs_adam_step(pv.data_ptr(), gv.data_ptr(), mv.data_ptr(), vv.data_ptr(), None, 0, vis.data_ptr(), P, pv.numel() // P,
                                          float(self.lrs[key]), 0.9, 0.999, float(self.eps), 1, 0, current_stream()))
                self.shs[:, lo:hi], self.m_shs[:, lo:hi], self.v_shs[:, lo:hi] = pv, mv, vv
        self._adam("means3D", vis)                        # last: the SH update reads the means the forward used
        if timed:
            ev[3].record()
        return ev, nbytes
