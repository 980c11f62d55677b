"""benchlib.py — the two arms of bench.py (kept separate so that each file stays small).

run_ours      : product path (diff_gaussian_rasterization._C -> liblodgs.so).
run_reference : the UNMODIFIED reference CUDA rasterizer rebuilt for sm_100a (oracle/_ref/libref_lodgs.so, built by
                oracle/Makefile from /root/reference) on the same inputs through the same protocol. The reference has
                no CPU implementation of this path, so its own code on the same B200 is the honest
                reference arm; the CPU restatement is reported separately as cpu_baseline.
"""
import json
import os
import time

import numpy as np
import torch

from bench import ClockSampler, DEG, H, M, ROOT, W, make_inputs

E = lambda dev, dt=torch.float32: torch.empty(0, device=dev, dtype=dt)  # noqa: E731
PARAM_KEYS = ("means3D", "shs", "opacities", "scales", "rotations")


def _pin(d):
    return {k: torch.from_numpy(np.ascontiguousarray(v)).pin_memory() for k, v in d.items() if isinstance(v, np.ndarray)}


def _barrier(world):
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()


def _max_over_ranks(x, world, dev):
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    return float(t.item())


def _sum_over_ranks(x, world, dev):
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.SUM)
    return float(t.item())


def _timed(step_fn, steps, warmup, world, dev, sampler=None, finish_fn=None):
    """W untimed warm-up steps, then exactly K steps between barrier+synchronize, CUDA events on the current stream."""
    for _ in range(warmup):
        step_fn()
    _barrier(world)
    if sampler is not None:
        sampler.start()
        time.sleep(0.25)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _barrier(world)
    ev0.record()
    for _ in range(steps):
        step_fn()
    if finish_fn is not None:
        finish_fn()
    ev1.record()
    _barrier(world)
    ms = ev0.elapsed_time(ev1)
    return _max_over_ranks(ms, world, dev)


class OursArm:
    def __init__(self, sc_host, cam_host, target_host, dev):
        from diff_gaussian_rasterization import _C
        self._C = _C
        self.dev = dev
        self.host = _pin(sc_host)
        self.target_host = torch.from_numpy(target_host).pin_memory()
        self.cam = {k: (torch.from_numpy(v).to(dev) if isinstance(v, np.ndarray) else v) for k, v in cam_host.items()}
        self.bg = torch.zeros(3, device=dev)
        self.d = {k: self.host[k].to(dev) for k in PARAM_KEYS}
        self.target = self.target_host.to(dev)
        self.ei, self.ef = E(dev, torch.int32), E(dev)
        self.out_host = None
        self.last = None
        self.pipe = None

    def fwd_bwd(self, d, target):
        c = self.cam
        fa = (self.bg, self.ei, self.ei, self.ef, self.ei, d["means3D"], self.ef, d["opacities"], d["scales"], d["rotations"], 1.0,
              self.ef, c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, d["shs"], DEG, c["campos"], False, False,
              True)
        R, color, radii, geom, binning, img, invdepth, seen = self._C.rasterize_gaussians(*fa)
        dL = (color - target) * (2.0 / color.numel())
        dLi = torch.zeros_like(invdepth)
        ba = (self.bg, self.ei, self.ei, self.ef, self.ei, d["means3D"], radii, self.ef, d["opacities"], d["scales"], d["rotations"],
              1.0, self.ef, c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], dL, dLi, d["shs"], DEG, c["campos"], geom, R,
              binning, img, False)
        grads = self._C.rasterize_gaussians_backward(*ba)
        self.last = (R, color, radii, img, grads)
        return color, grads

    def step_resident(self):
        self.fwd_bwd(self.d, self.target)

    # ---- training iteration (first half of "train iters/s"): fwd + bwd + sparse Adam on the visible Gaussians, the
    # update rule of scene/OurAdam.py Adam.step(relevant) with relevant = (radii > 0).nonzero() (train_single.py:171-176)
    GRAD_OF = {"means3D": 3, "shs": 5, "opacities": 2, "scales": 6, "rotations": 7}   # index into the backward's 8-tuple
    # the training learning rates x 1e-3: the cost of an Adam step does not depend on them, and the scene (hence the work of
    # the timed forward/backward) stays what it is over the timed iterations
    LR = {"means3D": 1.6e-7, "shs": 2.5e-6, "opacities": 5e-5, "scales": 5e-6, "rotations": 1e-6}

    def train_init(self):
        self.tp = {k: self.d[k].clone() for k in PARAM_KEYS}
        self.tm = {k: torch.zeros_like(v) for k, v in self.tp.items()}
        self.tv = {k: torch.zeros_like(v) for k, v in self.tp.items()}
        self.tstep = 0
        self._adam_tab = None

    def adam(self, relevant, grads):
        """All parameter tensors in one foreign call (lodgs_adam_step_multi), as lodgs.optim.Adam.step(relevant) issues it."""
        import ctypes as C

        import lodgs
        self.tstep += 1
        n = len(PARAM_KEYS)
        if getattr(self, "_adam_tab", None) is None:
            P4 = C.c_void_p * n
            rows = [self.tp[k].size(0) for k in PARAM_KEYS]
            self._adam_tab = (P4(*[self.tp[k].data_ptr() for k in PARAM_KEYS]), P4(*[self.tm[k].data_ptr() for k in PARAM_KEYS]),
                              P4(*[self.tv[k].data_ptr() for k in PARAM_KEYS]), (C.c_int64 * n)(*rows),
                              (C.c_int32 * n)(*[self.tp[k].numel() // r for k, r in zip(PARAM_KEYS, rows)]),
                              (C.c_double * n)(*[self.LR[k] for k in PARAM_KEYS]), (C.c_double * n)(*[0.9] * n), (C.c_double * n)(*[0.999] * n),
                              (C.c_double * n)(*[1e-15] * n))
        tp, tm, tv, rows, widths, lrs, b1, b2, eps = self._adam_tab
        gs = [grads[self.GRAD_OF[k]].contiguous() for k in PARAM_KEYS]
        lodgs.check(lodgs.lib.lodgs_adam_step_multi(n, tp, (C.c_void_p * n)(*[g.data_ptr() for g in gs]), tm, tv, relevant.data_ptr(),
                                                    relevant.numel(), None, rows, widths, lrs, b1, b2, eps, (C.c_int64 * n)(*[self.tstep] * n), 1,
                                                    lodgs.current_stream()))

    def step_train(self):
        """fwd + bwd + sparse Adam with the SH gradient kept in its factored form: the backward writes dL_drgb (12 B per Gaussian) instead
        of dL_dsh (192 B), `lodgs_sh_adam_from_rgb` rebuilds the row in registers while it applies the update (csrc/sh_adam.cu)."""
        import ctypes as C

        import lodgs
        from lodgs.optim import sh_adam_from_rgb
        c, d = self.cam, self.tp
        fa = (self.bg, self.ei, self.ei, self.ef, self.ei, d["means3D"], self.ef, d["opacities"], d["scales"], d["rotations"], 1.0,
              self.ef, c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, d["shs"], DEG, c["campos"], False, False, True)
        R, color, radii, geom, binning, img, invdepth, seen = self._C.rasterize_gaussians(*fa)
        dL = (color - self.target) * (2.0 / color.numel())
        ba = (self.bg, self.ei, self.ei, self.ef, self.ei, d["means3D"], radii, self.ef, d["opacities"], d["scales"], d["rotations"],
              1.0, self.ef, c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], dL, torch.zeros_like(invdepth), d["shs"], DEG,
              c["campos"], geom, R, binning, img, False)
        grads = self._C.rasterize_gaussians_backward_rgb(*ba)
        relevant = (radii > 0).nonzero().reshape(-1)
        self.tstep += 1
        sh_adam_from_rgb(grads[8], c["campos"], d["means3D"], d["shs"][:, :1], self.tm["shs"][:, :1], self.tv["shs"][:, :1], d["shs"][:, 1:],
                         self.tm["shs"][:, 1:], self.tv["shs"][:, 1:], DEG, self.LR["shs"], self.LR["shs"], eps=1e-15, step=self.tstep,
                         relevant=relevant)
        keys = ("opacities", "scales", "rotations", "means3D")        # the means last: the SH update reads the ones the forward used
        n = len(keys)
        P4 = C.c_void_p * n
        gs = [grads[self.GRAD_OF[k]] for k in keys]
        rows = d["means3D"].size(0)
        lodgs.check(lodgs.lib.lodgs_adam_step_multi(
            n, P4(*[d[k].data_ptr() for k in keys]), P4(*[g.data_ptr() for g in gs]), P4(*[self.tm[k].data_ptr() for k in keys]),
            P4(*[self.tv[k].data_ptr() for k in keys]), relevant.data_ptr(), relevant.numel(), None, (C.c_int64 * n)(*[rows] * n),
            (C.c_int32 * n)(*[d[k].numel() // rows for k in keys]), (C.c_double * n)(*[self.LR[k] for k in keys]), (C.c_double * n)(*[0.9] * n),
            (C.c_double * n)(*[0.999] * n), (C.c_double * n)(*[1e-15] * n), (C.c_int64 * n)(*[self.tstep] * n), 1, lodgs.current_stream()))

    def step_train_materialised(self):
        """The same iteration with dL_dsh written by the backward and read again by the optimizer (round-1 path; A/B and parity)."""
        color, grads = self.fwd_bwd(self.tp, self.target)
        relevant = (self.last[2] > 0).nonzero().reshape(-1)
        self.adam(relevant, grads)

    def _enqueue_h2d(self, slot):
        """Copy one step's inputs from pinned host memory into buffer slot `slot` on the H2D stream."""
        pp = self.pipe
        if pp["in"][slot] is None:
            pp["in"][slot] = ({k: torch.empty_like(self.d[k]) for k in PARAM_KEYS}, torch.empty_like(self.target))
        d, target = pp["in"][slot]
        with torch.cuda.stream(pp["h2d"]):
            if pp["ev_free"][slot] is not None:
                pp["h2d"].wait_event(pp["ev_free"][slot])       # the kernels of the step that last used this slot have finished
            for k in PARAM_KEYS:
                d[k].copy_(self.host[k], non_blocking=True)
            target.copy_(self.target_host, non_blocking=True)
            pp["ev_in"][slot] = pp["h2d"].record_event(self._ev())

    def step_e2e(self):
        """One end-to-end step: inputs leave pinned host memory, results land in pinned host memory. Three streams
        (H2D, compute, D2H) and two buffer slots: while step i computes, the inputs of step i+1 are already being copied
        into the other slot (a prefetching loader) and the results of step i-1 are on their way back. Every step moves its
        own inputs and its own results; the timed region ends only after the last D2H AND the last prefetch have landed, so
        K steps contain K complete input copies."""
        if self.pipe is None:
            self.pipe = {"h2d": torch.cuda.Stream(), "d2h": torch.cuda.Stream(), "i": 0, "prefetched": -1,
                         "in": [None, None], "ev_in": [None, None], "ev_free": [None, None], "ev_out": [None, None],
                         "out": [None, None], "keep": [None, None]}
        pp = self.pipe
        i = pp["i"]
        slot = i & 1
        pp["i"] += 1
        comp = torch.cuda.current_stream()
        if pp["prefetched"] != i:
            self._enqueue_h2d(slot)
        self._enqueue_h2d(slot ^ 1)
        pp["prefetched"] = i + 1
        d, target = pp["in"][slot]
        comp.wait_event(pp["ev_in"][slot])
        color, g = self.fwd_bwd(d, target)
        pp["ev_free"][slot] = comp.record_event(self._ev())
        outs = (color, g[3], g[5], g[2], g[6], g[7])  # image, dL_dmeans3D, dL_dsh, dL_dopacity, dL_dscales, dL_drotations
        if pp["out"][slot] is None:
            pp["out"][slot] = [torch.empty(o.shape, dtype=o.dtype).pin_memory() for o in outs]
            self.out_host = pp["out"][slot]
        elif pp["ev_out"][slot] is not None:
            pp["ev_out"][slot].synchronize()                    # host buffers of step i-2 are complete before they are overwritten
        with torch.cuda.stream(pp["d2h"]):
            pp["d2h"].wait_event(pp["ev_free"][slot])
            for h, o in zip(pp["out"][slot], outs):
                o.record_stream(pp["d2h"])
                h.copy_(o, non_blocking=True)
            pp["ev_out"][slot] = pp["d2h"].record_event(self._ev())
        pp["keep"][slot] = outs

    def _ev(self):
        return torch.cuda.Event(enable_timing=getattr(self, "ev_timing", False))  # timing only for tools/e2e_timeline.py

    def finish_e2e(self):
        """Called inside the timed region after the last step: the compute stream waits for every outstanding D2H and for
        the input copy prefetched by the last step."""
        if self.pipe is not None:
            for ev in self.pipe["ev_out"] + self.pipe["ev_in"]:
                if ev is not None:
                    torch.cuda.current_stream().wait_event(ev)

    def io_bytes(self):
        h2d = sum(self.host[k].numel() * 4 for k in PARAM_KEYS) + self.target_host.numel() * 4
        d2h = sum(h.numel() * 4 for h in self.out_host) if self.out_host else 0
        return h2d, d2h


class RefArm(OursArm):
    """Same protocol over the UNMODIFIED reference: its own torch extension (`_C.rasterize_gaussians` / `_backward` of
    HR/rasterize_points.cu + HR/ext.cpp, oracle/_ref/ref_dgr/_C.so), i.e. the stock code path including its allocations, zero fills and the
    mid-forward host synchronisation; the optimizer of the training section is scene/OurAdam.py's `Adam.step(relevant)` (the imported
    class where /root/reference exists, else oracle/ouradam.RestatedAdam, pinned bit-identical to it by tests/test_ouradam_pin.py)."""

    def __init__(self, sc_host, cam_host, target_host, dev):
        import benchsections
        self.ext = benchsections._ref_ext()
        if self.ext is None:
            raise FileNotFoundError("oracle/_ref/ref_dgr/_C.so")
        self.dev = dev
        self.host = _pin(sc_host)
        self.target_host = torch.from_numpy(target_host).pin_memory()
        self.cam = {k: (torch.from_numpy(v).to(dev) if isinstance(v, np.ndarray) else v) for k, v in cam_host.items()}
        self.bg = torch.zeros(3, device=dev)
        self.d = {k: self.host[k].to(dev) for k in PARAM_KEYS}
        self.target = self.target_host.to(dev)
        self.ei, self.ef = E(dev, torch.int32), E(dev)
        self.out_host = None
        self.last = None
        self.pipe = None

    def fwd_bwd(self, d, target):
        c = self.cam
        fa = (self.bg, self.ei, self.ei, self.ef, self.ei, d["means3D"], self.ef, d["opacities"], d["scales"], d["rotations"], 1.0,
              self.ef, c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, d["shs"], DEG, c["campos"], False, False, True)
        R, color, radii, geom, binning, img, invdepth, seen = self.ext.rasterize_gaussians(*fa)
        dL = (color - target) * (2.0 / color.numel())
        dLi = torch.zeros_like(invdepth)
        ba = (self.bg, self.ei, self.ei, self.ef, self.ei, d["means3D"], radii, self.ef, d["opacities"], d["scales"], d["rotations"],
              1.0, self.ef, c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], dL, dLi, d["shs"], DEG, c["campos"], geom, R,
              binning, img, False)
        grads = self.ext.rasterize_gaussians_backward(*ba)
        self.last = (R, color, radii, None, grads)
        return color, grads

    def train_init(self):
        from oracle import ouradam
        cls, self.opt_kind = ouradam.reference_or_restated()
        self.tp = {k: torch.nn.Parameter(self.d[k].clone()) for k in PARAM_KEYS}
        self.opt = cls([{"params": [self.tp[k]], "lr": self.LR[k]} for k in PARAM_KEYS], lr=0.0, eps=1e-15)

    def step_train(self):
        color, grads = self.fwd_bwd(self.tp, self.target)
        relevant = (self.last[2] > 0).nonzero().reshape(-1)
        for k in PARAM_KEYS:
            self.tp[k].grad = grads[self.GRAD_OF[k]]
        with torch.no_grad():
            self.opt.step(relevant)


# ------------------------------------------------------------------------------------------------ roofline
def _peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            p = json.load(fh)
        return float(p.get("hbm_gbs", 6650.0)), float(p.get("sm_max_mhz", 1965.0)), "measured"
    return 6650.0, 1965.0, "fallback"


def _work_model(arm, n_gauss):
    """Algorithmic work per step of every kernel (DESIGN.md 'Kernels and rooflines')."""
    import ctypes as C

    import lodgs
    R, color, radii, img, grads = arm.last
    p_vis = int((radii > 0).sum().item())
    cnt = torch.zeros(3, dtype=torch.int64, device=arm.dev)
    # counters need the buffers of the last forward: redo one forward and count on its buffers
    c, d = arm.cam, arm.d
    fa = (arm.bg, arm.ei, arm.ei, arm.ef, arm.ei, d["means3D"], arm.ef, d["opacities"], d["scales"], d["rotations"], 1.0, arm.ef,
          c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, d["shs"], DEG, c["campos"], False, False, True)
    R, color, radii, geom, binning, imgb, invdepth, seen = arm._C.rasterize_gaussians(*fa)
    lodgs.check(lodgs.lib.lodgs_blend_stats(radii.numel(), W, H, R, 0, geom.data_ptr(), binning.data_ptr(), imgb.data_ptr(), cnt.data_ptr(), lodgs.current_stream()))
    torch.cuda.synchronize()
    n_test, n_alpha, n_blend = (int(x) for x in cnt.tolist())
    work = {
        "preprocess_fwd_kernel": ("hbm", 20.0 * n_gauss + p_vis * (115.0 + 12.0 * M)),
        "sort_depth_hist_kernel": ("hbm", 4.0 * n_gauss),
        "sort_depth_pass_kernel": ("hbm", 16.0 * n_gauss),     # per launch (one 8-bit digit of the depth bits)
        "scan_sorted_kernel": ("hbm", 16.0 * n_gauss),
        "emit_balanced_kernel": ("hbm", 16.0 * n_gauss + 8.0 * R),
        "sort_tile_hist_kernel": ("hbm", 4.0 * R),
        "sort_tile_pass_kernel": ("hbm", 16.0 * R),            # per launch (one 8-bit digit of the tile id)
        "tile_ranges_kernel": ("hbm", 4.0 * R),
        "blend_fwd_kernel": ("fp32", 11.0 * n_test + 2.0 * n_alpha + 14.0 * n_blend),
        "blend_bwd_kernel": ("fp32", 70.0 * n_blend),
        "preprocess_bwd_kernel": ("hbm", p_vis * (115.0 + 68.0 + 24.0 * M)),
    }
    return work, {"num_rendered": R, "p_visible": p_vis, "n_test": n_test, "n_alpha": n_alpha, "n_blend": n_blend}


def _ncu_traffic():
    """DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of each kernel from the newest committed
    `ncu --set full` capture of this same workload (profiles/*ncu_traffic*.json); ({}, None) when there is none."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*ncu_traffic*.json")))
    if not files:
        return {}, None
    with open(files[-1]) as fh:
        t = json.load(fh)
    return {k: max(e["dram_bytes"] for e in v) for k, v in t.get("kernels", {}).items()}, os.path.relpath(files[-1], ROOT)


def _roofline(prof, steps, work, sm_count):
    hbm_peak, sm_max_mhz, which = _peaks()
    traffic, traffic_src = _ncu_traffic()
    fp32_peak = sm_count * 128 * 2 * sm_max_mhz * 1e6 / 1e12  # TFLOP/s at max clock
    per_kernel = []
    for name, (ms, n) in prof.items():
        avg_ms = ms / max(n, 1)
        ent = {"kernel": name, "launches_per_step": n / steps, "avg_ms": round(avg_ms, 5), "ms_per_step": round(ms / steps, 5)}
        if name in work:
            bound, amount = work[name]
            if bound == "hbm":
                ach = amount / (avg_ms * 1e-3) / 1e9
                ent.update({"bound": "hbm", "achieved": round(ach, 1), "peak": hbm_peak, "unit": "GB/s", "frac": round(ach / hbm_peak, 4)})
            else:
                ach = amount / (avg_ms * 1e-3) / 1e12
                ent.update({"bound": "fp32", "achieved": round(ach, 3), "peak": round(fp32_peak, 1), "unit": "TFLOP/s",
                            "frac": round(ach / fp32_peak, 4)})
        per_kernel.append(ent)
    per_kernel.sort(key=lambda e: -e["ms_per_step"])
    dom = next((e for e in per_kernel if "bound" in e), None)
    roof = None
    if dom is not None:
        roof = {"kernel": dom["kernel"], "bound": dom["bound"], "achieved": dom["achieved"], "peak": dom["peak"], "unit": dom["unit"],
                "frac": dom["frac"], "traffic": traffic.get(dom["kernel"]), "traffic_source": traffic_src,
                "peak_source": ("%s HBM copy bandwidth (MEASURED_PEAKS.json)" % which) if dom["bound"] == "hbm" else
                "nominal FP32 pipe: SMs x 128 lanes x 2 x max SM clock (no tensor cores on this path)"}
    return roof, per_kernel


def _cpu_baseline(no_cpu):
    """Oracle port (oracle/raster_oracle.c, OpenMP) timed on a bounded sample of workload A."""
    if no_cpu:
        return None
    try:
        from oracle import oracle as orc
        return orc.bench_sample(W, H)
    except Exception as e:  # noqa: BLE001
        return {"value": None, "unit": "Mpix/s", "cores": 0, "kind": "port", "sample": "unavailable: %s" % str(e)[:120]}


def _line(args, world, value, ms_per_step, clocks, extra):
    line = {"metric": "fwd+bwd Mpix/s @1080p", "value": round(value, 2), "unit": "Mpix/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "A: configs/vanilla.json fwd+bwd rasterization, %d synthetic Gaussians (SH deg 3), one 1920x1080 view "
                                   "per GPU" % args.P, "gaussians": args.P, "width": W, "height": H, "sh_degree": DEG,
                       "parallelism": "views sharded across GPUs, no collective", "l2": "inputs larger than L2 (params 236 MB + "
                       "grads 236 MB + binning state per step)"},
            "clocks": clocks}
    line.update(extra)
    return line


def _train_section(arm, args, world, dev, what):
    """fwd + bwd + sparse Adam per iteration, one view per iteration (configs/vanilla.json training step without the image loss
    terms outside the plugin); single GPU."""
    arm.train_init()
    ms = _timed(arm.step_train, args.steps, 3, world, dev)
    out = {"iters_per_s": round(args.steps / (ms * 1e-3), 2), "ms_per_iter": round(ms / args.steps, 4), "optimizer": what,
           "updated_floats_per_gaussian": 59}
    if isinstance(arm, OursArm) and not isinstance(arm, RefArm):
        import lodgs
        lodgs.profile(True)
        for _ in range(5):
            arm.step_train()
        torch.cuda.synchronize()
        prof = lodgs.profile_dump()
        lodgs.profile(False)
        out["per_kernel_ms"] = {name: round(t / 5, 4) for name, (t, n) in sorted(prof.items(), key=lambda kv: -kv[1][0])}
    return out


def _sections(impl, args, rank, world, dev):
    """The other BASELINE configurations (benchsections.py), run by every rank (eval_path and dp_train are collective), reported by
    rank 0. A section that cannot run reports why instead of aborting the headline line."""
    import benchsections as bs
    want = [x for x in (args.sections or "").split(",") if x]
    out = {}

    def guarded(name, fn):
        if name not in want:
            return
        try:
            torch.cuda.empty_cache()
            t0 = time.perf_counter()
            r = fn()
            if isinstance(r, dict):
                r["section_wall_s"] = round(time.perf_counter() - t0, 2)
            out[name] = r
        except Exception as e:  # noqa: BLE001
            out[name] = {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:300])}
            if world > 1:
                raise
    if world == 1:
        guarded("hier_step", lambda: bs.hier_step(impl, dev, iters=max(3, args.steps // 2), leaves=args.hier_leaves))
    guarded("eval_path", lambda: bs.eval_path(impl, rank, world, dev, frames=args.eval_frames, leaves=args.eval_leaves))
    guarded("dp_train", lambda: bs.dp_train(impl, rank, world, dev, steps=args.steps, warmup=args.warmup, n_gauss=args.P, mode=args.dp_mode))
    if world == 1 and impl == "ours" and not args.no_cpu_baseline:
        guarded("creator", lambda: bs.creator_and_cpu_cut(dev))
    return out


def run_ours(args, rank, world, dev):
    import lodgs
    sc, cam, target = make_inputs(rank, args.P)
    arm = OursArm(sc, cam, target, dev)
    sampler = ClockSampler(dev.index) if rank == 0 else None
    for _ in range(args.warmup):
        arm.step_resident()
    torch.cuda.synchronize()
    lodgs.profile(True)
    launches0 = lodgs.launch_count()
    ms = _timed(arm.step_resident, args.steps, 0, world, dev, sampler)
    launches = lodgs.launch_count() - launches0
    prof = lodgs.profile_dump()
    lodgs.profile(False)
    clocks = sampler.stop() if sampler else None
    ms_e2e = _timed(arm.step_e2e, args.steps, max(8, args.warmup), world, dev, finish_fn=arm.finish_e2e)
    h2d, d2h = arm.io_bytes()
    floor_ms = _copy_floor(arm, args.steps, world, dev)
    mpix = W * H / 1e6
    value = world * mpix * args.steps / (ms * 1e-3)
    e2e = world * mpix * args.steps / (ms_e2e * 1e-3)
    extra = {}
    if rank == 0:
        sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
        work, counters = _work_model(arm, args.P)
        roof, per_kernel = _roofline(prof, args.steps, work, sm_count)
        extra = {"impl": "ours", "e2e": {"value": round(e2e, 2), "unit": "Mpix/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                                         "ms_per_step": round(ms_e2e / args.steps, 4), "copy_floor_ms": round(floor_ms, 4),
                                         "frac_of_copy_floor": round(floor_ms / (ms_e2e / args.steps), 4)},
                 "gpu_launches": int(launches), "roofline": roof, "per_kernel": per_kernel, "workload_counters": counters,
                 "cpu_baseline": _cpu_baseline(args.no_cpu_baseline)}
    if world == 1:
        extra["train"] = _train_section(arm, args, world, dev, "sparse Adam on the rows with radii > 0: lodgs_sh_adam_from_rgb (SH rows rebuilt from "
                                                               "dL_drgb, no dL_dsh in HBM) + lodgs_adam_step_multi")
        arm.train_init()
        ms_mat = _timed(arm.step_train_materialised, args.steps, 3, world, dev)
        extra["train"]["ms_per_iter_with_materialised_dL_dsh"] = round(ms_mat / args.steps, 4)
    del arm
    extra.update(_sections("ours", args, rank, world, dev))
    if rank != 0:
        return None
    return _line(args, world, value, ms / args.steps, clocks, extra)


def _copy_floor(arm, steps, world, dev):
    """The host-link floor of the end-to-end step: the SAME per-step transfers (all inputs host->device, all results device->host, pinned
    memory, both directions concurrently on two streams) with no kernel in between; ms per step, max over ranks."""
    if arm.pipe is None or arm.out_host is None:
        return 0.0
    pp = arm.pipe
    d, target = pp["in"][0]
    outs = pp["keep"][0]
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _barrier(world)
    ev0.record()
    for _ in range(steps):
        with torch.cuda.stream(pp["h2d"]):
            for k in PARAM_KEYS:
                d[k].copy_(arm.host[k], non_blocking=True)
            target.copy_(arm.target_host, non_blocking=True)
        with torch.cuda.stream(pp["d2h"]):
            for h, o in zip(arm.out_host, outs):
                h.copy_(o, non_blocking=True)
    for st in (pp["h2d"], pp["d2h"]):
        torch.cuda.current_stream().wait_stream(st)
    ev1.record()
    _barrier(world)
    return _max_over_ranks(ev0.elapsed_time(ev1) / steps, world, dev)


def run_reference(args, rank, world, dev):
    sc, cam, target = make_inputs(rank, args.P)
    try:
        arm = RefArm(sc, cam, target, dev)
    except (FileNotFoundError, OSError, ImportError) as e:
        return {"impl": "reference", "unavailable": "oracle/_ref/ref_dgr/_C.so (the reference's torch extension) missing or not loadable (%s)" % str(e)[:80]}
    sampler = ClockSampler(dev.index) if rank == 0 else None
    ms = _timed(arm.step_resident, args.steps, args.warmup, world, dev, sampler)
    clocks = sampler.stop() if sampler else None
    ms_e2e = _timed(arm.step_e2e, args.steps, max(8, args.warmup), world, dev, finish_fn=arm.finish_e2e)
    h2d, d2h = arm.io_bytes()
    floor_ms = _copy_floor(arm, args.steps, world, dev)
    mpix = W * H / 1e6
    value = world * mpix * args.steps / (ms * 1e-3)
    e2e = world * mpix * args.steps / (ms_e2e * 1e-3)
    extra = {"impl": "reference",
             "e2e": {"value": round(e2e, 2), "unit": "Mpix/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                     "ms_per_step": round(ms_e2e / args.steps, 4), "copy_floor_ms": round(floor_ms, 4),
                     "frac_of_copy_floor": round(floor_ms / (ms_e2e / args.steps), 4)},
             "cpu_baseline": {"value": round(e2e, 2), "unit": "Mpix/s", "cores": 0, "kind": "reference",
                              "sample": "full workload; the reference has no CPU implementation of the rasterizer, so this arm runs its "
                                        "unmodified CUDA code through its own torch extension (oracle/_ref/ref_dgr, sm_100a rebuild) on the same "
                                        "B200; its CPU code (hierarchy creator) and the CPU cut are timed in the `creator` section of the other arm"}}
    if world == 1:
        extra["train"] = _train_section(arm, args, world, dev, "scene/OurAdam.py Adam.step(relevant) on the rows with radii > 0")
        extra["train"]["optimizer_class"] = getattr(arm, "opt_kind", None)
    del arm
    extra.update(_sections("reference", args, rank, world, dev))
    extra["native_code_note"] = ("timed regions of this arm run only the reference's code (oracle/_ref/ref_dgr/_C.so, libref_lodgs.so, scene/OurAdam.py or its "
                                 "pinned restatement); liblodgs.so is loaded by the hier_step / eval_path sections to BUILD the synthetic hierarchy both arms "
                                 "consume (lodgs_hier_build, untimed) — with --sections '' it is not loaded at all")
    if rank != 0:
        return None
    return _line(args, world, value, ms / args.steps, clocks, extra)
