"""Optimizers of the hot path, same call surface as the reference's, one fused kernel launch per parameter tensor.

* `Adam` mirrors scene/OurAdam.py `Adam.step(relevant)` (:106-176 -> `_single_tensor_adam`, :251-354): Adam with bias
  correction applied only to the rows listed in `relevant` (i64 indices into dim 0); an EMPTY list (or None) updates all
  rows (`_single_tensor_adam2`, :357-452, chosen by `relevant.size(0) == 0` at :214). The reference does ~10 eager elementwise kernels plus index gathers/scatters per
  group; here it is `lodgs_adam_step` (csrc/adam.cu), 28 B per updated scalar.
* `SparseGaussianAdam` mirrors AR/alt_gaussian_rasterization/__init__.py:246-271 (`_C.adamUpdate`, AR/cuda_rasterizer/
  adam.cu:9-38): rows masked by a bool visibility tensor, no bias correction.

There is no eager fallback: parameters must be contiguous float32 CUDA tensors.
"""
import ctypes as C

import torch

from . import check, current_stream, lib, ptr


def _require(p, what):
    if not (p.is_cuda and p.dtype == torch.float32 and p.is_contiguous()):
        raise RuntimeError("lodgs.optim: %s must be a contiguous float32 CUDA tensor" % what)


class Adam(torch.optim.Adam):
    """scene/OurAdam.py Adam: `step(relevant)` with an index list of the rows to update. All parameter tensors are updated by ONE
    foreign call (`lodgs_adam_step_multi`); the pointer tables are rebuilt only when a tensor was reallocated."""

    @torch.no_grad()
    def step(self, relevant=None, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if relevant is not None and relevant.size(0) == 0:
            relevant = None   # an empty list means "every row" in the reference (OurAdam.py:214 -> _single_tensor_adam2)
        if relevant is not None:
            relevant = relevant.to(torch.int64).contiguous()
        todo, keep = [], []
        for group in self.param_groups:
            if group.get("amsgrad") or group.get("weight_decay", 0) != 0 or group.get("maximize"):
                raise NotImplementedError("lodgs.optim.Adam: amsgrad / weight_decay / maximize are not used by the reference's training and "
                                          "are not implemented")
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise RuntimeError("Adam does not support sparse gradients, please consider SparseAdam instead")
                state = self.state[p]
                if len(state) == 0:
                    state["step"] = torch.tensor(0.)
                    state["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    state["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                state["step"] += 1
                _require(p, "parameter")
                grad = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                keep.append(grad)
                rows = p.size(0)
                todo.append((p.data_ptr(), grad.data_ptr(), state["exp_avg"].data_ptr(), state["exp_avg_sq"].data_ptr(), rows,
                             p.numel() // max(rows, 1), float(group["lr"]), float(beta1), float(beta2), float(group["eps"]),
                             int(state["step"].item())))
        n = len(todo)
        if n == 0:
            return loss
        cols = list(zip(*todo))
        P4 = C.c_void_p * n
        check(lib.lodgs_adam_step_multi(n, P4(*cols[0]), P4(*cols[1]), P4(*cols[2]), P4(*cols[3]),
                                        ptr(relevant) if relevant is not None else None, relevant.numel() if relevant is not None else 0, None,
                                        (C.c_int64 * n)(*cols[4]), (C.c_int32 * n)(*cols[5]), (C.c_double * n)(*cols[6]),
                                        (C.c_double * n)(*cols[7]), (C.c_double * n)(*cols[8]), (C.c_double * n)(*cols[9]),
                                        (C.c_int64 * n)(*cols[10]), 1, current_stream()))
        return loss


def sh_adam_from_rgb(dL_drgb, campos, means3D, dc, dc_m, dc_v, rest, rest_m, rest_v, degree, lr_dc, lr_rest, betas=(0.9, 0.999), eps=1e-15,
                     step=1, relevant=None, visible=None, bias_correction=True):
    """Sparse Adam on the SH coefficients straight from the colour gradient of `_C.rasterize_gaussians_backward_rgb` (`lodgs_sh_adam_from_rgb`,
    csrc/sh_adam.cu): dL_drgb [V,P,3] or [P,3], campos [V,3] or [3] (CUDA), means3D [P,3] as the forward saw them; dc [P,1,3] / rest [P,K,3]
    and their moments may be strided row views of one fused [P,1+K,3] tensor (inner dimensions contiguous). Rows: `relevant` (i64 list,
    OurAdam) or all, optionally masked by `visible` (bool [P], AR). No dL_dsh is ever materialised."""
    P = means3D.size(0)
    g = dL_drgb.reshape(-1, P, 3)
    if not g.is_contiguous():
        g = g.contiguous()
    cp = campos.reshape(-1, 3).to(device=means3D.device, dtype=torch.float32).contiguous()
    if cp.size(0) != g.size(0):
        raise RuntimeError("sh_adam_from_rgb: one camera centre per view")
    for t in (dc, dc_m, dc_v):
        if t.stride(-1) != 1 or t[0].numel() != 3 or t.dtype != torch.float32 or not t.is_cuda:
            raise RuntimeError("sh_adam_from_rgb: dc tensors must be float32 CUDA rows of 3 contiguous floats")
    rest_floats = rest[0].numel() if rest is not None and rest.numel() else 0
    if rest_floats:
        for t in (rest, rest_m, rest_v):
            if not t[0].is_contiguous() or t.dtype != torch.float32 or not t.is_cuda:
                raise RuntimeError("sh_adam_from_rgb: rest tensors must be float32 CUDA rows of contiguous floats")
    if relevant is not None and relevant.size(0) == 0:
        relevant = None
    if relevant is not None:
        relevant = relevant.to(torch.int64).contiguous()
    if visible is not None:
        visible = visible.contiguous()
        if visible.dtype not in (torch.bool, torch.uint8):
            visible = visible != 0
    same = lambda a, b: a.stride(0) == b.stride(0)  # noqa: E731
    if not (same(dc, dc_m) and same(dc, dc_v)) or (rest_floats and not (same(rest, rest_m) and same(rest, rest_v))):
        raise RuntimeError("sh_adam_from_rgb: moments must share the layout of their parameter")
    check(lib.lodgs_sh_adam_from_rgb(P, int(degree), g.size(0), g.data_ptr(), g.stride(0), cp.data_ptr(), means3D.contiguous().data_ptr(),
                                     ptr(relevant), relevant.numel() if relevant is not None else 0, ptr(visible), dc.data_ptr(), dc_m.data_ptr(),
                                     dc_v.data_ptr(), dc.stride(0), rest.data_ptr() if rest_floats else None,
                                     rest_m.data_ptr() if rest_floats else None, rest_v.data_ptr() if rest_floats else None,
                                     rest.stride(0) if rest_floats else 0, rest_floats, float(lr_dc), float(lr_rest), float(betas[0]),
                                     float(betas[1]), float(eps), int(step), 1 if bias_correction else 0, current_stream()))


def adam_update(param, grad, exp_avg, exp_avg_sq, visible, lr, b1, b2, eps, N, M):
    """AR `_C.adamUpdate(param, grad, exp_avg, exp_avg_sq, visible bool[N], lr, b1, b2, eps, N, M)` (AR/rasterize_points.cu adamUpdate)."""
    _require(param, "param")
    for t, name in ((exp_avg, "exp_avg"), (exp_avg_sq, "exp_avg_sq")):
        _require(t, name)
    grad = grad.contiguous()
    vis = visible.contiguous()
    if vis.dtype != torch.bool and vis.dtype != torch.uint8:
        vis = vis != 0
    check(lib.lodgs_adam_step(param.data_ptr(), grad.data_ptr(), exp_avg.data_ptr(), exp_avg_sq.data_ptr(), None, 0, vis.data_ptr(), int(N),
                              int(M), float(lr), float(b1), float(b2), float(eps), 1, 0, current_stream()))


class SparseGaussianAdam(torch.optim.Adam):
    """AR/alt_gaussian_rasterization/__init__.py:246-271."""

    def __init__(self, params, lr, eps):
        super().__init__(params=params, lr=lr, eps=eps)

    @torch.no_grad()
    def step(self, visibility, N):
        self.step_overlapped(visibility, N, None)

    @torch.no_grad()
    def step_overlapped(self, visibility, N, wait):
        """step() for data-parallel training (not in the reference): `wait` = {param: work handle of its gradient's all-reduce}; the
        handle is waited for on the current stream right before that parameter's update, so earlier groups update while later
        gradients are still being reduced (lodgs.parallel.allreduce_gradients_async)."""
        for group in self.param_groups:
            lr, eps = group["lr"], group["eps"]
            assert len(group["params"]) == 1, "more than one tensor in group"
            param = group["params"][0]
            if param.grad is None:
                continue
            if wait is not None and wait.get(param) is not None:
                wait[param].wait()
            state = self.state[param]
            if len(state) == 0:
                state["step"] = torch.tensor(0.0, dtype=torch.float32)
                state["exp_avg"] = torch.zeros_like(param, memory_format=torch.preserve_format)
                state["exp_avg_sq"] = torch.zeros_like(param, memory_format=torch.preserve_format)
            M = param.numel() // N
            adam_update(param, param.grad, state["exp_avg"], state["exp_avg_sq"], visibility, lr, 0.9, 0.999, eps, N, M)
