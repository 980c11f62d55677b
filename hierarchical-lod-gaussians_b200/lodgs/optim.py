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
