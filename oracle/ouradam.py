"""TEST INFRASTRUCTURE ONLY — the reference's sparse Adam (scene/OurAdam.py) as a checker.

`load_reference()` imports the UNMODIFIED /root/reference/scene/OurAdam.py by file path (pure torch, no package import of
`scene/`), which works wherever /root/reference exists (the build container). On the GPU box the reference tree is
absent, so `RestatedAdam` restates `Adam.step(relevant)` op by op:

    Adam.step(relevant)                       scene/OurAdam.py:106-176  (lazy state, one `step` tensor per parameter)
    adam(...)                                 scene/OurAdam.py:179-248  (relevant.size(0) == 0 -> all rows, else the index list)
    _single_tensor_adam  (relevant rows)      scene/OurAdam.py:251-354  (non-capturable branch :329-354)
    _single_tensor_adam2 (all rows)           scene/OurAdam.py:357-452  (non-capturable branch :431-452)

It issues the same torch ops in the same order, so on one device it is BIT-IDENTICAL to the reference class; that is what
tests/test_ouradam_pin.py asserts (against the imported class when /root/reference is present, against
tests/golden/ouradam.npz — produced by the imported class, tests/golden/make_golden_adam.py — everywhere).
Only tests/, __graft_entry__.smoke() and bench.py's reference arm may import this file.
"""
import importlib.util
import math
import os

import torch

REF_PATH = "/root/reference/scene/OurAdam.py"


def load_reference():
    """-> the reference module (its `Adam` class, `_single_tensor_adam`, ...) or None when /root/reference is absent."""
    if not os.path.exists(REF_PATH):
        return None
    spec = importlib.util.spec_from_file_location("_ref_OurAdam", REF_PATH)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    # torch 2.11 renamed Optimizer._cuda_graph_capture_health_check (called at OurAdam.py:113) to
    # _accelerator_graph_capture_health_check; alias it on the loaded class so the unmodified step() runs on this torch
    if not hasattr(mod.Adam, "_cuda_graph_capture_health_check"):
        mod.Adam._cuda_graph_capture_health_check = getattr(mod.Adam, "_accelerator_graph_capture_health_check", lambda self: None)
    return mod


def single_tensor_adam(params, grads, exp_avgs, exp_avg_sqs, state_steps, *, beta1, beta2, lr, relevant, eps):
    """scene/OurAdam.py:251-354, amsgrad=False, weight_decay=0, maximize=False, capturable=False."""
    for i, parami in enumerate(params):
        grad = grads[i][relevant]                                   # :283
        exp_avg = exp_avgs[i][relevant]                             # :284
        exp_avg_sq = exp_avg_sqs[i][relevant]                       # :285
        param = parami[relevant]                                    # :286
        step_t = state_steps[i]
        step_t += 1                                                 # :293
        exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)             # :300
        exp_avg_sq.mul_(beta2).addcmul_(grad, grad.conj(), value=1 - beta2)   # :301
        step = step_t.item()                                        # :330
        bias_correction1 = 1 - beta1 ** step
        bias_correction2 = 1 - beta2 ** step
        step_size = lr / bias_correction1                           # :335
        bias_correction2_sqrt = math.sqrt(bias_correction2)         # :338
        denom = (exp_avg_sq.sqrt() / bias_correction2_sqrt).add_(eps)   # :346
        param.addcdiv_(exp_avg, denom, value=-step_size)            # :349
        exp_avgs[i][relevant] = exp_avg                             # :352
        exp_avg_sqs[i][relevant] = exp_avg_sq                       # :353
        parami[relevant] = param                                    # :355


def single_tensor_adam2(params, grads, exp_avgs, exp_avg_sqs, state_steps, *, beta1, beta2, lr, eps):
    """scene/OurAdam.py:357-452 (every row, in place)."""
    for i, param in enumerate(params):
        grad, exp_avg, exp_avg_sq, step_t = grads[i], exp_avgs[i], exp_avg_sqs[i], state_steps[i]
        step_t += 1
        exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
        exp_avg_sq.mul_(beta2).addcmul_(grad, grad.conj(), value=1 - beta2)
        step = step_t.item()
        bias_correction1 = 1 - beta1 ** step
        bias_correction2 = 1 - beta2 ** step
        step_size = lr / bias_correction1
        denom = (exp_avg_sq.sqrt() / math.sqrt(bias_correction2)).add_(eps)
        param.addcdiv_(exp_avg, denom, value=-step_size)


class RestatedAdam(torch.optim.Optimizer):
    """Same constructor and `step(relevant)` as scene/OurAdam.py `Adam` (:74-176) for the options the training scripts use."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))

    @torch.no_grad()
    def step(self, relevant):
        for group in self.param_groups:
            ps, gs, ms, vs, steps = [], [], [], [], []
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                state = self.state[p]
                if len(state) == 0:                                 # :139-149
                    state["step"] = torch.tensor(0.)
                    state["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    state["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                ps.append(p)
                gs.append(p.grad)
                ms.append(state["exp_avg"])
                vs.append(state["exp_avg_sq"])
                steps.append(state["step"])
            if relevant.size(0) == 0:                               # :214-229
                single_tensor_adam2(ps, gs, ms, vs, steps, beta1=beta1, beta2=beta2, lr=group["lr"], eps=group["eps"])
            else:
                single_tensor_adam(ps, gs, ms, vs, steps, beta1=beta1, beta2=beta2, lr=group["lr"], relevant=relevant,
                                   eps=group["eps"])


def reference_or_restated():
    """The class bench.py's reference arm optimises with: the real one where the reference tree exists, else the restatement
    (pinned bit-identical to it by tests/test_ouradam_pin.py). -> (class, "reference" | "restated")."""
    mod = load_reference()
    if mod is not None:
        return mod.Adam, "reference"
    return RestatedAdam, "restated"


# ---- shared by the golden generator and the tests: one seeded multi-step scenario per (width layout, eps)
SHAPES = {"xyz": (3,), "f_dc": (1, 3), "f_rest": (15, 3), "opacity": (1,), "scaling": (3,), "rotation": (4,)}   # 59 floats per row
LRS = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 1.25e-4, "opacity": 0.05, "scaling": 5e-3, "rotation": 1e-3}


def scenario(rows, eps, seed, device="cpu"):
    """-> (params dict of leaf tensors, list of (relevant i64, grads dict)) for 4 steps: three index lists (sorted, unsorted
    permutation, single row) and one empty list (= all rows, OurAdam.py:214)."""
    g = torch.Generator().manual_seed(seed)
    params = {k: torch.randn((rows,) + s, generator=g).to(device).requires_grad_(True) for k, s in SHAPES.items()}
    steps = []
    for it in range(4):
        grads = {k: (torch.randn((rows,) + s, generator=g) * (10.0 ** -(it + 1))).to(device) for k, s in SHAPES.items()}
        if it == 0:
            rel = torch.randperm(rows, generator=g)[: rows // 3].sort().values
        elif it == 1:
            rel = torch.randperm(rows, generator=g)[: rows // 2]
        elif it == 2:
            rel = torch.tensor([rows - 1])
        else:
            rel = torch.empty(0, dtype=torch.int64)
        steps.append((rel.to(device), grads))
    return params, steps


def run_scenario(adam_cls, rows, eps, seed, device="cpu"):
    """Runs the scenario through an optimizer class with the reference's six parameter groups (gaussian_model.py:925-937).
    -> {name: (param, exp_avg, exp_avg_sq)} after the last step."""
    params, steps = scenario(rows, eps, seed, device)
    groups = [{"params": [params[k]], "lr": LRS[k], "name": k} for k in SHAPES]
    opt = adam_cls(groups, lr=0.0, eps=eps)
    for rel, grads in steps:
        for k in SHAPES:
            params[k].grad = grads[k]
        opt.step(rel)
    return {k: (params[k].detach(), opt.state[params[k]]["exp_avg"], opt.state[params[k]]["exp_avg_sq"]) for k in SHAPES}
