#!/usr/bin/env python3
"""Times lodgs_sh_adam_from_rgb (csrc/sh_adam.cu) in isolation next to lodgs_adam_step on a materialised [P,48] gradient: 1 M rows, 88 % of
them selected, index list / mask / dense, with and without bias correction, 1 and 8 views. Prints one JSON line."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "hierarchical-lod-gaussians_b200"), ROOT]
import torch  # noqa: E402

import lodgs  # noqa: E402
from lodgs.optim import sh_adam_from_rgb  # noqa: E402

P = 1_000_000
g = torch.Generator(device="cuda").manual_seed(0)
means = torch.randn(P, 3, device="cuda", generator=g) * 10
shs = torch.randn(P, 16, 3, device="cuda", generator=g)
m, v = torch.zeros_like(shs), torch.zeros_like(shs)
sel = torch.rand(P, device="cuda", generator=g) < 0.88
rel = sel.nonzero().reshape(-1)
scale = float(os.environ.get("GRAD_SCALE", "1e-7"))


def timed(fn, it=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it


out = {"rows": P, "selected": int(sel.sum()), "grad_scale": scale}
for views in (1, 8):
    drgb = torch.randn(views, P, 3, device="cuda", generator=g) * scale * sel[None, :, None]
    campos = torch.randn(views, 3, device="cuda", generator=g) * 30
    for bias, rows in ((True, "list"), (False, "mask"), (True, "dense"), (False, "dense")):
        kw = {"relevant": rel} if rows == "list" else ({"visible": sel} if rows == "mask" else {})
        ms = timed(lambda: sh_adam_from_rgb(drgb, campos, means, shs[:, :1], m[:, :1], v[:, :1], shs[:, 1:], m[:, 1:], v[:, 1:], 3, 1e-6, 1e-7, step=5,
                                            bias_correction=bias, **kw))
        n = P if rows == "dense" else int(sel.sum())
        out["sh_adam views=%d bias=%d %s" % (views, bias, rows)] = {"ms": round(ms, 4), "GBps_24B_per_scalar": round(n * 48 * 24 / ms / 1e6, 1)}
grad = torch.randn(P, 48, device="cuda", generator=g) * scale * sel[:, None]
p2, m2, v2 = shs.view(P, 48).clone(), torch.zeros(P, 48, device="cuda"), torch.zeros(P, 48, device="cuda")
for bias, rows in ((True, "list"), (False, "mask"), (True, "dense")):
    ms = timed(lambda: lodgs.check(lodgs.lib.lodgs_adam_step(p2.data_ptr(), grad.data_ptr(), m2.data_ptr(), v2.data_ptr(), rel.data_ptr() if rows == "list" else None,
                                                              rel.numel() if rows == "list" else 0, sel.data_ptr() if rows == "mask" else None, P, 48, 1e-6, 0.9,
                                                              0.999, 1e-15, 5, 1 if bias else 0, lodgs.current_stream())))
    n = P if rows == "dense" else int(sel.sum())
    out["adam_rows bias=%d %s" % (bias, rows)] = {"ms": round(ms, 4), "GBps_28B_per_scalar": round(n * 48 * 28 / ms / 1e6, 1)}
print(json.dumps(out))
