"""Adam as one kernel launch per parameter group (host side of csrc/adam.cu).

`FusedAdam` IS a `torch.optim.Adam`: same constructor, same `param_groups`, same per-parameter
state keys (`step`, `exp_avg`, `exp_avg_sq`), so `state_dict()` / `load_state_dict()` interchange
with the stock optimizer and `ReduceLROnPlateau` drives it unchanged.  Only `step()` differs: the
reference's drivers build `torch.optim.Adam(model.parameters(), lr, weight_decay)`
(train.py:331-335, cross_validation.py:93-97) whose step walks ~250 small tensors; here the
parameters are cut into 16K-element chunks listed in a device table and updated by
`b2seg_adam_step` in one launch (28 B of HBM traffic per parameter element).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib as L

CHUNK = 16384
_CHUNK_DT = np.dtype([("p", "<u8"), ("g", "<u8"), ("m", "<u8"), ("v", "<u8"), ("n", "<i4"), ("pad", "<i4")])


class FusedAdam(torch.optim.Adam):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False, **kw):
        kw.pop("fused", None)
        kw.pop("foreach", None)
        super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=amsgrad, **kw)
        self._tables = {}

    def _table(self, gi, ps, gs, ms, vs):
        key = tuple(t.data_ptr() for ts in (ps, gs, ms, vs) for t in ts)
        hit = self._tables.get(gi)
        if hit is not None and hit[0] == key:
            return hit[1], hit[2]
        rows = []
        for p, g, m, v in zip(ps, gs, ms, vs):
            n = p.numel()
            for off in range(0, n, CHUNK):
                b = 4 * off
                rows.append((p.data_ptr() + b, g.data_ptr() + b, m.data_ptr() + b, v.data_ptr() + b,
                             min(CHUNK, n - off), 0))
        host = np.array(rows, dtype=_CHUNK_DT)
        dev = torch.from_numpy(host.view(np.uint8).copy()).to(ps[0].device)
        self._tables[gi] = (key, dev, len(rows))
        return dev, len(rows)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for gi, group in enumerate(self.param_groups):
            if group.get("amsgrad") or group.get("maximize") or group.get("capturable") or \
                    group.get("differentiable"):
                raise L.B2SegError("FusedAdam supports amsgrad=maximize=capturable=differentiable=False")
            by_step = {}
            for p in group["params"]:
                if p.grad is None:
                    continue
                g = p.grad
                if g.is_sparse or p.dtype != torch.float32 or g.dtype != torch.float32 or not p.is_cuda:
                    raise L.B2SegError("FusedAdam needs dense fp32 CUDA parameters and gradients")
                if not p.is_contiguous():
                    raise L.B2SegError("FusedAdam needs contiguous parameters")
                if not g.is_contiguous():
                    g = p.grad = g.contiguous()
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["step"] += 1
                by_step.setdefault(float(st["step"]), []).append((p, g, st["exp_avg"], st["exp_avg_sq"]))
            lr = group["lr"]
            lr = float(lr) if not torch.is_tensor(lr) else float(lr.item())
            b1, b2 = group["betas"]
            for si, (t, items) in enumerate(sorted(by_step.items())):
                ps, gs, ms, vs = zip(*items)
                dev, n = self._table((gi, si), ps, gs, ms, vs)
                L.check(L.lib().b2seg_adam_step(dev.data_ptr(), n, lr, float(b1), float(b2), float(group["eps"]),
                                                float(group["weight_decay"]), t, L.stream_ptr()))
        return loss


LAUNCHES_PER_GROUP = 1
