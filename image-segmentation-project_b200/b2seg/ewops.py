"""Planned ops for the HBM-bound layer kernels (section 4 of include/b2seg.h): BatchNorm
forward/backward, stem, max-pool, 1-channel convs, attention gate, squeeze-excite, packing."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib as L
from .convops import Op


def _ptr(v):
    if v is None:
        return None
    if isinstance(v, int):
        return v
    if torch.is_tensor(v):
        return v.data_ptr()
    return v.ptr()  # convops.View


def ew(kind, p=(), i=(), f=(), keep=()):
    """Create one planned op.  `p` entries may be tensors, Views, raw addresses or None."""
    d = L.EwDesc()
    d.kind = kind
    for k, v in enumerate(p):
        d.p[k] = _ptr(v)
    for k, v in enumerate(i):
        d.i[k] = int(v)
    for k, v in enumerate(f):
        d.f[k] = float(v)
    h = C.c_void_p()
    L.check(L.lib().b2seg_ew_create(C.byref(d), C.byref(h)))
    return Op(h, tuple(keep) + tuple(v for v in p if torch.is_tensor(v)), f"ew{kind}")


def partial_rows(kind, P, Cn) -> int:
    return L.lib().b2seg_ew_partial_rows(kind, int(P), int(Cn))


def stem_stat_rows(n, h, w) -> int:
    return L.lib().b2seg_stem_stat_rows(n, h, w)


def stem_wgrad_rows(n, h, w) -> int:
    return L.lib().b2seg_stem_wgrad_rows(n, h, w)
