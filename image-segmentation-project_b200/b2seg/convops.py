"""Host-side descriptors for the tcgen05 convolution kernels.

Every convolution flavour of the U-Net (advanced_models.py:76-160, torchvision
BasicBlock/Bottleneck) is expressed as a *tap-gather GEMM* (forward and data
gradient) or a *pixel-reduction GEMM* (weight gradient) over NHWC bf16 views:

  3x3 s1 p1 / 1x1 s1   one source view, taps shift the pixel coordinate
  3x3 s2 p1 / 1x1 s2   four stride-2 phase views of the input (no im2col, no gather)
  ConvTranspose 2x2 s2 one tap, N = 4*Cout, pixel-shuffle store into the concat slice
  dgrad                same kernel, weights repacked [tap][cin][cout], taps negated
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib as L


class View:
    """A channel slice [c0, c0+c) of an NHWC bf16 buffer (channel stride 1)."""

    __slots__ = ("t", "c0", "c", "n", "h", "w", "sn", "sh", "sw", "off")

    def __init__(self, t: torch.Tensor, c0: int = 0, c: int | None = None):
        assert t.dtype == torch.bfloat16 and t.dim() == 4 and t.stride(3) == 1, (t.dtype, t.shape, t.stride())
        self.t = t
        self.c0 = c0
        self.c = t.shape[3] - c0 if c is None else c
        self.n, self.h, self.w = t.shape[0], t.shape[1], t.shape[2]
        self.sn, self.sh, self.sw = t.stride(0), t.stride(1), t.stride(2)
        self.off = 0  # extra element offset (phases)

    def ptr(self) -> int:
        return self.t.data_ptr() + 2 * (self.c0 + self.off)

    def t4(self) -> L.Tensor4:
        return L.Tensor4(self.ptr(), self.n, self.h, self.w, self.c, self.sn, self.sh, self.sw)

    def phase(self, py: int, px: int) -> "View":
        """Pixels (2i+py, 2j+px) as a view of half resolution."""
        assert self.h % 2 == 0 and self.w % 2 == 0, "stride-2 layers need even H, W"
        v = View.__new__(View)
        v.t, v.c0, v.c = self.t, self.c0, self.c
        v.n, v.h, v.w = self.n, self.h // 2, self.w // 2
        v.sn, v.sh, v.sw = self.sn, 2 * self.sh, 2 * self.sw
        v.off = self.off + py * self.sh + px * self.sw
        return v

    def torch(self) -> torch.Tensor:
        assert self.off == 0
        return self.t[..., self.c0:self.c0 + self.c]


class Op:
    """Owns one planned op of libb2seg (and keeps its tensors alive)."""

    def __init__(self, handle, keep, kind, flops=0.0):
        self.handle = handle
        self.keep = keep
        self.kind = kind
        self.flops = float(flops)  # algorithmic 2*M*N*K of a tensor-core op (bench roofline)
        self.desc = ""
        self.side = False  # True: nothing later in the sequence reads its output (see Seq)

    def run(self):
        L.check(L.lib().b2seg_op_run(self.handle, L.stream_ptr()))

    @property
    def launches(self):
        return L.lib().b2seg_op_launches(self.handle)

    def __del__(self):
        try:
            if self.handle:
                L.lib().b2seg_op_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


class Seq:
    """Ordered list of ops replayed with one C call."""

    def __init__(self, ops, side_kinds=(), join_before_kinds=()):
        """`side_kinds`: kind prefixes of ops (plus every op with `op.side` set) the graph replay
        runs on a side branch
        (b2seg_seq_append_lane): weight gradients, which nothing later in the list reads, and the
        weight repack, which is joined before the first op whose kind starts with one of
        `join_before_kinds` (b2seg_seq_join)."""
        self.ops = list(ops)
        h = C.c_void_p()
        L.check(L.lib().b2seg_seq_create(C.byref(h)))
        self.handle = h
        pending = False
        for op in self.ops:
            lane = 1 if op.side or any(op.kind.startswith(k) for k in side_kinds) else 0
            if pending and lane == 0 and any(op.kind.startswith(k) for k in join_before_kinds):
                L.check(L.lib().b2seg_seq_join(self.handle))
                pending = False
            pending = pending or lane == 1
            L.check(L.lib().b2seg_seq_append_lane(self.handle, op.handle, lane))

    def run(self):
        L.check(L.lib().b2seg_seq_run(self.handle, L.stream_ptr()))

    def run_timed(self):
        """Replay with an event pair around every op; returns per-op milliseconds (synchronises)."""
        ms = (C.c_float * len(self.ops))()
        L.check(L.lib().b2seg_seq_run_timed(self.handle, L.stream_ptr(), ms, len(self.ops)))
        return list(ms)

    @property
    def launches(self):
        return L.lib().b2seg_seq_launches(self.handle)

    def __del__(self):
        try:
            if self.handle:
                L.lib().b2seg_seq_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


# ----------------------------------------------------------- weight packing ----
def pack_conv_fprop(w: torch.Tensor) -> torch.Tensor:
    """OIHW fp32 -> [kh*kw][O][I] bf16."""
    o, i, kh, kw = w.shape
    return w.detach().permute(2, 3, 0, 1).reshape(kh * kw, o, i).to(torch.bfloat16).contiguous()


def pack_conv_dgrad(w: torch.Tensor) -> torch.Tensor:
    """OIHW fp32 -> [kh*kw][I][O] bf16."""
    o, i, kh, kw = w.shape
    return w.detach().permute(2, 3, 1, 0).reshape(kh * kw, i, o).to(torch.bfloat16).contiguous()


def pack_convT_fprop(w: torch.Tensor) -> torch.Tensor:
    """ConvTranspose IOHW (k=2) fp32 -> [1][q*O + o][I] bf16, q = qy*2+qx."""
    i, o, kh, kw = w.shape
    return w.detach().permute(2, 3, 1, 0).reshape(1, kh * kw * o, i).to(torch.bfloat16).contiguous()


def pack_convT_dgrad(w: torch.Tensor) -> torch.Tensor:
    """ConvTranspose IOHW fp32 -> [q][I][O] bf16."""
    i, o, kh, kw = w.shape
    return w.detach().permute(2, 3, 0, 1).reshape(kh * kw, i, o).to(torch.bfloat16).contiguous()


# ------------------------------------------------------------------ tap tables ----
def _s2_tap(r):
    """Filter row r of a 3x3/s2/p1 conv -> (phase, offset in the phase view)."""
    return ((r - 1) & 1, (r - 1) >> 1)  # r=0 -> (1,-1), r=1 -> (0,0), r=2 -> (1,0)


def fprop_taps(k: int, stride: int):
    taps = []
    for r in range(k):
        for s in range(k):
            if stride == 1:
                taps.append((0, r - k // 2, s - k // 2, r * k + s))
            else:
                if k == 1:
                    taps.append((0, 0, 0, 0))
                else:
                    pr, dh = _s2_tap(r)
                    ps, dw = _s2_tap(s)
                    taps.append((pr * 2 + ps, dh, dw, r * k + s))
    return taps


def _fill_sub(sub, taps, out_offset=0):
    sub.ntaps = len(taps)
    for i, (src, dh, dw, wt) in enumerate(taps):
        sub.taps[i] = L.Tap(src, dh, dw, wt)
    sub.out_offset = out_offset


def _create_tapgemm(d, keep, kind):
    h = C.c_void_p()
    L.check(L.lib().b2seg_tapgemm_create(C.byref(d), C.byref(h)))
    taps = sum(d.sub[s].ntaps for s in range(d.nsub))
    op = Op(h, keep, kind, flops=2.0 * d.on * d.oh * d.ow * d.ngemm * d.cin * taps)
    op.desc = f"M={d.on * d.oh * d.ow} ({d.on}x{d.oh}x{d.ow}) N={d.ngemm} K={taps}x{d.cin} nsub={d.nsub}"
    return op


def _srcs(d, views):
    d.nsrc = len(views)
    for i, v in enumerate(views):
        d.src[i] = v.t4()


def stat_rows(n, h, w) -> int:
    return L.lib().b2seg_tapgemm_stat_rows(n, h, w)


def conv_fprop(x: View, wp: torch.Tensor, k: int, stride: int, out: View, bias=None, stats=None):
    """out = conv(x, W) (+bias); optional per-channel partial (sum, sumsq) rows in `stats` (fp32
    [rows][2][C]), or -- an int64 `stats` [2][C], zeroed by the caller -- fixed-point sums added with atomics."""
    d = L.TapGemmDesc()
    views = [x] if stride == 1 else ([x.phase(0, 0)] if k == 1 else
                                     [x.phase(a, b) for a in (0, 1) for b in (0, 1)])
    _srcs(d, views)
    d.nsub = 1
    _fill_sub(d.sub[0], fprop_taps(k, stride))
    d.weights = wp.data_ptr()
    d.wtaps, d.ngemm, d.cin = wp.shape[0], wp.shape[1], wp.shape[2]
    assert d.cin == x.c and d.ngemm == out.c and d.wtaps == k * k
    d.on, d.oh, d.ow = out.n, out.h, out.w
    assert (out.h, out.w) == (x.h // stride, x.w // stride)
    d.out, d.osn, d.osh, d.osw = out.ptr(), out.sn, out.sh, out.sw
    d.cq = 0
    d.bias = bias.data_ptr() if bias is not None else None
    d.stats = stats.data_ptr() if stats is not None else None
    d.stats_fixed = 0
    if stats is not None and stats.dtype == torch.int64:
        assert stats.numel() >= 2 * out.c
        d.stats_fixed = 1
    elif stats is not None:
        assert stats.dtype == torch.float32 and stats.numel() >= stat_rows(out.n, out.h, out.w) * 2 * out.c
    return _create_tapgemm(d, (x.t, wp, out.t, bias, stats), f"conv{k}x{k}s{stride}")


def conv_dgrad(dy: View, wd: torch.Tensor, k: int, stride: int, dx: View, accumulate=False):
    """dx (+)= conv_transpose(dy, W).  wd = pack_conv_dgrad(W)."""
    d = L.TapGemmDesc()
    _srcs(d, [dy])
    d.weights = wd.data_ptr()
    d.wtaps, d.ngemm, d.cin = wd.shape[0], wd.shape[1], wd.shape[2]
    assert d.cin == dy.c and d.ngemm == dx.c
    d.out = dx.ptr()
    d.cq = 0
    d.accumulate = 1 if accumulate else 0
    if stride == 1:
        d.nsub = 1
        c = k // 2
        _fill_sub(d.sub[0], [(0, c - r, c - s, r * k + s) for r in range(k) for s in range(k)])
        d.on, d.oh, d.ow = dx.n, dx.h, dx.w
        d.osn, d.osh, d.osw = dx.sn, dx.sh, dx.sw
    else:
        # output phase (ph,pw) of dx gathers the filter rows r with r = ph+1 (mod 2)
        d.on, d.oh, d.ow = dx.n, dx.h // 2, dx.w // 2
        d.osn, d.osh, d.osw = dx.sn, 2 * dx.sh, 2 * dx.sw
        if k == 1:
            d.nsub = 1
            _fill_sub(d.sub[0], [(0, 0, 0, 0)])
        else:
            rows = {0: [(1, 0)], 1: [(0, 1), (2, 0)]}  # phase -> [(r, offset into dy)]
            d.nsub = 4
            for ph in (0, 1):
                for pw in (0, 1):
                    taps = [(0, dh, dw, r * 3 + s) for (r, dh) in rows[ph] for (s, dw) in rows[pw]]
                    _fill_sub(d.sub[ph * 2 + pw], taps, ph * dx.sh + pw * dx.sw)
    return _create_tapgemm(d, (dy.t, wd, dx.t), f"dgrad{k}x{k}s{stride}")


def convT_fprop(x: View, wp: torch.Tensor, out: View, bias=None):
    """ConvTranspose2d(k=2,s=2): out[2h+qy, 2w+qx, :] = x[h,w,:] @ W[:,:,qy,qx] + b, written
    straight into `out` (which may be a channel slice of the concat buffer)."""
    d = L.TapGemmDesc()
    _srcs(d, [x])
    d.nsub = 1
    _fill_sub(d.sub[0], [(0, 0, 0, 0)])
    d.weights = wp.data_ptr()
    d.wtaps, d.ngemm, d.cin = 1, wp.shape[1], wp.shape[2]
    cout = d.ngemm // 4
    assert d.cin == x.c and cout == out.c and (out.h, out.w) == (2 * x.h, 2 * x.w)
    d.on, d.oh, d.ow = x.n, x.h, x.w
    d.out, d.osn, d.osh, d.osw = out.ptr(), out.sn, 2 * out.sh, 2 * out.sw
    d.cq, d.oqh, d.oqw = cout, out.sh, out.sw
    d.bias = bias.data_ptr() if bias is not None else None
    return _create_tapgemm(d, (x.t, wp, out.t, bias), "convT2x2")


def convT_dgrad(dy: View, wd: torch.Tensor, dx: View, accumulate=False):
    """dx[h,w,:] = sum_q dy[2h+qy, 2w+qx, :] @ W[:,:,q]^T."""
    d = L.TapGemmDesc()
    _srcs(d, [dy.phase(a, b) for a in (0, 1) for b in (0, 1)])
    d.nsub = 1
    _fill_sub(d.sub[0], [(q, 0, 0, q) for q in range(4)])
    d.weights = wd.data_ptr()
    d.wtaps, d.ngemm, d.cin = 4, wd.shape[1], wd.shape[2]
    assert d.cin == dy.c and d.ngemm == dx.c
    d.on, d.oh, d.ow = dx.n, dx.h, dx.w
    d.out, d.osn, d.osh, d.osw = dx.ptr(), dx.sn, dx.sh, dx.sw
    d.cq = 0
    d.accumulate = 1 if accumulate else 0
    return _create_tapgemm(d, (dy.t, wd, dx.t), "convT_dgrad")


# ---------------------------------------------------------------------- wgrad ----
def wgrad_workspace_bytes(a_ch, b_ch, ntaps, n, h, w) -> int:
    return L.lib().b2seg_wgrad_workspace_bytes(a_ch, b_ch, ntaps, n, h, w)


def _create_wgrad(d, keep, kind):
    h = C.c_void_p()
    L.check(L.lib().b2seg_wgrad_create(C.byref(d), C.byref(h)))
    op = Op(h, keep, kind, flops=2.0 * d.pn * d.ph * d.pw * d.a_src[0].c * d.b_src.c * d.ntaps)
    op.desc = f"pix={d.pn * d.ph * d.pw} ({d.pn}x{d.ph}x{d.pw}) A={d.ntaps}x{d.a_src[0].c} B={d.b_src.c}"
    return op


def conv_wgrad(x: View, dy: View, k: int, stride: int, grad: torch.Tensor, ws: torch.Tensor,
               accumulate=False):
    """grad[O,I,kh,kw] (fp32) (+)= sum_pix dy[pix,o] * x[pix*stride + tap, i]."""
    d = L.WgradDesc()
    views = [x] if stride == 1 else ([x.phase(0, 0)] if k == 1 else
                                     [x.phase(a, b) for a in (0, 1) for b in (0, 1)])
    d.na_src = len(views)
    for i, v in enumerate(views):
        d.a_src[i] = v.t4()
    d.b_src = dy.t4()
    taps = fprop_taps(k, stride)
    d.ntaps = len(taps)
    for i, (src, dh, dw, wt) in enumerate(taps):
        d.taps[i] = L.WgradTap(src, dh, dw, 0, wt)
    d.pn, d.ph, d.pw = dy.n, dy.h, dy.w
    assert grad.dtype == torch.float32 and grad.is_contiguous()
    assert tuple(grad.shape) == (dy.c, x.c, k, k), (grad.shape, dy.c, x.c, k)
    d.grad = grad.data_ptr()
    d.row_stride, d.col_stride = k * k, x.c * k * k
    d.accumulate = 1 if accumulate else 0
    d.workspace, d.workspace_bytes = ws.data_ptr(), ws.numel() * ws.element_size()
    return _create_wgrad(d, (x.t, dy.t, grad, ws), f"wgrad{k}x{k}s{stride}")


def convT_wgrad(x: View, dy: View, grad: torch.Tensor, ws: torch.Tensor, accumulate=False):
    """grad[I,O,2,2] (fp32) (+)= sum_pix x[pix,i] * dy[2pix+q, o]."""
    d = L.WgradDesc()
    views = [dy.phase(a, b) for a in (0, 1) for b in (0, 1)]
    d.na_src = 4
    for i, v in enumerate(views):
        d.a_src[i] = v.t4()
    d.b_src = x.t4()
    d.ntaps = 4
    for q in range(4):
        d.taps[q] = L.WgradTap(q, 0, 0, 0, q)
    d.pn, d.ph, d.pw = x.n, x.h, x.w
    assert grad.dtype == torch.float32 and grad.is_contiguous()
    assert tuple(grad.shape) == (x.c, dy.c, 2, 2)
    d.grad = grad.data_ptr()
    d.row_stride, d.col_stride = 4, dy.c * 4
    d.accumulate = 1 if accumulate else 0
    d.workspace, d.workspace_bytes = ws.data_ptr(), ws.numel() * ws.element_size()
    return _create_wgrad(d, (x.t, dy.t, grad, ws), "wgradT")
