"""Fused segmentation loss / metrics on the GPU (host side of csrc/seg_loss.cu).

`FusedSegLoss` is the autograd entry the reference-compatible loss classes in
``losses.py`` delegate to; one forward launch produces the loss, the gradient
coefficients and the confusion counts, one backward launch the logits gradient.
No host synchronisation happens unless the caller reads a value.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib as L


def make_spec(**kw) -> L.LossSpec:
    s = L.LossSpec()
    s.dice_smooth = 1.0
    s.tv_alpha = s.tv_beta = 0.5
    s.tv_smooth = 1.0
    s.tv_gamma = 1.0
    s.ss_w, s.ss_eps = 0.5, 1e-7
    s.bce_beta = 1.0
    s.focal_alpha, s.focal_gamma = 0.25, 2.0
    for k, v in kw.items():
        if not hasattr(s, k):
            raise KeyError(k)
        setattr(s, k, v)
    return s


class _Scratch:
    """Reduction workspace (ticket counter + block partials; zero-initialised once, kernels leave it
    zeroed), one per (device, stream): reductions in flight on different streams or from different
    threads never share tickets or partials."""
    _by_dev = {}

    @classmethod
    def get(cls, device) -> torch.Tensor:
        key = (device.type, device.index if device.index is not None else torch.cuda.current_device(),
               int(torch.cuda.current_stream(device).cuda_stream))
        ws = cls._by_dev.get(key)
        if ws is None:
            n = L.lib().b2seg_seg_reduce_workspace_bytes()
            ws = torch.zeros((n + 7) // 8, dtype=torch.int64, device=device)
            cls._by_dev[key] = ws
        return ws


def _flat_f32(t: torch.Tensor) -> torch.Tensor:
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def loss_forward(logits, target, spec: L.LossSpec, weight_map=None, group=None):
    """Returns (loss[0-dim f32], stats[16 f64], coef[8 f32]) — all device tensors.
    `group`: a torch.distributed process group of data-parallel replicas; the eight sums and the
    pixel count are all-reduced between the reduction and the loss evaluation, so the value and
    the gradient coefficients are those of the GLOBAL batch (exact single-process semantics for
    Dice / Tversky, whose ratios do not decompose over shards)."""
    L.require_cuda(logits, target, weight_map)
    x, t = _flat_f32(logits), _flat_f32(target)
    if x.numel() != t.numel():
        raise ValueError(f"logits {tuple(logits.shape)} and target {tuple(target.shape)} differ in size")
    if x.numel() == 0:
        raise ValueError("empty input to segmentation loss")
    dev = x.device
    out = torch.empty(L.STATS_LEN, dtype=torch.float64, device=dev)
    small = torch.empty(1 + L.COEF_LEN, dtype=torch.float32, device=dev)
    wm = _flat_f32(weight_map) if weight_map is not None else None
    L.check(L.lib().b2seg_seg_loss_fwd(
        x.data_ptr(), t.data_ptr(), wm.data_ptr() if wm is not None else None, x.numel(),
        C.byref(spec), _Scratch.get(dev).data_ptr(), out.data_ptr(), small.data_ptr(),
        small.data_ptr() + 4, L.stream_ptr()))
    if group is not None:
        import torch.distributed as dist
        dist.all_reduce(out[:L.S_N + 1], op=dist.ReduceOp.SUM, group=group)
        L.check(L.lib().b2seg_seg_loss_finish(C.byref(spec), 0, out.data_ptr(), small.data_ptr(),
                                              small.data_ptr() + 4, L.stream_ptr()))
    return small[0], out, small[1:], (x, t, wm)


def loss_backward(saved, coef, grad_out, shape):
    x, t, wm = saved
    g = torch.empty_like(x)
    go = grad_out.contiguous().float() if grad_out is not None else None
    L.check(L.lib().b2seg_seg_loss_bwd(
        x.data_ptr(), t.data_ptr(), wm.data_ptr() if wm is not None else None, x.numel(),
        coef.data_ptr(), go.data_ptr() if go is not None else None, g.data_ptr(), L.stream_ptr()))
    return g.view(shape)


class FusedSegLoss(torch.autograd.Function):
    """loss, stats = FusedSegLoss.apply(logits, target, spec, weight_map)"""

    @staticmethod
    def forward(ctx, logits, target, spec, weight_map=None, group=None):
        loss, stats, coef, saved = loss_forward(logits, target, spec, weight_map, group)
        ctx.saved = saved
        ctx.coef = coef
        ctx.shape = logits.shape
        ctx.mark_non_differentiable(stats)
        return loss, stats

    @staticmethod
    def backward(ctx, grad_loss, _grad_stats):
        g = loss_backward(ctx.saved, ctx.coef, grad_loss, ctx.shape)
        return g, None, None, None, None


def metrics_from_probs(probs, target) -> torch.Tensor:
    """Device stats[16] (f64) for utils.calculate_metrics semantics on probabilities."""
    L.require_cuda(probs, target)
    p, t = _flat_f32(probs), _flat_f32(target)
    if p.numel() != t.numel() or p.numel() == 0:
        raise ValueError("calculate_metrics: pred/target size mismatch or empty")
    out = torch.empty(L.STATS_LEN, dtype=torch.float64, device=p.device)
    L.check(L.lib().b2seg_seg_metrics_from_probs(p.data_ptr(), t.data_ptr(), p.numel(),
                                                 _Scratch.get(p.device).data_ptr(), out.data_ptr(),
                                                 L.stream_ptr()))
    return out


def _planes_f32(t: torch.Tensor):
    if t.dim() < 2:
        raise ValueError(f"expected [..., H, W] planes, got {tuple(t.shape)}")
    t = _flat_f32(t)
    h, w = t.shape[-2], t.shape[-1]
    return t, t.numel() // (h * w), h, w


def boundary_weight_map(target, theta0: int = 3, theta: int = 5, gauss=None, gain: float = 10.0) -> torch.Tensor:
    """1 + gain * gauss(dilate(t) - erode(t)) of BoundaryLoss (losses.py:283-301), one launch.
    `gauss`: the theta x theta kernel as a CPU float32 tensor (BoundaryLoss._gaussian_kernel)."""
    L.require_cuda(target)
    t, planes, h, w = _planes_f32(target)
    k = gauss.detach().to("cpu", torch.float32).contiguous()
    if tuple(k.shape) != (theta, theta):
        raise ValueError(f"gaussian kernel must be {theta} x {theta}, got {tuple(k.shape)}")
    out = torch.empty_like(t)
    L.check(L.lib().b2seg_loss_boundary_weight_map(
        t.data_ptr(), out.data_ptr(), planes, h, w, int(theta0), int(theta),
        C.cast(k.data_ptr(), C.POINTER(C.c_float)), float(gain), L.stream_ptr()))
    return out


def distance_transform(src, mode: int = 0, alpha: float = 1.0) -> torch.Tensor:
    """scipy.ndimage.distance_transform_edt per [H, W] plane of `src`, on the device.
    mode 0: the distances; mode 1: 1 + alpha * d / (max_plane(d) + 1e-7) (losses.py:203-232)."""
    L.require_cuda(src)
    t, planes, h, w = _planes_f32(src)
    ws = torch.empty(L.lib().b2seg_loss_edt_workspace_bytes(planes, h, w), dtype=torch.uint8, device=t.device)
    out = torch.empty_like(t)
    L.check(L.lib().b2seg_loss_edt_f32(t.data_ptr(), out.data_ptr(), planes, h, w, int(mode), float(alpha),
                                       ws.data_ptr(), L.stream_ptr()))
    return out


def metrics_dict(stats_host, eps: float = 1e-7) -> dict:
    """utils.py:136-151 from a host copy of the stats block."""
    tp, fp, fn, tn = (float(stats_host[i]) for i in (L.S_TP, L.S_FP, L.S_FN, L.S_TN))
    precision = tp / (tp + fp + eps)
    recall = tp / (tp + fn + eps)
    return {
        "precision": precision,
        "recall": recall,
        "f1": 2 * precision * recall / (precision + recall + eps),
        "iou": tp / (tp + fp + fn + eps),
        "accuracy": (tp + tn) / (tp + tn + fp + fn + eps),
    }


# kernel launches one loss forward / backward enqueues (bench accounting; csrc/seg_loss.cu)
LAUNCHES_FWD = 1
LAUNCHES_BWD = 1
