"""GPU input pipeline (host side of csrc/preproc.cu): the arithmetic of
CellSegmentationDataset.__getitem__ (reference dataset.py:30-66) on batches of uint8 planes that
are already on the device.

    images_f32, masks_f32 = dataset_items(images_u8 [N,H,W], masks_u8 [N,H,W], (256, 256))

cv2.resize coefficient tables are computed once per (source size, target size) with OpenCV's own
formulas (modules/imgproc/src/resize.cpp) and cached on the device.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np
import torch

from . import _lib as L

NEAREST, LINEAR, AREA = 0, 1, 3   # cv2.INTER_NEAREST / INTER_LINEAR / INTER_AREA
_F = np.float32
_tables = {}


def _dev(arr, device):
    return torch.from_numpy(np.ascontiguousarray(arr)).to(device)


def _linear_axis(ssize, dsize, area_mode):
    scale, inv = ssize / dsize, dsize / ssize
    ofs = np.zeros(dsize, np.int32)
    coef = np.zeros((dsize, 2), np.int16)
    for d in range(dsize):
        if not area_mode:
            fx = _F((d + 0.5) * scale - 0.5)
            sx = int(math.floor(fx))
            fx = _F(fx - _F(sx))
        else:
            sx = int(math.floor(d * scale))
            fx = _F((d + 1) - (sx + 1) * inv)
            fx = _F(0.0) if fx <= 0 else _F(fx - _F(math.floor(fx)))
        if sx < 0:
            fx, sx = _F(0.0), 0
        if sx >= ssize - 1:
            fx, sx = _F(0.0), ssize - 1
        ofs[d] = sx
        coef[d, 0] = int(np.clip(np.rint(_F(_F(1.0) - fx) * _F(2048.0)), -32768, 32767))
        coef[d, 1] = int(np.clip(np.rint(_F(fx) * _F(2048.0)), -32768, 32767))
    return ofs, coef


def _area_axis(ssize, dsize):
    scale = ssize / dsize
    start, si, alpha = [0], [], []
    for dx in range(dsize):
        fsx1 = dx * scale
        fsx2 = fsx1 + scale
        cell = min(scale, ssize - fsx1)
        sx1, sx2 = math.ceil(fsx1), math.floor(fsx2)
        sx2 = min(sx2, ssize - 1)
        sx1 = min(sx1, sx2)
        if sx1 - fsx1 > 1e-3:
            si.append(sx1 - 1)
            alpha.append(_F((sx1 - fsx1) / cell))
        for sx in range(sx1, sx2):
            si.append(sx)
            alpha.append(_F(1.0 / cell))
        if fsx2 - sx2 > 1e-3:
            si.append(sx2)
            alpha.append(_F(min(min(fsx2 - sx2, 1.0), cell) / cell))
        start.append(len(si))
    return np.array(start, np.int32), np.array(si, np.int32), np.array(alpha, np.float32)


def _nearest_axis(ssize, dsize):
    inv = 1.0 / (dsize / ssize)
    return np.minimum(np.floor(np.arange(dsize) * inv).astype(np.int64), ssize - 1).astype(np.int32)


def _table(kind, sh, sw, dh, dw, device):
    key = (kind, sh, sw, dh, dw, str(device))
    t = _tables.get(key)
    if t is None:
        if kind == "nearest":
            t = (_dev(_nearest_axis(sw, dw), device), _dev(_nearest_axis(sh, dh), device))
        elif kind in ("linear", "linear_area"):
            xo, xa = _linear_axis(sw, dw, kind == "linear_area")
            yo, ya = _linear_axis(sh, dh, kind == "linear_area")
            t = (_dev(xo, device), _dev(xa, device), _dev(yo, device), _dev(ya, device))
        else:
            t = tuple(_dev(a, device) for a in (*_area_axis(sw, dw), *_area_axis(sh, dh)))
        _tables[key] = t
    return t


def _planes(src):
    L.require_cuda(src)
    if src.dtype != torch.uint8 or src.dim() != 3 or src.numel() == 0:
        raise TypeError(f"expected a non-empty uint8 [planes,H,W] CUDA tensor, got {src.dtype} {tuple(src.shape)}")
    return src.contiguous()


def resize(src, size, interpolation=AREA, binary_f32=False):
    """cv2.resize(plane, (dw, dh), interpolation=...) for every plane of a uint8 [planes,H,W] CUDA
    tensor.  `binary_f32=True` (nearest only) returns (plane > 0) as float32 -- the mask path."""
    s = _planes(src)
    n, sh, sw = s.shape
    dw, dh = int(size[0]), int(size[1])
    dev, st, lib = s.device, L.stream_ptr(), L.lib()
    if interpolation == NEAREST:
        xo, yo = _table("nearest", sh, sw, dh, dw, dev)
        out = torch.empty((n, dh, dw), dtype=torch.float32 if binary_f32 else torch.uint8, device=dev)
        L.check(lib.b2seg_resize_nearest_u8(s.data_ptr(), n, sh, sw, None if binary_f32 else out.data_ptr(),
                                            out.data_ptr() if binary_f32 else None, dh, dw,
                                            xo.data_ptr(), yo.data_ptr(), st))
        return out
    if binary_f32:
        raise ValueError("binary_f32 is the nearest-neighbour mask path")
    out = torch.empty((n, dh, dw), dtype=torch.uint8, device=dev)
    if interpolation == AREA and sw >= dw and sh >= dh:
        if sw % dw == 0 and sh % dh == 0:
            L.check(lib.b2seg_resize_areafast_u8(s.data_ptr(), n, sh, sw, out.data_ptr(), dh, dw, st))
        else:
            t = _table("area", sh, sw, dh, dw, dev)
            L.check(lib.b2seg_resize_area_u8(s.data_ptr(), n, sh, sw, out.data_ptr(), dh, dw,
                                             *(a.data_ptr() for a in t), st))
        return out
    if interpolation not in (LINEAR, AREA):
        raise NotImplementedError(f"interpolation {interpolation}: INTER_NEAREST, INTER_LINEAR and INTER_AREA are implemented")
    xo, xa, yo, ya = _table("linear_area" if interpolation == AREA else "linear", sh, sw, dh, dw, dev)
    L.check(lib.b2seg_resize_linear_u8(s.data_ptr(), n, sh, sw, out.data_ptr(), dh, dw, xo.data_ptr(),
                                       xa.data_ptr(), yo.data_ptr(), ya.data_ptr(), st))
    return out


def histograms(planes):
    """uint32 [N,256] histograms of a uint8 [N,H,W] CUDA tensor."""
    s = _planes(planes)
    n = s.shape[0]
    hist = torch.empty((n, 256), dtype=torch.int32, device=s.device)
    L.check(L.lib().b2seg_aug_hist_u8(s.data_ptr(), s.shape[1] * s.shape[2], n, hist.data_ptr(), L.stream_ptr()))
    return hist


def normalize_microscopy(planes, q=(2.0, 98.0), clip_limit=2.0, tiles=(8, 8), return_stages=False):
    """normalize_microscopy_image (dataset.py:30-43) for every plane of a uint8 [N,H,W] CUDA tensor:
    percentile clip, CLAHE, min-max -> float32 [N,H,W]."""
    s = _planes(planes)
    n, h, w = s.shape
    dev, st, lib = s.device, L.stream_ptr(), L.lib()
    hist = histograms(s)
    lut = torch.empty((n, 256), dtype=torch.uint8, device=dev)
    pct = torch.empty((n, 2), dtype=torch.float64, device=dev)
    L.check(lib.b2seg_pre_percentile_lut(hist.data_ptr(), n, h * w, float(q[0]), float(q[1]), lut.data_ptr(),
                                         pct.data_ptr(), st))
    nbytes = lib.b2seg_pre_clahe_workspace_bytes(n, tiles[0], tiles[1])
    ws = torch.empty((nbytes + 7) // 8, dtype=torch.int64, device=dev)
    clipped, eq = torch.empty_like(s), torch.empty_like(s)
    out = torch.empty((n, h, w), dtype=torch.float32, device=dev)
    L.check(lib.b2seg_pre_clahe_normalize(s.data_ptr(), n, h, w, lut.data_ptr(), float(clip_limit), tiles[0],
                                          tiles[1], clipped.data_ptr(), eq.data_ptr(), out.data_ptr(),
                                          ws.data_ptr(), nbytes, st))
    if return_stages:
        return out, {"percentiles": pct, "clipped": clipped, "clahe": eq}
    return out


def scale255(planes):
    s = _planes(planes)
    out = torch.empty(s.shape, dtype=torch.float32, device=s.device)
    L.check(L.lib().b2seg_pre_scale255_f32(s.data_ptr(), s.shape[0], s.shape[1] * s.shape[2], out.data_ptr(),
                                           L.stream_ptr()))
    return out


def dataset_items(images_u8, masks_u8, img_size=(256, 256), normalize=True):
    """CellSegmentationDataset.__getitem__ for a whole stack: uint8 [N,H,W] image / mask planes ->
    (images [N,1,h,w] fp32 in [0,1], masks [N,1,h,w] fp32 {0,1}); 8 launches per stack."""
    im = resize(images_u8, img_size, AREA)
    im = normalize_microscopy(im) if normalize else scale255(im)
    mk = resize(masks_u8, img_size, NEAREST, binary_f32=True)
    return im.unsqueeze(1), mk.unsqueeze(1)
