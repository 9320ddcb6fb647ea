"""Geometric augmentation on the GPU (host side of csrc/augment.cu).

All functions take a uint8 CUDA tensor of shape [planes, H, W] (image and mask planes
stacked) and return a new tensor.  Index permutations are bit-exact; warps reproduce
cv2.warpAffine / cv2.remap fixed-point arithmetic (see csrc/augment.cu).
"""
from __future__ import annotations

import ctypes as C
import math

import torch

from . import _lib as L

NEAREST, LINEAR = L.INTERP_NEAREST, L.INTERP_LINEAR
CONSTANT, REFLECT, REFLECT101 = L.BORDER_CONSTANT, L.BORDER_REFLECT, L.BORDER_REFLECT101


def _planes(src: torch.Tensor) -> torch.Tensor:
    L.require_cuda(src)
    if src.dtype != torch.uint8:
        raise TypeError(f"augment kernels take uint8 planes, got {src.dtype}")
    if src.dim() == 2:
        src = src.unsqueeze(0)
    if src.dim() != 3 or src.numel() == 0:
        raise ValueError(f"expected non-empty [planes,H,W], got {tuple(src.shape)}")
    return src.contiguous()


def _permute(src, mode, k=0):
    s = _planes(src)
    p, h, w = s.shape
    swap = mode == L.AUG_TRANSPOSE or (mode == L.AUG_ROT90 and k % 2 == 1)
    dst = torch.empty((p, w, h) if swap else (p, h, w), dtype=torch.uint8, device=s.device)
    L.check(L.lib().b2seg_aug_permute_u8(s.data_ptr(), dst.data_ptr(), p, h, w, mode, k, L.stream_ptr()))
    return dst


def hflip(src):
    return _permute(src, L.AUG_HFLIP)


def vflip(src):
    return _permute(src, L.AUG_VFLIP)


def rot90(src, k=1):
    """np.rot90(plane, k): counter-clockwise quarter turns."""
    return _permute(src, L.AUG_ROT90, int(k) % 4)


def transpose(src):
    return _permute(src, L.AUG_TRANSPOSE)


def invert_affine(m):
    """The double-precision inversion cv2.warpAffine applies to a forward 2x3 matrix."""
    m00, m01, m02, m10, m11, m12 = (float(v) for v in (m[0][0], m[0][1], m[0][2], m[1][0], m[1][1], m[1][2]))
    d = m00 * m11 - m01 * m10
    d = 1.0 / d if d != 0 else 0.0
    a11, a22 = m11 * d, m00 * d
    i00, i01, i10, i11 = a11, m01 * -d, m10 * -d, a22
    b1 = -i00 * m02 - i01 * m12
    b2 = -i10 * m02 - i11 * m12
    return [i00, i01, b1, i10, i11, b2]


def warp_affine(src, m_forward, interp=LINEAR, border=CONSTANT, cval=0):
    s = _planes(src)
    p, h, w = s.shape
    dst = torch.empty_like(s)
    minv = (C.c_double * 6)(*invert_affine(m_forward))
    L.check(L.lib().b2seg_aug_warp_affine_u8(s.data_ptr(), dst.data_ptr(), p, h, w, minv, interp,
                                             border, int(cval), L.stream_ptr()))
    return dst


def remap(src, map_x, map_y, interp=LINEAR, border=REFLECT, cval=0):
    s = _planes(src)
    p, h, w = s.shape
    L.require_cuda(map_x, map_y)
    mx = map_x.to(torch.float32).contiguous()
    my = map_y.to(torch.float32).contiguous()
    if tuple(mx.shape) != (h, w) or tuple(my.shape) != (h, w):
        raise ValueError(f"maps must be [{h},{w}], got {tuple(mx.shape)} / {tuple(my.shape)}")
    dst = torch.empty_like(s)
    L.check(L.lib().b2seg_aug_remap_u8(s.data_ptr(), dst.data_ptr(), p, h, w, mx.data_ptr(),
                                       my.data_ptr(), interp, border, int(cval), L.stream_ptr()))
    return dst


def affine_matrix(h, w, scale=1.0, translate=(0.0, 0.0), rotate_deg=0.0, shear_deg=(0.0, 0.0)):
    """Forward 2x3 matrix about the image centre (w/2-0.5, h/2-0.5): scale, shear, rotation,
    then a translation given as a fraction of the image size (A.Affine parameterisation used at
    dataset.py:150-151 / forward_selection_integration.py:75-83)."""
    cx, cy = w / 2.0 - 0.5, h / 2.0 - 0.5
    th = math.radians(rotate_deg)
    shx, shy = math.tan(math.radians(shear_deg[0])), math.tan(math.radians(shear_deg[1]))
    c, s = math.cos(th), math.sin(th)
    # rot @ shear * scale
    l00 = (c * 1.0 + -s * shy) * scale
    l01 = (c * shx + -s * 1.0) * scale
    l10 = (s * 1.0 + c * shy) * scale
    l11 = (s * shx + c * 1.0) * scale
    tx = cx + translate[0] * w - (l00 * cx + l01 * cy)
    ty = cy + translate[1] * h - (l10 * cx + l11 * cy)
    return [[l00, l01, tx], [l10, l11, ty]]
