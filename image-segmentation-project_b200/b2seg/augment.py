"""Geometric augmentation on the GPU (host side of csrc/augment.cu).

All functions take a uint8 CUDA tensor of shape [planes, H, W] (image and mask planes
stacked) and return a new tensor.  Index permutations are bit-exact; warps reproduce
cv2.warpAffine / cv2.remap fixed-point arithmetic (see csrc/augment.cu).
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

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
    """Forward 2x3 matrix about the image centre (w/2-0.5, h/2-0.5): scale (a number, or (sx, sy)
    per axis as A.Affine samples it with keep_ratio=False), shear, rotation, then a translation
    given as a fraction of the image size (A.Affine parameterisation used at dataset.py:150-151 /
    forward_selection_integration.py:75-83)."""
    cx, cy = w / 2.0 - 0.5, h / 2.0 - 0.5
    th = math.radians(rotate_deg)
    shx, shy = math.tan(math.radians(shear_deg[0])), math.tan(math.radians(shear_deg[1]))
    c, s = math.cos(th), math.sin(th)
    sx, sy = (scale, scale) if not isinstance(scale, (tuple, list)) else (float(scale[0]), float(scale[1]))
    # rot @ shear @ diag(sx, sy)
    l00 = (c * 1.0 + -s * shy) * sx
    l01 = (c * shx + -s * 1.0) * sy
    l10 = (s * 1.0 + c * shy) * sx
    l11 = (s * shx + c * 1.0) * sy
    tx = cx + translate[0] * w - (l00 * cx + l01 * cy)
    ty = cy + translate[1] * h - (l10 * cx + l11 * cy)
    return [[l00, l01, tx], [l10, l11, ty]]


# ------------------------------------------------------------ batched warps ----
def _item_table(entries, device):
    """[(minv[6] | None, interp per plane, cval per plane, border)] -> device table of b2seg_warp_item."""
    tab = (L.WarpItem * len(entries))()
    for k, (minv, interp, cval, border) in enumerate(entries):
        for j in range(6):
            tab[k].minv[j] = float(minv[j]) if minv is not None else 0.0
        for j in range(4):
            tab[k].interp[j] = int(interp[j]) if j < len(interp) else int(interp[-1])
            tab[k].cval[j] = int(cval[j]) if j < len(cval) else int(cval[-1])
        tab[k].border = int(border)
    return torch.frombuffer(bytearray(bytes(tab)), dtype=torch.uint8).to(device)


def warp_affine_batch(src, matrices, interp=(LINEAR, NEAREST), border=CONSTANT, cval=(0, 0)):
    """src uint8 [items, planes_per_item, H, W]; item i is warped by the forward 2x3 matrix
    matrices[i]; plane j of every item uses interp[j] / cval[j] (image bilinear, mask nearest).
    ONE launch for the whole batch (b2seg_aug_warp_affine_batch_u8)."""
    L.require_cuda(src)
    if src.dtype != torch.uint8 or src.dim() != 4:
        raise TypeError(f"expected uint8 [items,planes,H,W], got {src.dtype} {tuple(src.shape)}")
    s = src.contiguous()
    n, ppi, h, w = s.shape
    if len(matrices) != n:
        raise ValueError(f"{len(matrices)} matrices for {n} items")
    tab = _item_table([(invert_affine(m), interp, cval, border) for m in matrices], s.device)
    dst = torch.empty_like(s)
    L.check(L.lib().b2seg_aug_warp_affine_batch_u8(s.data_ptr(), dst.data_ptr(), n, ppi, h, w, tab.data_ptr(),
                                                   L.stream_ptr()))
    return dst


def remap_batch(src, map_x, map_y, interp=(LINEAR, NEAREST), border=REFLECT, cval=(0, 0)):
    """src uint8 [items, planes_per_item, H, W]; maps float32 [H,W] (shared) or [items,H,W]."""
    L.require_cuda(src, map_x, map_y)
    if src.dtype != torch.uint8 or src.dim() != 4:
        raise TypeError(f"expected uint8 [items,planes,H,W], got {src.dtype} {tuple(src.shape)}")
    s = src.contiguous()
    n, ppi, h, w = s.shape
    mx, my = map_x.to(torch.float32).contiguous(), map_y.to(torch.float32).contiguous()
    shared = mx.dim() == 2
    if tuple(mx.shape[-2:]) != (h, w) or mx.shape != my.shape or (not shared and mx.shape[0] != n):
        raise ValueError(f"maps {tuple(mx.shape)} / {tuple(my.shape)} do not fit planes {tuple(s.shape)}")
    tab = _item_table([(None, interp, cval, border)] * n, s.device)
    dst = torch.empty_like(s)
    L.check(L.lib().b2seg_aug_remap_batch_u8(s.data_ptr(), dst.data_ptr(), n, ppi, h, w, mx.data_ptr(), my.data_ptr(),
                                             0 if shared else h * w, tab.data_ptr(), L.stream_ptr()))
    return dst


def warp_affine_pair(planes, m_forward, interp_image=LINEAR, interp_mask=NEAREST, border=CONSTANT, cval=0,
                     cval_mask=0):
    """[planes,H,W] with plane 0 = image, planes 1.. = masks: one launch, per-plane interpolation."""
    s = _planes(planes)
    return warp_affine_batch(s.unsqueeze(0), [m_forward], (interp_image,) + (interp_mask,) * 3, border,
                             (cval,) + (cval_mask,) * 3)[0]


def remap_pair(planes, map_x, map_y, interp_image=LINEAR, interp_mask=NEAREST, border=REFLECT, cval=0, cval_mask=0):
    s = _planes(planes)
    return remap_batch(s.unsqueeze(0), map_x, map_y, (interp_image,) + (interp_mask,) * 3, border,
                       (cval,) + (cval_mask,) * 3)[0]


# --------------------------------------------------- coordinate maps on the device ----
def grid_distortion_maps(h, w, xsteps, ysteps, device):
    """Remap maps of A.GridDistortion(num_steps=len(steps)-1): cell i of the regular grid is
    stretched by steps[i] (piecewise linear); the per-axis profiles (w + h floats) are laid out on
    the host, the [H,W] maps are expanded on the device."""
    def axis(n, steps):
        num = len(steps)
        step = n // num
        out = np.zeros(n, np.float32)
        prev = 0.0
        for idx, x in enumerate(range(0, n, step)):
            start, end = x, min(x + step, n)
            cur = prev + step * steps[min(idx, num - 1)] if end < n else float(n)
            out[start:end] = np.linspace(prev, cur, end - start)
            prev = cur
        return out
    xx = torch.from_numpy(axis(w, list(xsteps))).to(device)
    yy = torch.from_numpy(axis(h, list(ysteps))).to(device)
    return xx[None, :].expand(h, w).contiguous(), yy[:, None].expand(h, w).contiguous()


def optical_distortion_maps(h, w, k, dx, dy, device):
    """Radial camera-model maps of A.OpticalDistortion: fx = fy = w, principal point (w/2+dx, h/2+dy),
    x' = x (1 + k r^2 + k r^4).  float64 on the device, rounded once to float32."""
    fx = fy = float(w)
    cx, cy = w * 0.5 + dx, h * 0.5 + dy
    xs = torch.arange(w, dtype=torch.float64, device=device)[None, :].expand(h, w)
    ys = torch.arange(h, dtype=torch.float64, device=device)[:, None].expand(h, w)
    x, y = (xs - cx) / fx, (ys - cy) / fy
    r2 = x * x + y * y
    f = 1 + k * r2 + k * r2 * r2
    return (x * f * fx + cx).to(torch.float32).contiguous(), (y * f * fy + cy).to(torch.float32).contiguous()


def gaussian_filter_f32(field, sigma, gain=1.0, add_coord=(0, 0)):
    """scipy.ndimage.gaussian_filter(field, sigma) (mode='reflect', truncate=4) of float32 [planes,H,W]
    on the device: two 1-D passes (b2seg_aug_blur1d_f32).  The second pass applies `gain` and can
    add the x / y pixel coordinate per plane (add_coord[plane] in {0, 1, 2})."""
    L.require_cuda(field)
    f = field.to(torch.float32).contiguous()
    n, h, w = f.shape
    rad = int(4.0 * float(sigma) + 0.5)
    ax = np.arange(-rad, rad + 1, dtype=np.float64)
    wts = np.exp(-0.5 * ax * ax / (float(sigma) ** 2))
    wts = torch.from_numpy((wts / wts.sum()).astype(np.float32)).to(f.device)
    tmp, out = torch.empty_like(f), torch.empty_like(f)
    lib, st = L.lib(), L.stream_ptr()
    L.check(lib.b2seg_aug_blur1d_f32(f.data_ptr(), tmp.data_ptr(), n, h, w, wts.data_ptr(), rad, 1, 1.0, 0, st))
    for pl in range(n):  # per-plane coordinate term
        L.check(lib.b2seg_aug_blur1d_f32(tmp[pl].data_ptr(), out[pl].data_ptr(), 1, h, w, wts.data_ptr(), rad, 0,
                                         float(gain), int(add_coord[pl] if pl < len(add_coord) else 0), st))
    return out


def uniform_field(shape, seed, device):
    """U(-1, 1) float32 field from the in-kernel Philox generator."""
    out = torch.empty(shape, dtype=torch.float32, device=device)
    L.check(L.lib().b2seg_aug_uniform_f32(out.data_ptr(), out.numel(), int(seed) & (2 ** 64 - 1), L.stream_ptr()))
    return out


def elastic_maps(h, w, alpha, sigma, seed, device, same_dxdy=False):
    """ElasticTransform maps: map_x = x + alpha * G_sigma(U(-1,1)), map_y = y + alpha * G_sigma(U(-1,1))."""
    noise = uniform_field((1 if same_dxdy else 2, h, w), seed, device)
    if same_dxdy:
        noise = noise.expand(2, h, w).contiguous()
    maps = gaussian_filter_f32(noise, sigma, gain=alpha, add_coord=(1, 2))
    return maps[0], maps[1]


# ------------------------------------------------------ non-geometric transforms ----
def _image_only(planes, fn):
    """Apply fn to plane 0 (the image); masks pass through."""
    s = _planes(planes)
    out = s.clone()
    fn(s[0], out[0])
    return out


def lut_image(planes, table):
    lut = torch.from_numpy(np.ascontiguousarray(table, dtype=np.uint8)).to(planes.device)
    if lut.numel() != 256:
        raise ValueError("look-up table must have 256 entries")
    return _image_only(planes, lambda a, b: L.check(L.lib().b2seg_aug_lut_u8(
        a.data_ptr(), b.data_ptr(), a.numel(), lut.data_ptr(), L.stream_ptr())))


def brightness_contrast(planes, brightness, contrast, contrast_first=False):
    """ColorJitter on a grayscale uint8 tile: v*b, and v*c + mean*(1-c), as uint8 look-up tables
    (truncating cast), in the sampled order; the mean is the exact integer mean of the CURRENT image."""
    from . import preproc
    s = _planes(planes)
    out = s.clone()
    lib, st = L.lib(), L.stream_ptr()

    def bright(img):
        L.check(lib.b2seg_aug_affine_lut_u8(img.data_ptr(), img.data_ptr(), img.numel(), None, float(brightness), 0.0, st))

    def contr(img):
        hist = preproc.histograms(img.unsqueeze(0))
        L.check(lib.b2seg_aug_affine_lut_u8(img.data_ptr(), img.data_ptr(), img.numel(), hist.data_ptr(),
                                            float(contrast), 1.0 - float(contrast), st))

    for step in ((contr, bright) if contrast_first else (bright, contr)):
        step(out[0])
    return out


def gauss_noise(planes, mean, sigma, seed):
    return _image_only(planes, lambda a, b: L.check(L.lib().b2seg_aug_gauss_noise_u8(
        a.data_ptr(), b.data_ptr(), a.numel(), float(mean), float(sigma), int(seed) & (2 ** 64 - 1), L.stream_ptr())))


def fill_rects(planes, rects, value=0, mask_value=None):
    """Dropout: fill (y1, x1, y2, x2) rectangles of the image with `value` (and of the masks with
    `mask_value` when given)."""
    s = _planes(planes)
    out = s.clone()
    if not len(rects):
        return out
    tab = torch.tensor([list(map(int, r)) for r in rects], dtype=torch.int32, device=s.device)
    h, w = s.shape[-2:]
    lib, st = L.lib(), L.stream_ptr()
    L.check(lib.b2seg_aug_fill_rects_u8(out[0].data_ptr(), h, w, tab.data_ptr(), tab.shape[0], int(value), st))
    if mask_value is not None:
        for pl in range(1, s.shape[0]):
            L.check(lib.b2seg_aug_fill_rects_u8(out[pl].data_ptr(), h, w, tab.data_ptr(), tab.shape[0], int(mask_value), st))
    return out


def filter2d_image(planes, kernel):
    k = np.ascontiguousarray(kernel, dtype=np.float32)
    if k.ndim != 2 or k.shape[0] != k.shape[1]:
        raise ValueError("square kernel expected")
    kd = torch.from_numpy(k).to(planes.device)
    s = _planes(planes)
    out = s.clone()
    h, w = s.shape[-2:]
    L.check(L.lib().b2seg_aug_filter2d_u8(s[0].data_ptr(), out[0].data_ptr(), 1, h, w, kd.data_ptr(), k.shape[0],
                                          L.stream_ptr()))
    return out


def crop(planes, y0, x0, height, width):
    s = _planes(planes)
    return s[:, y0:y0 + height, x0:x0 + width].contiguous()


def pad(planes, top, bottom, left, right, border=REFLECT101, cval=0):
    """cv2.copyMakeBorder for every plane."""
    s = _planes(planes)
    n, h, w = s.shape
    out = torch.empty((n, h + top + bottom, w + left + right), dtype=torch.uint8, device=s.device)
    L.check(L.lib().b2seg_aug_pad_u8(s.data_ptr(), out.data_ptr(), n, h, w, top, bottom, left, right, border,
                                     int(cval), L.stream_ptr()))
    return out


def crop_and_pad(planes, sides, border, cval, cval_mask, keep_size, interp):
    """A.CropAndPad: sides = (top, right, bottom, left), negative crops / positive pads; then resize
    back to the input size (image: `interp`, masks: nearest)."""
    from . import preproc
    s = _planes(planes)
    n, h, w = s.shape
    top, right, bottom, left = (int(v) for v in sides)
    ct, cb, cl, cr = max(-top, 0), max(-bottom, 0), max(-left, 0), max(-right, 0)
    if ct + cb >= h or cl + cr >= w:   # albumentations keeps at least one pixel
        ct = cb = cl = cr = 0
    out = s[:, ct:h - cb, cl:w - cr].contiguous()
    pt, pb, pl_, pr = max(top, 0), max(bottom, 0), max(left, 0), max(right, 0)
    if pt or pb or pl_ or pr:
        img = pad(out[0:1], pt, pb, pl_, pr, border, cval)
        msk = pad(out[1:], pt, pb, pl_, pr, border, cval_mask) if n > 1 else None
        out = img if msk is None else torch.cat([img, msk])
    if keep_size and tuple(out.shape[-2:]) != (h, w):
        img = preproc.resize(out[0:1], (w, h), preproc.LINEAR if interp == LINEAR else preproc.NEAREST)
        out = img if n == 1 else torch.cat([img, preproc.resize(out[1:], (w, h), preproc.NEAREST)])
    return out


def downscale_image(planes, scale, interp_down=NEAREST, interp_up=NEAREST):
    """A.Downscale: image resized by `scale` and back (cv2.resize arithmetic); masks untouched."""
    from . import preproc
    s = _planes(planes)
    n, h, w = s.shape
    # cv2.resize(img, None, fx=scale, fy=scale): dsize = round(w*scale), round(h*scale)
    dw, dh = max(1, int(round(w * scale))), max(1, int(round(h * scale)))
    code = lambda i: preproc.LINEAR if i == LINEAR else preproc.NEAREST
    small = preproc.resize(s[0:1], (dw, dh), code(interp_down))
    big = preproc.resize(small, (w, h), code(interp_up))
    return big if n == 1 else torch.cat([big, s[1:]])
