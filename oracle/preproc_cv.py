"""CPU restatement (numpy) of the OpenCV / numpy arithmetic behind CellSegmentationDataset.__getitem__
(reference dataset.py:30-66) -- oracle; test infrastructure only.

  image: cv2.resize(INTER_AREA) -> np.percentile(2, 98) clip -> uint8 -> CLAHE(2.0, 8x8) -> min-max
  mask:  cv2.resize(INTER_NEAREST) -> (> 0)

The algorithms live in OpenCV (opencv-python >= 4.5.5, requirements.txt:6; modules/imgproc/src/
resize.cpp, clahe.cpp) and numpy (np.percentile, method 'linear'), neither vendored by the
reference.  Pinning: cv2 and numpy are installed on both machines, so tests/test_oracle_cpu.py
checks every function here against cv2.resize / cv2.createCLAHE / np.percentile themselves.
"""
import math

import numpy as np

F = np.float32


def _cv_round(x):
    """cvRound / saturate_cast<int>(float): round half to even."""
    return np.rint(x).astype(np.int64)


# ---------------------------------------------------------------- resize ----
def resize_nearest_u8(src, dh, dw):
    h, w = src.shape
    ifx, ify = 1.0 / (dw / w), 1.0 / (dh / h)
    xs = np.minimum(np.floor(np.arange(dw) * ifx).astype(np.int64), w - 1)
    ys = np.minimum(np.floor(np.arange(dh) * ify).astype(np.int64), h - 1)
    return src[ys[:, None], xs[None, :]]


def _linear_tab(ssize, dsize, area_mode):
    """(offsets, int16 coefficients [d][2]) of resize.cpp's INTER_LINEAR (or INTER_AREA upscaling)."""
    scale = ssize / dsize
    inv_scale = dsize / ssize
    ofs = np.zeros(dsize, np.int64)
    coef = np.zeros((dsize, 2), np.int64)
    for d in range(dsize):
        if not area_mode:
            fx = F((d + 0.5) * scale - 0.5)
            sx = int(math.floor(fx))
            fx = F(fx - F(sx))
        else:
            sx = int(math.floor(d * scale))
            fx = F((d + 1) - (sx + 1) * inv_scale)
            fx = F(0.0) if fx <= 0 else F(fx - F(math.floor(fx)))
        if sx < 0:
            fx, sx = F(0.0), 0
        if sx >= ssize - 1:
            fx, sx = F(0.0), ssize - 1
        ofs[d] = sx
        c0, c1 = F(1.0) - fx, fx
        coef[d, 0] = int(np.clip(np.rint(F(c0) * F(2048.0)), -32768, 32767))
        coef[d, 1] = int(np.clip(np.rint(F(c1) * F(2048.0)), -32768, 32767))
    return ofs, coef


def _resize_linear_core(src, dh, dw, area_mode):
    h, w = src.shape
    xo, xa = _linear_tab(w, dw, area_mode)
    yo, ya = _linear_tab(h, dh, area_mode)
    s = src.astype(np.int64)
    x1 = np.minimum(xo + 1, w - 1)
    hor = s[:, xo] * xa[None, :, 0] + s[:, x1] * xa[None, :, 1]          # [h][dw], x 2048
    y1 = np.minimum(yo + 1, h - 1)
    r0, r1 = hor[yo], hor[y1]
    out = (((ya[:, None, 0] * (r0 >> 4)) >> 16) + ((ya[:, None, 1] * (r1 >> 4)) >> 16) + 2) >> 2
    return np.clip(out, 0, 255).astype(np.uint8)


def resize_linear_u8(src, dh, dw):
    return _resize_linear_core(src, dh, dw, False)


def _area_tab(ssize, dsize, scale):
    tab = []
    for dx in range(dsize):
        fsx1 = dx * scale
        fsx2 = fsx1 + scale
        cell = min(scale, ssize - fsx1)
        sx1, sx2 = math.ceil(fsx1), math.floor(fsx2)
        sx2 = min(sx2, ssize - 1)
        sx1 = min(sx1, sx2)
        if sx1 - fsx1 > 1e-3:
            tab.append((dx, sx1 - 1, F((sx1 - fsx1) / cell)))
        for sx in range(sx1, sx2):
            tab.append((dx, sx, F(1.0 / cell)))
        if fsx2 - sx2 > 1e-3:
            tab.append((dx, sx2, F(min(min(fsx2 - sx2, 1.0), cell) / cell)))
    return tab


def resize_area_u8(src, dh, dw):
    """cv2.resize(src, (dw, dh), interpolation=cv2.INTER_AREA) on a uint8 plane."""
    h, w = src.shape
    sx, sy = w / dw, h / dh
    if sx >= 1 and sy >= 1:
        ix, iy = int(sx), int(sy)   # saturate_cast<int>(double) rounds; equal to truncation when integral
        if abs(sx - round(sx)) < np.finfo(np.float64).eps and abs(sy - round(sy)) < np.finfo(np.float64).eps:
            ix, iy = int(round(sx)), int(round(sy))
            blocks = src[:dh * iy, :dw * ix].reshape(dh, iy, dw, ix).astype(np.int64).sum(axis=(1, 3))
            if ix == 2 and iy == 2:
                return ((blocks + 2) >> 2).astype(np.uint8)
            scale = F(1.0) / F(ix * iy)
            return np.clip(_cv_round(blocks.astype(F) * scale), 0, 255).astype(np.uint8)
        xtab, ytab = _area_tab(w, dw, sx), _area_tab(h, dh, sy)
        out = np.zeros((dh, dw), np.uint8)
        s = src.astype(F)
        summ = np.zeros(dw, F)
        prev = ytab[0][0]
        for (dy, syi, beta) in ytab:
            buf = np.zeros(dw, F)
            for (dx, sxi, alpha) in xtab:
                buf[dx] = F(buf[dx] + F(s[syi, sxi] * alpha))
            if dy != prev:
                out[prev] = np.clip(_cv_round(summ), 0, 255)
                summ = (beta * buf).astype(F)
                prev = dy
            else:
                summ = (summ + (beta * buf).astype(F)).astype(F)
        out[prev] = np.clip(_cv_round(summ), 0, 255)
        return out
    return _resize_linear_core(src, dh, dw, True)


# ----------------------------------------------------------------- CLAHE ----
def _reflect101(p, n):
    if n == 1:
        return 0
    while p < 0 or p >= n:
        p = -p if p < 0 else 2 * (n - 1) - p
    return p


def clahe_u8(src, clip_limit=2.0, tiles=(8, 8)):
    """cv2.createCLAHE(clipLimit, tileGridSize=tiles).apply(src) on a uint8 plane (clahe.cpp)."""
    tx, ty = tiles
    h, w = src.shape
    if w % tx == 0 and h % ty == 0:
        ext = src
    else:
        eh, ew = h + (ty - h % ty), w + (tx - w % tx)
        ys = [_reflect101(y, h) for y in range(eh)]
        xs = [_reflect101(x, w) for x in range(ew)]
        ext = src[np.array(ys)[:, None], np.array(xs)[None, :]]
    th, tw = ext.shape[0] // ty, ext.shape[1] // tx
    area = th * tw
    lut_scale = F(255.0) / F(area)
    clip = 0
    if clip_limit > 0.0:
        clip = max(int(clip_limit * area / 256), 1)
    luts = np.zeros((ty, tx, 256), np.uint8)
    for j in range(ty):
        for i in range(tx):
            hist = np.bincount(ext[j * th:(j + 1) * th, i * tw:(i + 1) * tw].ravel(), minlength=256).astype(np.int64)
            if clip > 0:
                clipped = int(np.maximum(hist - clip, 0).sum())
                hist = np.minimum(hist, clip)
                batch = clipped // 256
                residual = clipped - batch * 256
                hist = hist + batch
                if residual != 0:
                    step = max(256 // residual, 1)
                    k = 0
                    while k < 256 and residual > 0:
                        hist[k] += 1
                        k += step
                        residual -= 1
            luts[j, i] = np.clip(_cv_round(np.cumsum(hist).astype(F) * lut_scale), 0, 255)
    inv_tw, inv_th = F(1.0) / F(tw), F(1.0) / F(th)
    xs = np.arange(w, dtype=F) * inv_tw - F(0.5)
    tx1 = np.floor(xs).astype(np.int64)
    xa = (xs - tx1.astype(F)).astype(F)
    xa1 = (F(1.0) - xa).astype(F)
    tx2 = np.minimum(tx1 + 1, tx - 1)
    tx1 = np.maximum(tx1, 0)
    ysf = np.arange(h, dtype=F) * inv_th - F(0.5)
    ty1 = np.floor(ysf).astype(np.int64)
    ya = (ysf - ty1.astype(F)).astype(F)
    ya1 = (F(1.0) - ya).astype(F)
    ty2 = np.minimum(ty1 + 1, ty - 1)
    ty1 = np.maximum(ty1, 0)
    v = src.astype(np.int64)
    l11 = luts[ty1[:, None], tx1[None, :], v].astype(F)
    l12 = luts[ty1[:, None], tx2[None, :], v].astype(F)
    l21 = luts[ty2[:, None], tx1[None, :], v].astype(F)
    l22 = luts[ty2[:, None], tx2[None, :], v].astype(F)
    top = (l11 * xa1[None, :]).astype(F) + (l12 * xa[None, :]).astype(F)
    bot = (l21 * xa1[None, :]).astype(F) + (l22 * xa[None, :]).astype(F)
    res = (top.astype(F) * ya1[:, None]).astype(F) + (bot.astype(F) * ya[:, None]).astype(F)
    return np.clip(_cv_round(res.astype(F)), 0, 255).astype(np.uint8)


# ------------------------------------------------------ the whole pipeline ----
def percentile_clip_u8(image, lo=2.0, hi=98.0):
    """np.clip(image, *np.percentile(image, [lo, hi])).astype(np.uint8)  (dataset.py:33-34,37)."""
    a = np.sort(image.ravel())
    n = a.size
    out = []
    for q in (lo, hi):
        vi = q / 100.0 * (n - 1)
        k = int(math.floor(vi))
        t = vi - k
        x0, x1 = float(a[k]), float(a[min(k + 1, n - 1)])
        d = x1 - x0
        p = x0 + d * t
        if t >= 0.5:
            p = x1 - d * (1 - t)
        out.append(p)
    return np.clip(image, out[0], out[1]).astype(np.uint8), out


def normalize_microscopy_image(image):
    """dataset.py:30-43 on an already resized uint8 plane -> float32 in [0, 1]."""
    clipped, _ = percentile_clip_u8(image)
    eq = clahe_u8(clipped, 2.0, (8, 8))
    mn, mx = eq.min(), eq.max()
    return ((eq - mn) / (np.uint8(mx - mn) + 1e-8)).astype(np.float32)


def dataset_item(image, mask, img_size=(256, 256), normalize=True):
    """CellSegmentationDataset.__getitem__ (dataset.py:44-66) from decoded uint8 planes."""
    dw, dh = img_size
    im = resize_area_u8(image, dh, dw)
    mk = resize_nearest_u8(mask, dh, dw)
    im = normalize_microscopy_image(im) if normalize else im.astype(np.float32) / np.float32(255.0)
    return im.astype(np.float32), (mk > 0).astype(np.float32)
