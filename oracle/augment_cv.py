"""numpy restatement of the OpenCV arithmetic the reference's geometric augmentations
bottom out in (oracle; tests only).

The reference calls albumentations (dataset.py:148-154,184;
forward_selection_integration.py:73-105), an un-vendored dependency
(requirements.txt:7, `albumentations>=1.1.0`) which in turn calls OpenCV
(`opencv-python>=4.5.5`): cv2.flip / np.rot90 / transpose / cv2.warpAffine /
cv2.remap.  What is restated here is OpenCV's published algorithm
(modules/imgproc/src/imgwarp.cpp): warpAffine inverts the matrix in double,
evaluates destination->source coordinates in 10-bit fixed point and reduces them
to 1/32 px; remap rounds float maps to 1/32 px; bilinear taps are blended with
15-bit integer weights (BilinearTab_i) and rounded with +2^14 >> 15.  Pinned
against cv2 itself in tests/test_oracle_cpu.py.
"""
import numpy as np

NEAREST, LINEAR = 0, 1
CONSTANT, REFLECT, REFLECT101 = 0, 1, 2


def hflip(a):
    return a[..., ::-1].copy()


def vflip(a):
    return a[..., ::-1, :].copy()


def rot90(a, k):
    return np.ascontiguousarray(np.rot90(a, k, axes=(-2, -1)))


def transpose(a):
    return np.ascontiguousarray(np.swapaxes(a, -1, -2))


def invert_affine(m):
    """cv2.warpAffine's own inversion of the forward 2x3 matrix (double)."""
    m = np.asarray(m, dtype=np.float64).reshape(2, 3).copy()
    d = m[0, 0] * m[1, 1] - m[0, 1] * m[1, 0]
    d = 1.0 / d if d != 0 else 0.0
    a11, a22 = m[1, 1] * d, m[0, 0] * d
    m[0, 0] = a11
    m[0, 1] *= -d
    m[1, 0] *= -d
    m[1, 1] = a22
    b1 = -m[0, 0] * m[0, 2] - m[0, 1] * m[1, 2]
    b2 = -m[1, 0] * m[0, 2] - m[1, 1] * m[1, 2]
    m[0, 2], m[1, 2] = b1, b2
    return m


def _border(p, n, mode):
    """cv::borderInterpolate; returns -1 for constant-border misses."""
    p = np.asarray(p, dtype=np.int64).copy()
    inside = (p >= 0) & (p < n)
    if mode == CONSTANT:
        return np.where(inside, p, -1)
    if n == 1:
        return np.zeros_like(p)
    delta = 1 if mode == REFLECT101 else 0
    for _ in range(64):
        bad = (p < 0) | (p >= n)
        if not bad.any():
            break
        p = np.where(p < 0, -p - 1 + delta, p)
        p = np.where(p >= n, n - 1 - (p - n) - delta, p)
    return p


def _fetch(src, x, y, border, cval):
    h, w = src.shape
    xx, yy = _border(x, w, border), _border(y, h, border)
    miss = (xx < 0) | (yy < 0)
    v = src[np.clip(yy, 0, h - 1), np.clip(xx, 0, w - 1)].astype(np.int64)
    return np.where(miss, cval, v)


def _blend(src, sx, sy, fx, fy, interp, border, cval):
    if interp == NEAREST:
        return _fetch(src, sx, sy, border, cval).astype(np.uint8)
    w00 = (32 - fy) * (32 - fx) * 32
    w01 = (32 - fy) * fx * 32
    w10 = fy * (32 - fx) * 32
    w11 = fy * fx * 32
    acc = (_fetch(src, sx, sy, border, cval) * w00 + _fetch(src, sx + 1, sy, border, cval) * w01
           + _fetch(src, sx, sy + 1, border, cval) * w10
           + _fetch(src, sx + 1, sy + 1, border, cval) * w11 + (1 << 14)) >> 15
    return np.clip(acc, 0, 255).astype(np.uint8)


def warp_affine_u8(src, m_forward, interp=LINEAR, border=CONSTANT, cval=0):
    """cv2.warpAffine(src, M, (w, h), flags=interp, borderMode=border, borderValue=cval) for one
    uint8 plane (WarpAffineInvoker: AB_BITS = 10, INTER_BITS = 5)."""
    h, w = src.shape
    m = invert_affine(m_forward)
    xs = np.arange(w, dtype=np.float64)
    ys = np.arange(h, dtype=np.float64)
    rd = 512 if interp == NEAREST else 16
    adelta = np.rint(m[0, 0] * xs * 1024.0).astype(np.int64)
    bdelta = np.rint(m[1, 0] * xs * 1024.0).astype(np.int64)
    x0 = np.rint((m[0, 1] * ys + m[0, 2]) * 1024.0).astype(np.int64) + rd
    y0 = np.rint((m[1, 1] * ys + m[1, 2]) * 1024.0).astype(np.int64) + rd
    X = x0[:, None] + adelta[None, :]
    Y = y0[:, None] + bdelta[None, :]
    if interp == NEAREST:
        sx, sy, fx, fy = X >> 10, Y >> 10, 0, 0
    else:
        X, Y = X >> 5, Y >> 5
        sx, sy, fx, fy = X >> 5, Y >> 5, X & 31, Y & 31
    sx, sy = np.clip(sx, -32768, 32767), np.clip(sy, -32768, 32767)
    return _blend(src, sx, sy, fx, fy, interp, border, cval)


def remap_u8(src, map_x, map_y, interp=LINEAR, border=REFLECT, cval=0):
    """cv2.remap(src, map_x, map_y, interp, borderMode=border) for float32 absolute maps."""
    mx = np.asarray(map_x, dtype=np.float32)
    my = np.asarray(map_y, dtype=np.float32)
    if interp == NEAREST:
        sx, sy, fx, fy = np.rint(mx).astype(np.int64), np.rint(my).astype(np.int64), 0, 0
    else:
        X = np.rint(mx * np.float32(32.0)).astype(np.int64)
        Y = np.rint(my * np.float32(32.0)).astype(np.int64)
        sx, sy, fx, fy = X >> 5, Y >> 5, X & 31, Y & 31
    sx, sy = np.clip(sx, -32768, 32767), np.clip(sy, -32768, 32767)
    return _blend(src, sx, sy, fx, fy, interp, border, cval)


# ---- parameter -> matrix / map builders (host-side; the sampling itself is unpinned) ----------
def affine_matrix(h, w, scale=1.0, translate=(0.0, 0.0), rotate_deg=0.0, shear_deg=(0.0, 0.0)):
    """Forward 2x3 matrix about the image centre (w/2 - 0.5, h/2 - 0.5): scale, shear, rotate,
    then translate by a fraction of the image size (the parameterisation of A.Affine in
    dataset.py:150-151)."""
    cx, cy = w / 2.0 - 0.5, h / 2.0 - 0.5
    th = np.deg2rad(rotate_deg)
    shx, shy = np.tan(np.deg2rad(shear_deg[0])), np.tan(np.deg2rad(shear_deg[1]))
    rot = np.array([[np.cos(th), -np.sin(th)], [np.sin(th), np.cos(th)]])
    shear = np.array([[1.0, shx], [shy, 1.0]])
    lin = rot @ shear * scale
    t = np.array([cx + translate[0] * w, cy + translate[1] * h]) - lin @ np.array([cx, cy])
    return np.concatenate([lin, t[:, None]], axis=1)


def grid_distortion_maps(h, w, xsteps, ysteps):
    """Piecewise-linear remap maps of A.GridDistortion(num_steps=len(xsteps)): cell i of the
    regular grid is stretched by xsteps[i] / ysteps[i] (forward_selection_integration.py:89-95)."""
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
    xx, yy = axis(w, xsteps), axis(h, ysteps)
    return np.tile(xx[None, :], (h, 1)).astype(np.float32), np.tile(yy[:, None], (1, w)).astype(np.float32)


def optical_distortion_maps(h, w, k, dx, dy):
    """Radial (camera-model) remap maps of A.OpticalDistortion with coefficient k and principal
    point shift (dx, dy) (forward_selection_integration.py:99-105)."""
    fx = fy = float(w)
    cx, cy = w * 0.5 + dx, h * 0.5 + dy
    xs, ys = np.meshgrid(np.arange(w, dtype=np.float64), np.arange(h, dtype=np.float64))
    x, y = (xs - cx) / fx, (ys - cy) / fy
    r2 = x * x + y * y
    f = 1 + k * r2 + k * r2 * r2
    return (x * f * fx + cx).astype(np.float32), (y * f * fy + cy).astype(np.float32)


def elastic_maps(h, w, alpha, sigma, rng):
    """Elastic displacement maps (Gaussian-smoothed uniform noise x alpha).  The north star asks
    for an elastic warp but the reference has no call site for one: parity unpinned."""
    from scipy.ndimage import gaussian_filter
    dx = gaussian_filter(rng.uniform(-1, 1, (h, w)), sigma) * alpha
    dy = gaussian_filter(rng.uniform(-1, 1, (h, w)), sigma) * alpha
    xs, ys = np.meshgrid(np.arange(w), np.arange(h))
    return (xs + dx).astype(np.float32), (ys + dy).astype(np.float32)
