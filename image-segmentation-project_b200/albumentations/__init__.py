"""`albumentations` stand-in for the reference's augmentation call sites, on the B200 gather kernels.

The reference imports `albumentations as A` (dataset.py:12, forward_selection_integration.py:8) --
an un-vendored dependency that is installed on neither machine -- and uses a small part of it:
`A.Compose`, the geometric transforms of dataset.py:148-154 and of the selection candidates
(forward_selection_integration.py:73-105), and eleven non-geometric candidates (:107-178).  This
package provides those names with the keyword arguments the reference passes, so the reference's
scripts import and run unchanged with `image-segmentation-project_b200/` first on PYTHONPATH.

Every transform works on uint8 planes ON THE GPU (b2seg.augment -> csrc/augment.cu): flips /
rot90 / transpose are exact index permutations, Affine is cv2.warpAffine's fixed-point warp,
GridDistortion / OpticalDistortion / ElasticTransform build their coordinate maps on the device and
go through cv2.remap's arithmetic.  `Compose(image=..., mask=...)` uploads the pair once, runs the
chain on the device and downloads once; `Compose.apply_planes` keeps everything on the device.

Parameter sampling uses Python's `random` module like albumentations 1.x (`random.seed(n)` makes a
run repeatable).  The sampling DISTRIBUTIONS follow the albumentations documentation; they are
host-side and "parity unpinned" (SURVEY 8c): the kernels are pinned against cv2 for any parameter.
"""
from __future__ import annotations

import math
import random

import numpy as np
import torch

from b2seg import augment as _ag

__version__ = "1.4.0+b2seg"

_CV_NEAREST, _CV_LINEAR = 0, 1          # cv2.INTER_NEAREST / INTER_LINEAR
_CV_CONSTANT, _CV_REFLECT, _CV_REFLECT101 = 0, 2, 4   # cv2.BORDER_CONSTANT / _REFLECT / _REFLECT_101


def _interp(code):
    if code in (None, _CV_LINEAR):
        return _ag.LINEAR
    if code == _CV_NEAREST:
        return _ag.NEAREST
    raise NotImplementedError(f"interpolation {code}: the B200 warp implements cv2.INTER_NEAREST and INTER_LINEAR")


def _border(code):
    try:
        return {_CV_CONSTANT: _ag.CONSTANT, _CV_REFLECT: _ag.REFLECT, _CV_REFLECT101: _ag.REFLECT101}[code]
    except KeyError:
        raise NotImplementedError(f"border_mode {code}: BORDER_CONSTANT, BORDER_REFLECT and BORDER_REFLECT_101 "
                                  "are implemented") from None


def _pair(v, low=None):
    """albumentations' to_tuple: scalar s -> (-s, s) (or (low, s)); 2-sequence kept."""
    if isinstance(v, (tuple, list)):
        return float(v[0]), float(v[1])
    return (-float(v), float(v)) if low is None else (float(low), float(v))


class BasicTransform:
    """Base: probability `p`, parameters sampled per call, applied to a [planes, H, W] uint8 CUDA
    tensor whose plane 0 is the image and plane 1 (when present) the mask."""

    geometric = True   # applies to the mask as well

    def __init__(self, always_apply=False, p=0.5):
        self.always_apply, self.p = bool(always_apply), float(p)

    def get_params(self, h, w):
        return {}

    def apply_planes(self, planes, **params):
        raise NotImplementedError

    def __call__(self, force_apply=False, **data):
        return Compose([self])(force_apply=force_apply, **data)

    def _fire(self, force_apply=False):
        return force_apply or self.always_apply or random.random() < self.p

    def __repr__(self):
        kw = ", ".join(f"{k}={v!r}" for k, v in vars(self).items())
        return f"{type(self).__name__}({kw})"


class DualTransform(BasicTransform):
    pass


class ImageOnlyTransform(BasicTransform):
    geometric = False


class Compose:
    def __init__(self, transforms, p=1.0, **_ignored):
        self.transforms = list(transforms)
        self.p = float(p)

    def __len__(self):
        return len(self.transforms)

    def __getitem__(self, i):
        return self.transforms[i]

    def apply_planes(self, planes, force_apply=False):
        """[planes,H,W] uint8 CUDA tensor -> same (possibly different H,W), all on the device."""
        if self.transforms and (force_apply or random.random() < self.p):
            for t in self.transforms:
                if t._fire(force_apply):
                    planes = t.apply_planes(planes, **t.get_params(planes.shape[-2], planes.shape[-1]))
        return planes

    def __call__(self, force_apply=False, **data):
        if "image" not in data:
            raise KeyError("Compose needs image=...")
        image = np.ascontiguousarray(data["image"])
        if image.dtype != np.uint8 or image.ndim != 2:
            raise TypeError("the B200 augmentation path takes 2-D uint8 images (grayscale microscopy tiles), got "
                            f"{image.dtype} {image.shape}")
        stack = [image]
        mask = data.get("mask")
        if mask is not None:
            mask = np.ascontiguousarray(mask)
            if mask.dtype != np.uint8 or mask.shape != image.shape:
                raise TypeError("mask must be a uint8 array of the image's shape")
            stack.append(mask)
        if not self.transforms:
            return dict(data)
        dev = torch.device("cuda", torch.cuda.current_device())
        planes = torch.from_numpy(np.stack(stack)).to(dev)
        planes = self.apply_planes(planes, force_apply)
        host = planes.cpu().numpy()
        out = dict(data)
        out["image"] = host[0]
        if mask is not None:
            out["mask"] = host[1]
        return out


# ------------------------------------------------------------------ geometric ----
class HorizontalFlip(DualTransform):
    def apply_planes(self, planes):
        return _ag.hflip(planes)


class VerticalFlip(DualTransform):
    def apply_planes(self, planes):
        return _ag.vflip(planes)


class Transpose(DualTransform):
    def apply_planes(self, planes):
        return _ag.transpose(planes)


class RandomRotate90(DualTransform):
    def get_params(self, h, w):
        return {"factor": random.randint(0, 3)}

    def apply_planes(self, planes, factor=0):
        return _ag.rot90(planes, factor) if factor else planes


class Affine(DualTransform):
    """A.Affine as the reference calls it (dataset.py:150-151, forward_selection_integration.py:75-83):
    scale / translate_percent / rotate / shear ranges, x and y sampled independently
    (keep_ratio=False), bilinear image + nearest mask, `border_mode` for both, constant value 0."""

    def __init__(self, scale=None, translate_percent=None, translate_px=None, rotate=None, shear=None,
                 interpolation=_CV_LINEAR, mask_interpolation=_CV_NEAREST, cval=0, cval_mask=0,
                 mode=_CV_CONSTANT, border_mode=None, fit_output=False, keep_ratio=False,
                 always_apply=False, p=0.5, **_ignored):
        super().__init__(always_apply, p)
        if fit_output or translate_px is not None:
            raise NotImplementedError("Affine(fit_output / translate_px) is not used by the reference")
        self.scale = _pair(scale, 1.0) if scale is not None and not isinstance(scale, dict) else scale or (1.0, 1.0)
        self.translate_percent = (_pair(translate_percent) if translate_percent is not None
                                  and not isinstance(translate_percent, dict) else translate_percent or (0.0, 0.0))
        self.rotate = _pair(rotate) if rotate is not None else (0.0, 0.0)
        self.shear = _pair(shear) if shear is not None and not isinstance(shear, dict) else shear or (0.0, 0.0)
        self.interpolation, self.mask_interpolation = interpolation, mask_interpolation
        self.border_mode = mode if border_mode is None else border_mode
        self.cval, self.cval_mask, self.keep_ratio = cval, cval_mask, keep_ratio

    @staticmethod
    def _xy(spec, same=False):
        if isinstance(spec, dict):
            return random.uniform(*_pair(spec["x"])), random.uniform(*_pair(spec["y"]))
        x = random.uniform(*spec)
        return (x, x) if same else (x, random.uniform(*spec))

    def get_params(self, h, w):
        sx, sy = self._xy(self.scale, self.keep_ratio)
        tx, ty = self._xy(self.translate_percent)
        shx, shy = self._xy(self.shear)
        return {"matrix": _ag.affine_matrix(h, w, scale=(sx, sy), translate=(tx, ty),
                                            rotate_deg=random.uniform(*self.rotate), shear_deg=(shx, shy))}

    def apply_planes(self, planes, matrix=None):
        return _ag.warp_affine_pair(planes, matrix, _interp(self.interpolation), _interp(self.mask_interpolation),
                                    _border(self.border_mode), int(self.cval), int(self.cval_mask))


class _RemapTransform(DualTransform):
    def __init__(self, interpolation=_CV_LINEAR, border_mode=_CV_REFLECT101, value=0, mask_value=0,
                 always_apply=False, p=0.5):
        super().__init__(always_apply, p)
        self.interpolation, self.border_mode, self.value, self.mask_value = interpolation, border_mode, value, mask_value

    def _remap(self, planes, map_x, map_y):
        # the mask takes nearest-neighbour sampling (albumentations' mask interpolation default)
        return _ag.remap_pair(planes, map_x, map_y, _interp(self.interpolation), _ag.NEAREST,
                              _border(self.border_mode), int(self.value or 0), int(self.mask_value or 0))


class GridDistortion(_RemapTransform):
    """Piecewise-linear stretch of a num_steps x num_steps grid (forward_selection_integration.py:89-95)."""

    def __init__(self, num_steps=5, distort_limit=0.3, interpolation=_CV_LINEAR, border_mode=_CV_REFLECT101,
                 value=None, mask_value=None, normalized=False, always_apply=False, p=0.5, **_ignored):
        super().__init__(interpolation, border_mode, value, mask_value, always_apply, p)
        self.num_steps, self.distort_limit = int(num_steps), _pair(distort_limit)

    def get_params(self, h, w):
        lo, hi = self.distort_limit
        return {"xsteps": [1 + random.uniform(lo, hi) for _ in range(self.num_steps + 1)],
                "ysteps": [1 + random.uniform(lo, hi) for _ in range(self.num_steps + 1)]}

    def apply_planes(self, planes, xsteps=(), ysteps=()):
        h, w = planes.shape[-2:]
        return self._remap(planes, *_ag.grid_distortion_maps(h, w, xsteps, ysteps, planes.device))


class OpticalDistortion(_RemapTransform):
    """Radial camera-model distortion (forward_selection_integration.py:99-105)."""

    def __init__(self, distort_limit=0.05, shift_limit=0.05, interpolation=_CV_LINEAR,
                 border_mode=_CV_REFLECT101, value=None, mask_value=None, always_apply=False, p=0.5, **_ignored):
        super().__init__(interpolation, border_mode, value, mask_value, always_apply, p)
        self.distort_limit, self.shift_limit = _pair(distort_limit), _pair(shift_limit)

    def get_params(self, h, w):
        return {"k": random.uniform(*self.distort_limit), "dx": round(random.uniform(*self.shift_limit)),
                "dy": round(random.uniform(*self.shift_limit))}

    def apply_planes(self, planes, k=0.0, dx=0, dy=0):
        h, w = planes.shape[-2:]
        return self._remap(planes, *_ag.optical_distortion_maps(h, w, k, dx, dy, planes.device))


class ElasticTransform(_RemapTransform):
    """Elastic warp: Gaussian-smoothed uniform displacement x alpha.  `north_star` names it; the
    reference has no call site, so its parameterisation follows albumentations' docs."""

    def __init__(self, alpha=1.0, sigma=50.0, alpha_affine=None, interpolation=_CV_LINEAR,
                 border_mode=_CV_REFLECT101, value=None, mask_value=None, always_apply=False,
                 approximate=False, same_dxdy=False, p=0.5, **_ignored):
        super().__init__(interpolation, border_mode, value, mask_value, always_apply, p)
        self.alpha, self.sigma, self.same_dxdy = float(alpha), float(sigma), bool(same_dxdy)

    def get_params(self, h, w):
        return {"seed": random.getrandbits(62)}

    def apply_planes(self, planes, seed=0):
        h, w = planes.shape[-2:]
        return self._remap(planes, *_ag.elastic_maps(h, w, self.alpha, self.sigma, seed, planes.device,
                                                     same_dxdy=self.same_dxdy))


class RandomCrop(DualTransform):
    def __init__(self, height, width, always_apply=False, p=1.0, **_ignored):
        super().__init__(always_apply, p)
        self.height, self.width = int(height), int(width)

    def get_params(self, h, w):
        if h < self.height or w < self.width:
            raise ValueError(f"Requested crop size ({self.height}, {self.width}) is larger than the image size ({h}, {w})")
        return {"y0": int((h - self.height + 1) * random.random()), "x0": int((w - self.width + 1) * random.random())}

    def apply_planes(self, planes, y0=0, x0=0):
        return _ag.crop(planes, y0, x0, self.height, self.width)


class CropAndPad(DualTransform):
    """CropAndPad(percent=(lo, hi), pad_mode, keep_size=True) (forward_selection_integration.py:174-178):
    one fraction per side; negative crops, positive pads with `pad_mode`; resized back to the input size."""

    def __init__(self, px=None, percent=None, pad_mode=_CV_CONSTANT, pad_cval=0, pad_cval_mask=0, keep_size=True,
                 sample_independently=True, interpolation=_CV_LINEAR, always_apply=False, p=1.0, **_ignored):
        super().__init__(always_apply, p)
        if px is not None or percent is None:
            raise NotImplementedError("CropAndPad(px=...) is not used by the reference")
        self.percent, self.pad_mode = _pair(percent), pad_mode
        self.pad_cval, self.pad_cval_mask, self.keep_size = pad_cval, pad_cval_mask, keep_size
        self.sample_independently, self.interpolation = sample_independently, interpolation

    def get_params(self, h, w):
        if self.sample_independently:
            fr = [random.uniform(*self.percent) for _ in range(4)]
        else:
            fr = [random.uniform(*self.percent)] * 4
        top, right, bottom, left = (int(fr[0] * h), int(fr[1] * w), int(fr[2] * h), int(fr[3] * w))
        return {"sides": (top, right, bottom, left)}

    def apply_planes(self, planes, sides=(0, 0, 0, 0)):
        return _ag.crop_and_pad(planes, sides, _border(self.pad_mode), int(self.pad_cval), int(self.pad_cval_mask),
                                self.keep_size, _interp(self.interpolation))


# -------------------------------------------------------------- non-geometric ----
class InvertImg(ImageOnlyTransform):
    def apply_planes(self, planes):
        return _ag.lut_image(planes, np.arange(255, -1, -1, dtype=np.uint8))


class Solarize(ImageOnlyTransform):
    def __init__(self, threshold=128, always_apply=False, p=0.5, **_ignored):
        super().__init__(always_apply, p)
        self.threshold = _pair(threshold, threshold) if not isinstance(threshold, (tuple, list)) else _pair(threshold)

    def get_params(self, h, w):
        return {"threshold": random.uniform(*self.threshold)}

    def apply_planes(self, planes, threshold=128):
        v = np.arange(256)
        return _ag.lut_image(planes, np.where(v >= threshold, 255 - v, v).astype(np.uint8))


class RandomGamma(ImageOnlyTransform):
    def __init__(self, gamma_limit=(80, 120), eps=None, always_apply=False, p=0.5, **_ignored):
        super().__init__(always_apply, p)
        self.gamma_limit = _pair(gamma_limit)

    def get_params(self, h, w):
        return {"gamma": random.uniform(*self.gamma_limit) / 100.0}

    def apply_planes(self, planes, gamma=1.0):
        table = (np.arange(0, 256.0 / 255, 1.0 / 255) ** gamma) * 255   # albumentations' gamma_transform (uint8)
        return _ag.lut_image(planes, table.astype(np.uint8))


class ColorJitter(ImageOnlyTransform):
    """Brightness / contrast jitter on a grayscale tile (saturation and hue do nothing there):
    img*b and (img - mean)*c + mean in random order, each rounded and clipped to uint8."""

    def __init__(self, brightness=0.2, contrast=0.2, saturation=0.2, hue=0.2, always_apply=False, p=0.5, **_ignored):
        super().__init__(always_apply, p)
        self.brightness = (max(0.0, 1 - brightness), 1 + brightness) if not isinstance(brightness, (tuple, list)) else _pair(brightness)
        self.contrast = (max(0.0, 1 - contrast), 1 + contrast) if not isinstance(contrast, (tuple, list)) else _pair(contrast)

    def get_params(self, h, w):
        return {"brightness": random.uniform(*self.brightness), "contrast": random.uniform(*self.contrast),
                "contrast_first": random.random() < 0.5}

    def apply_planes(self, planes, brightness=1.0, contrast=1.0, contrast_first=False):
        return _ag.brightness_contrast(planes, brightness, contrast, contrast_first)


class GaussNoise(ImageOnlyTransform):
    def __init__(self, var_limit=(10.0, 50.0), mean=0, per_channel=True, always_apply=False, p=0.5, **_ignored):
        super().__init__(always_apply, p)
        self.var_limit, self.mean = _pair(var_limit, 0.0), float(mean)

    def get_params(self, h, w):
        return {"sigma": math.sqrt(random.uniform(*self.var_limit)), "seed": random.getrandbits(62)}

    def apply_planes(self, planes, sigma=0.0, seed=0):
        return _ag.gauss_noise(planes, self.mean, sigma, seed)


class CoarseDropout(DualTransform):
    def __init__(self, max_holes=8, max_height=8, max_width=8, min_holes=None, min_height=None, min_width=None,
                 fill_value=0, mask_fill_value=None, always_apply=False, p=0.5, **_ignored):
        super().__init__(always_apply, p)
        self.holes = (min_holes if min_holes is not None else max_holes, max_holes)
        self.hh = (min_height if min_height is not None else max_height, max_height)
        self.ww = (min_width if min_width is not None else max_width, max_width)
        self.fill_value, self.mask_fill_value = fill_value, mask_fill_value

    def get_params(self, h, w):
        rects = []
        for _ in range(random.randint(*self.holes)):
            hh, ww = random.randint(*self.hh), random.randint(*self.ww)
            y1, x1 = random.randint(0, max(h - hh, 0)), random.randint(0, max(w - ww, 0))
            rects.append((y1, x1, min(y1 + hh, h), min(x1 + ww, w)))
        return {"rects": rects}

    def apply_planes(self, planes, rects=()):
        return _ag.fill_rects(planes, rects, int(self.fill_value), self.mask_fill_value)


class GridDropout(DualTransform):
    def __init__(self, ratio=0.5, unit_size_min=None, unit_size_max=None, holes_number_x=None, holes_number_y=None,
                 shift_x=0, shift_y=0, random_offset=False, fill_value=0, mask_fill_value=None,
                 always_apply=False, p=0.5, **_ignored):
        super().__init__(always_apply, p)
        self.ratio, self.unit = float(ratio), (unit_size_min, unit_size_max)
        self.holes_number = (holes_number_x, holes_number_y)
        self.shift, self.random_offset = (shift_x, shift_y), random_offset
        self.fill_value, self.mask_fill_value = fill_value, mask_fill_value

    def get_params(self, h, w):
        umin, umax = self.unit
        if umin is not None and umax is not None:
            uw = uh = random.randint(umin, min(umax, min(h, w) - 1) + 1 - 1)
        else:
            nx = self.holes_number[0] or 10
            ny = self.holes_number[1] or 10
            uw, uh = max(2, w // nx), max(2, h // ny)
        hole_w, hole_h = int(uw * self.ratio), int(uh * self.ratio)
        hole_w, hole_h = min(max(hole_w, 1), uw - 1), min(max(hole_h, 1), uh - 1)
        if self.random_offset:
            sx, sy = random.randint(0, uw - hole_w), random.randint(0, uh - hole_h)
        else:
            sx, sy = min(max(0, self.shift[0]), uw - hole_w), min(max(0, self.shift[1]), uh - hole_h)
        rects = []
        for i in range(w // uw + 1):
            for j in range(h // uh + 1):
                x1, y1 = min(sx + uw * i, w), min(sy + uh * j, h)
                rects.append((y1, x1, min(y1 + hole_h, h), min(x1 + hole_w, w)))
        return {"rects": rects}

    def apply_planes(self, planes, rects=()):
        return _ag.fill_rects(planes, rects, int(self.fill_value), self.mask_fill_value)


class AdvancedBlur(ImageOnlyTransform):
    """Generalised-Gaussian blur (dataset.py:153, forward_selection_integration.py:136-144): a k x k
    kernel exp(-(|x/sx|^b + |y/sy|^b)) rotated by `rotate`, multiplied by uniform noise, normalised,
    applied with cv2.filter2D's arithmetic (csrc/augment.cu filter2d_u8)."""

    def __init__(self, blur_limit=(3, 7), sigmaX_limit=(0.2, 1.0), sigmaY_limit=(0.2, 1.0), sigma_x_limit=None,
                 sigma_y_limit=None, rotate_limit=90, beta_limit=(0.5, 8.0), noise_limit=(0.9, 1.1),
                 always_apply=False, p=0.5, **_ignored):
        super().__init__(always_apply, p)
        self.blur_limit = tuple(int(v) for v in _pair(blur_limit, 3))
        self.sx = _pair(sigma_x_limit if sigma_x_limit is not None else sigmaX_limit, 0.0)
        self.sy = _pair(sigma_y_limit if sigma_y_limit is not None else sigmaY_limit, 0.0)
        self.rotate, self.beta, self.noise = _pair(rotate_limit), _pair(beta_limit, 1.0), _pair(noise_limit, 1.0)

    def get_params(self, h, w):
        k = random.randrange(self.blur_limit[0] | 1, (self.blur_limit[1] | 1) + 1, 2)
        sx, sy = random.uniform(*self.sx), random.uniform(*self.sy)
        angle = math.radians(random.uniform(*self.rotate))
        beta = random.uniform(self.beta[0], 1) if random.random() < 0.5 else random.uniform(1, self.beta[1])
        noise = np.array([[random.uniform(*self.noise) for _ in range(k)] for _ in range(k)], np.float32)
        ax = np.arange(-k // 2 + 1.0, k // 2 + 1.0)
        xx, yy = np.meshgrid(ax, ax)
        xr = xx * math.cos(angle) + yy * math.sin(angle)
        yr = -xx * math.sin(angle) + yy * math.cos(angle)
        kern = np.exp(-(np.abs(xr / max(sx, 1e-6)) ** beta + np.abs(yr / max(sy, 1e-6)) ** beta)) * noise
        kern = (kern / kern.sum()).astype(np.float32)
        return {"kernel": kern}

    def apply_planes(self, planes, kernel=None):
        return _ag.filter2d_image(planes, kernel)


class Downscale(ImageOnlyTransform):
    def __init__(self, scale_min=0.25, scale_max=0.25, interpolation=_CV_NEAREST, scale_range=None,
                 always_apply=False, p=0.5, **_ignored):
        super().__init__(always_apply, p)
        self.scale = _pair(scale_range) if scale_range is not None else (float(scale_min), float(scale_max))
        if isinstance(interpolation, dict):
            self.down, self.up = interpolation["downscale"], interpolation["upscale"]
        else:
            self.down = self.up = interpolation

    def get_params(self, h, w):
        return {"scale": random.uniform(*self.scale)}

    def apply_planes(self, planes, scale=1.0):
        return _ag.downscale_image(planes, scale, _interp(self.down), _interp(self.up))


__all__ = [n for n, v in list(globals().items()) if isinstance(v, type) and issubclass(v, (BasicTransform, Compose))]
