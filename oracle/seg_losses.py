"""CPU restatement of the reference's closed-form losses and metrics (oracle; tests only).

Computed in float64 from first principles so that it is a *truth* the fp32
reference and the CUDA kernels can both be compared with.  Each branch cites the
reference lines it restates.
"""
import math

import numpy as np
import torch

FACTORY_DEFAULTS = {  # losses.py:345-396 (factory defaults, which differ from class defaults)
    "smooth": 1.0, "beta": 2.0, "focal_alpha": 0.25, "focal_gamma": 2.0, "loss_alpha": 0.5,
    "alpha_dice": 0.33, "alpha_bce": 0.33, "alpha_focal": 0.34, "tversky_alpha": 0.5,
    "tversky_beta": 0.5, "focal_tversky_gamma": 0.75, "sensitivity_weight": 0.5,
    "w_dice": 0.5, "w_cross": 0.5, "gamma_dice": 0.3, "gamma_cross": 0.3, "distance_alpha": 1.0,
    "boundary_theta0": 3, "boundary_theta": 5,
}

CLOSED_FORM = ("dice", "bce", "weighted_bce", "balanced_bce", "focal", "combo", "triple_combo",
               "tversky", "tversky_balanced", "tversky_recall", "focal_tversky",
               "sensitivity_specificity", "log_cosh_dice")


def _cfg(config, key):
    return config.get(key, FACTORY_DEFAULTS[key])


def _terms(x, t):
    x = x.double().reshape(-1)
    t = t.double().reshape(-1)
    p = torch.sigmoid(x)
    # F.binary_cross_entropy_with_logits, elementwise
    bce = torch.clamp(x, min=0) - x * t + torch.log1p(torch.exp(-x.abs()))
    return x, t, p, bce


def _dice(p, t, smooth):  # losses.py:19-28, utils.py:26-43
    inter = (p * t).sum()
    return 1 - (2 * inter + smooth) / (p.sum() + t.sum() + smooth)


def _tversky(p, t, a, b, smooth=1.0):  # losses.py:97-109
    tp = (p * t).sum()
    fp = ((1 - t) * p).sum()
    fn = (t * (1 - p)).sum()
    return 1 - (tp + smooth) / (tp + a * fp + b * fn + smooth)


def _focal(bce, alpha, gamma):  # losses.py:76-82 (alpha is NOT class dependent)
    pt = torch.exp(-bce)
    return (alpha * (1 - pt) ** gamma * bce).mean()


def loss_value(name, logits, target, config=None):
    """Scalar float64 tensor (differentiable w.r.t. logits) for a reference loss name."""
    config = config or {}
    x, t, p, bce = _terms(logits, target)
    if name == "dice":
        return _dice(p, t, _cfg(config, "smooth"))
    if name == "bce":  # losses.py:36-37
        return bce.mean()
    if name == "weighted_bce":  # losses.py:46-50
        beta = _cfg(config, "beta")
        return (bce * (t * beta + (1 - t))).mean()
    if name == "balanced_bce":  # losses.py:58-65
        beta = 1 - t.sum() / t.numel()
        return (bce * (t * beta + (1 - t) * (1 - beta))).mean()
    if name == "focal":
        return _focal(bce, _cfg(config, "focal_alpha"), _cfg(config, "focal_gamma"))
    if name == "combo":  # losses.py:168-171, utils.py:45-56 (ComboLoss smooth stays 1.0)
        a = _cfg(config, "loss_alpha")
        return a * bce.mean() + (1 - a) * _dice(p, t, 1.0)
    if name == "triple_combo":  # losses.py:335-342
        return (_cfg(config, "alpha_dice") * _dice(p, t, 1.0) + _cfg(config, "alpha_bce") * bce.mean()
                + _cfg(config, "alpha_focal") * _focal(bce, _cfg(config, "focal_alpha"),
                                                       _cfg(config, "focal_gamma")))
    if name == "tversky":
        return _tversky(p, t, _cfg(config, "tversky_alpha"), _cfg(config, "tversky_beta"))
    if name == "tversky_balanced":  # losses.py:370
        return _tversky(p, t, 0.5, 0.5)
    if name == "tversky_recall":  # losses.py:371
        return _tversky(p, t, 0.3, 0.7)
    if name == "focal_tversky":  # losses.py:119-122
        tv = _tversky(p, t, _cfg(config, "tversky_alpha"), _cfg(config, "tversky_beta"))
        return tv ** _cfg(config, "focal_tversky_gamma")
    if name == "sensitivity_specificity":  # losses.py:132-147
        w, eps = _cfg(config, "sensitivity_weight"), 1e-7
        tp = (p * t).sum()
        tn = ((1 - p) * (1 - t)).sum()
        fp = ((1 - t) * p).sum()
        fn = (t * (1 - p)).sum()
        return 1 - (w * tp / (tp + fn + eps) + (1 - w) * tn / (tn + fp + eps))
    if name == "log_cosh_dice":  # losses.py:156-158
        return torch.log(torch.cosh(_dice(p, t, _cfg(config, "smooth"))))
    if name == "exponential_logarithmic":  # losses.py:184-194
        # In float32 ON PURPOSE, through probabilities as the reference does: for confident pixels
        # `1 - cross_entropy` rounds to a few float32 steps below 1, and the slope of (-log .)^0.3 there
        # is steep, so the reference's own rounding is part of its answer (it moves the gradient's L1
        # norm by 0.6 % on the known-answer inputs); an exact-arithmetic restatement would not be it.
        import torch.nn.functional as F
        xf, tf = logits.float().reshape(-1), target.float().reshape(-1)
        pf = torch.sigmoid(xf)
        inter = (pf * tf).sum()
        d = 1 - (2.0 * inter + 1.0) / (pf.sum() + tf.sum() + 1.0)
        e_dice = torch.pow(-torch.log(torch.clamp(1 - d, min=1e-7)), _cfg(config, "gamma_dice"))
        ce = F.binary_cross_entropy(pf, tf, reduction="none")
        e_cross = torch.pow(-torch.log(torch.clamp(1 - ce, min=1e-7)), _cfg(config, "gamma_cross")).mean()
        return (_cfg(config, "w_dice") * e_dice + _cfg(config, "w_cross") * e_cross).double()
    if name == "distance_map_bce":  # losses.py:203-214
        return ((1 + _cfg(config, "distance_alpha") * distance_map(target).double().reshape(-1)) * bce).mean()
    if name == "boundary":  # losses.py:283-307
        wm = boundary_weight_map(target, _cfg(config, "boundary_theta0"), _cfg(config, "boundary_theta"))
        return (wm.double().reshape(-1) * bce).mean()
    raise KeyError(name)


MAP_LOSSES = ("distance_map_bce", "boundary", "exponential_logarithmic")


def edt(plane):
    """scipy.ndimage.distance_transform_edt, the call the reference makes (losses.py:226, 270;
    scipy >= 1.7, requirements.txt): float64 distances of the non-zero pixels to the nearest zero."""
    from scipy.ndimage import distance_transform_edt
    return distance_transform_edt(np.asarray(plane))


def edt_brute(plane):
    """The definition, by exhaustion (small planes only): pins what `edt` means, including scipy's
    answer for a plane without zeros (distances to an imaginary pixel at row -1, column 0)."""
    a = np.asarray(plane)
    h, w = a.shape
    zy, zx = np.nonzero(a == 0)
    if zy.size == 0:
        zy, zx = np.array([-1]), np.array([0])
    ys, xs = np.mgrid[0:h, 0:w]
    d2 = ((ys[..., None] - zy) ** 2 + (xs[..., None] - zx) ** 2).min(axis=-1)
    return np.where(a != 0, np.sqrt(d2.astype(np.float64)), 0.0)


def distance_map(target):
    """DistanceMapPenalizedCrossEntropy._compute_distance_map (losses.py:216-232): float32 [B,C,H,W]."""
    host = target.detach().cpu().numpy()
    out = np.empty(host.shape, dtype=np.float64)
    for b in range(host.shape[0]):
        for c in range(host.shape[1]):
            d = edt(host[b, c])
            out[b, c] = d / (d.max() + 1e-7)
    return torch.from_numpy(out).float()


def hausdorff_dtm(img):
    """HausdorffDistanceLoss._compute_dtm (losses.py:262-273)."""
    host = img.detach().cpu().numpy()
    field = torch.zeros(host.shape, dtype=torch.float32)
    for b in range(host.shape[0]):
        for c in range(host.shape[1]):
            if host[b, c].max() > 0:
                field[b, c] = torch.from_numpy(edt(host[b, c])).float()
    return field / (field.max() + 1e-7)


def gaussian_kernel(size, sigma):
    """BoundaryLoss._gaussian_kernel (losses.py:309-318)."""
    c = torch.arange(size).float() - (size - 1) / 2.0
    g = c ** 2
    g = (-(g.unsqueeze(0) + g.unsqueeze(1)) / (2 * sigma ** 2)).exp()
    return g / g.sum()


def boundary_weight_map(target, theta0=3, theta=5):
    """BoundaryLoss weight map (losses.py:285-301), float32 on the CPU."""
    import torch.nn.functional as F
    t = target.detach().float().cpu()
    dil = F.max_pool2d(t, kernel_size=theta0, stride=1, padding=theta0 // 2)
    ero = -F.max_pool2d(-t, kernel_size=theta0, stride=1, padding=theta0 // 2)
    k = gaussian_kernel(theta, theta / 3.0).view(1, 1, theta, theta).repeat(t.size(1), 1, 1, 1)
    return 1 + 10 * F.conv2d(dil - ero, k, padding=theta // 2, groups=t.size(1))


def loss_and_grad(name, logits, target, config=None):
    x = logits.detach().double().clone().requires_grad_(True)
    val = loss_value(name, x, target, config)
    (g,) = torch.autograd.grad(val, x)
    return float(val), g.reshape(logits.shape)


def confusion_metrics(probs, target):
    """utils.py:120-151 on probabilities: threshold > 0.5 (on the fp32 value), eps 1e-7."""
    pb = (probs.float() > 0.5).double().reshape(-1)
    t = target.double().reshape(-1)
    tp = float((pb * t).sum())
    fp = float((pb * (1 - t)).sum())
    fn = float(((1 - pb) * t).sum())
    tn = float(((1 - pb) * (1 - t)).sum())
    return metrics_from_counts(tp, fp, fn, tn)


def metrics_from_counts(tp, fp, fn, tn, eps=1e-7):  # utils.py:136-151
    precision = tp / (tp + fp + eps)
    recall = tp / (tp + fn + eps)
    return {
        "precision": precision,
        "recall": recall,
        "f1": 2 * precision * recall / (precision + recall + eps),
        "iou": tp / (tp + fp + fn + eps),
        "accuracy": (tp + tn) / (tp + tn + fp + fn + eps),
    }


def kat_inputs(shape=(4, 1, 256, 256), seed=1234):
    """The SURVEY section-4 known-answer inputs."""
    torch.manual_seed(seed)
    logits = torch.randn(*shape) * 2.0
    target = (torch.rand(*shape) > 0.9).float()
    return logits, target
