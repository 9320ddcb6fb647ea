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
    raise KeyError(name)


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
