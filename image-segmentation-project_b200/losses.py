"""Segmentation losses — API mirror of the reference `losses.py`, B200-native underneath.

Same class names, constructor arguments, defaults and `get_loss_function(config)` keys as
the reference (losses.py:13-403), so `cross_validation.py`, `train.py` and the experiment
scripts use it unchanged.  The thirteen closed-form losses are thin front-ends of ONE fused
CUDA reduction (b2seg.lossfn / csrc/seg_loss.cu): each instance only describes its weights
in a `b2seg_loss_spec`.  The map-weighted losses (distance-map, boundary) feed a per-pixel
weight map to the same kernel; the two remaining exotic losses are composites.

Every forward also leaves the batch confusion counts on the device in `last_stats`, which
`train.train_epoch` / `utils.calculate_metrics_from_stats` read instead of re-reducing.
"""
import math

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from b2seg import lossfn
from b2seg._lib import B2SegError


class _FusedLoss(nn.Module):
    """Base: subclasses return the kernel spec; forward launches the fused kernel."""

    def __init__(self):
        super().__init__()
        self.last_stats = None
        self.last_logits_ref = None
        self.dp_group = None  # set by b2seg.dataparallel.attach: global-batch loss across replicas

    def _spec(self):
        raise NotImplementedError

    def _weight_map(self, pred, target):
        return None

    def forward(self, pred, target):
        loss, stats = lossfn.FusedSegLoss.apply(pred, target, self._spec(), self._weight_map(pred, target),
                                                self.dp_group)
        self.last_stats = stats
        self.last_logits_ref = pred
        return loss


class DiceLoss(_FusedLoss):
    def __init__(self, smooth=1.0):
        super().__init__()
        self.smooth = smooth

    def _spec(self):
        return lossfn.make_spec(w_dice=1.0, dice_smooth=self.smooth)


class BCELoss(_FusedLoss):
    def _spec(self):
        return lossfn.make_spec(w_bce=1.0, bce_mode=0)


class WeightedBCELoss(_FusedLoss):
    def __init__(self, beta=1.0):
        super().__init__()
        self.beta = beta

    def _spec(self):
        return lossfn.make_spec(w_bce=1.0, bce_mode=1, bce_beta=self.beta)


class BalancedBCELoss(_FusedLoss):
    def _spec(self):
        return lossfn.make_spec(w_bce=1.0, bce_mode=2)


class FocalLoss(_FusedLoss):
    def __init__(self, alpha=0.25, gamma=2.0, reduction='mean'):
        super().__init__()
        self.alpha, self.gamma, self.reduction = alpha, gamma, reduction

    def _spec(self):
        return lossfn.make_spec(w_focal=1.0, focal_alpha=self.alpha, focal_gamma=self.gamma)

    def forward(self, pred, target):
        if self.reduction == 'mean':
            return super().forward(pred, target)
        if self.reduction == 'sum':
            return super().forward(pred, target) * pred.numel()
        # reduction='none' returns the per-pixel map: not a reduction, no kernel to fuse
        bce = F.binary_cross_entropy_with_logits(pred, target, reduction='none')
        return self.alpha * (1 - torch.exp(-bce)) ** self.gamma * bce


class TverskyLoss(_FusedLoss):
    def __init__(self, alpha=0.5, beta=0.5, smooth=1.0):
        super().__init__()
        self.alpha, self.beta, self.smooth = alpha, beta, smooth

    def _spec(self):
        return lossfn.make_spec(w_tversky=1.0, tv_alpha=self.alpha, tv_beta=self.beta,
                                tv_smooth=self.smooth, tv_gamma=1.0)


class FocalTverskyLoss(_FusedLoss):
    def __init__(self, alpha=0.5, beta=0.5, gamma=1.0, smooth=1.0):
        super().__init__()
        self.tversky = TverskyLoss(alpha=alpha, beta=beta, smooth=smooth)
        self.gamma = gamma

    def _spec(self):
        t = self.tversky
        return lossfn.make_spec(w_tversky=1.0, tv_alpha=t.alpha, tv_beta=t.beta, tv_smooth=t.smooth,
                                tv_gamma=self.gamma)


class SensitivitySpecificityLoss(_FusedLoss):
    def __init__(self, w=0.5, smooth=1e-7):
        super().__init__()
        self.w, self.smooth = w, smooth

    def _spec(self):
        return lossfn.make_spec(w_sens_spec=1.0, ss_w=self.w, ss_eps=self.smooth)


class LogCoshDiceLoss(_FusedLoss):
    def __init__(self, smooth=1.0):
        super().__init__()
        self.dice_loss = DiceLoss(smooth=smooth)

    def _spec(self):
        return lossfn.make_spec(w_logcosh_dice=1.0, dice_smooth=self.dice_loss.smooth)


class ComboLoss(_FusedLoss):
    """alpha * BCE + (1 - alpha) * Dice — the default criterion."""

    def __init__(self, alpha=0.5, smooth=1.0):
        super().__init__()
        self.alpha = alpha
        self.dice_loss = DiceLoss(smooth=smooth)

    def _spec(self):
        return lossfn.make_spec(w_bce=self.alpha, bce_mode=0, w_dice=1.0 - self.alpha,
                                dice_smooth=self.dice_loss.smooth)


class TripleComboLoss(_FusedLoss):
    def __init__(self, alpha_dice=0.33, alpha_bce=0.33, alpha_focal=0.34,
                 focal_alpha=0.25, gamma=2.0, smooth=1.0):
        super().__init__()
        self.alpha_dice, self.alpha_bce, self.alpha_focal = alpha_dice, alpha_bce, alpha_focal
        self.dice_loss = DiceLoss(smooth=smooth)
        self.bce_loss = BCELoss()
        self.focal_loss = FocalLoss(alpha=focal_alpha, gamma=gamma)

    def _spec(self):
        return lossfn.make_spec(w_dice=self.alpha_dice, dice_smooth=self.dice_loss.smooth,
                                w_bce=self.alpha_bce, bce_mode=0, w_focal=self.alpha_focal,
                                focal_alpha=self.focal_loss.alpha, focal_gamma=self.focal_loss.gamma)


# ---- losses whose per-pixel weights come from a target-only map ------------------------------
class _MapWeightedBCE(_FusedLoss):
    """mean(weight_map * bce): the map depends on the target only (no gradient), the
    reduction and the gradient run in the fused kernel (bce_mode 3)."""

    def _spec(self):
        return lossfn.make_spec(w_bce=1.0, bce_mode=3)


class DistanceMapPenalizedCrossEntropy(_MapWeightedBCE):
    """mean((1 + alpha * edt(t) / max(edt(t))) * bce)  (losses.py:197-232).  The exact Euclidean
    distance transform runs on the device (csrc/map_loss.cu), bit-identical to the scipy path."""

    def __init__(self, alpha=1.0):
        super().__init__()
        self.alpha = alpha

    def _compute_distance_map(self, target):
        return (lossfn.distance_transform(target, mode=1, alpha=1.0) - 1.0).view(target.shape)

    def _weight_map(self, pred, target):
        return lossfn.distance_transform(target, mode=1, alpha=self.alpha).view(target.shape)


class BoundaryLoss(_MapWeightedBCE):
    """mean((1 + 10 * gauss(dilate(t) - erode(t))) * bce)  (losses.py:276-318); the weight map is one
    fused kernel (csrc/map_loss.cu)."""

    def __init__(self, theta0=3, theta=5):
        super().__init__()
        self.theta0, self.theta = theta0, theta

    def _gaussian_kernel(self, size, sigma):
        ax = torch.arange(size).float()
        ax -= (size - 1) / 2.0
        g = ax ** 2
        g = (-(g.unsqueeze(0) + g.unsqueeze(1)) / (2 * sigma ** 2)).exp()
        g /= g.sum()
        return g

    def _weight_map(self, pred, target):
        k = self._gaussian_kernel(self.theta, self.theta / 3.0)
        return lossfn.boundary_weight_map(target, self.theta0, self.theta, k, 10.0).view(target.shape)


class ExponentialLogarithmicLoss(_FusedLoss):
    """w_dice * (-log(1 - dice_loss))^g_d + w_cross * mean((-log(clamp(1 - bce_i, 1e-7)))^g_c)
    (losses.py:174-194): both parts, and the gradient, in the fused kernel."""

    def __init__(self, w_dice=0.5, w_cross=0.5, gamma_dice=0.3, gamma_cross=0.3):
        super().__init__()
        self.w_dice, self.w_cross = w_dice, w_cross
        self.gamma_dice, self.gamma_cross = gamma_dice, gamma_cross

    def _spec(self):
        return lossfn.make_spec(w_explog_dice=self.w_dice, w_explog_cross=self.w_cross,
                                el_gamma_dice=self.gamma_dice, el_gamma_cross=self.gamma_cross, dice_smooth=1.0)


class HausdorffDistanceLoss(nn.Module):
    """(sigmoid(x) - t)^2 * (dtm(pred)^a + dtm(target)^a)  (losses.py:235-273).  The reference version
    cannot be differentiated (it calls .numpy() on a tensor that requires grad, losses.py:268); here
    the distance maps are constants, which is what the formula intends.  The distance transforms run
    on the device, including scipy's answer for the zero-free sigmoid planes."""

    def __init__(self, alpha=2.0, reduction='mean'):
        super().__init__()
        self.alpha, self.reduction = alpha, reduction
        self.last_stats = None
        self.last_logits_ref = None

    @torch.no_grad()
    def _compute_dtm(self, img):
        field = lossfn.distance_transform(img, mode=0).view(img.shape)
        keep = (img.flatten(2).amax(dim=2) > 0).to(field.dtype)[:, :, None, None]  # losses.py:269
        field = field * keep
        return field / (field.max() + 1e-7)

    def forward(self, pred, target):
        prob = torch.sigmoid(pred)
        dist = self._compute_dtm(prob) ** self.alpha + self._compute_dtm(target) ** self.alpha
        hd = (prob - target) ** 2 * dist
        if self.reduction == 'mean':
            return hd.mean()
        if self.reduction == 'sum':
            return hd.sum()
        return hd


_FACTORY = {
    'dice': lambda c: DiceLoss(smooth=c.get('smooth', 1.0)),
    'bce': lambda c: BCELoss(),
    'weighted_bce': lambda c: WeightedBCELoss(beta=c.get('beta', 2.0)),
    'balanced_bce': lambda c: BalancedBCELoss(),
    'focal': lambda c: FocalLoss(alpha=c.get('focal_alpha', 0.25), gamma=c.get('focal_gamma', 2.0)),
    'combo': lambda c: ComboLoss(alpha=c.get('loss_alpha', 0.5)),
    'triple_combo': lambda c: TripleComboLoss(
        alpha_dice=c.get('alpha_dice', 0.33), alpha_bce=c.get('alpha_bce', 0.33),
        alpha_focal=c.get('alpha_focal', 0.34), focal_alpha=c.get('focal_alpha', 0.25),
        gamma=c.get('focal_gamma', 2.0)),
    'tversky': lambda c: TverskyLoss(alpha=c.get('tversky_alpha', 0.5), beta=c.get('tversky_beta', 0.5)),
    'tversky_balanced': lambda c: TverskyLoss(alpha=0.5, beta=0.5),
    'tversky_recall': lambda c: TverskyLoss(alpha=0.3, beta=0.7),
    'focal_tversky': lambda c: FocalTverskyLoss(
        alpha=c.get('tversky_alpha', 0.5), beta=c.get('tversky_beta', 0.5),
        gamma=c.get('focal_tversky_gamma', 0.75)),
    'sensitivity_specificity': lambda c: SensitivitySpecificityLoss(w=c.get('sensitivity_weight', 0.5)),
    'log_cosh_dice': lambda c: LogCoshDiceLoss(smooth=c.get('smooth', 1.0)),
    'exponential_logarithmic': lambda c: ExponentialLogarithmicLoss(
        w_dice=c.get('w_dice', 0.5), w_cross=c.get('w_cross', 0.5),
        gamma_dice=c.get('gamma_dice', 0.3), gamma_cross=c.get('gamma_cross', 0.3)),
    'distance_map_bce': lambda c: DistanceMapPenalizedCrossEntropy(alpha=c.get('distance_alpha', 1.0)),
    'hausdorff': lambda c: HausdorffDistanceLoss(alpha=c.get('hausdorff_alpha', 2.0)),
    'boundary': lambda c: BoundaryLoss(theta0=c.get('boundary_theta0', 3), theta=c.get('boundary_theta', 5)),
}


def get_loss_function(config):
    """config['loss_fn'] -> loss module; unknown names fall back to ComboLoss with a warning."""
    name = config.get('loss_fn', 'combo')
    make = _FACTORY.get(name)
    if make is None:
        print(f"Warning: Unknown loss function '{name}', defaulting to ComboLoss")
        return ComboLoss(alpha=config.get('loss_alpha', 0.5))
    return make(config)


LOSS_NAMES = tuple(_FACTORY)
