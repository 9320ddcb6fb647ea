"""Metrics and helpers — API mirror of the reference `utils.py` on the B200 kernels.

Names, signatures and return types follow utils.py:9-195 so `train.py`,
`cross_validation.py` and the experiment scripts import it unchanged.  The functional losses
and `calculate_metrics` run on the fused CUDA reduction (csrc/seg_loss.cu): one pass, one
device->host copy of a 128-byte block instead of four `.item()` syncs.
"""
import os

import numpy as np
import torch

from b2seg import lossfn


def get_device():
    """The training device.  This build is B200-only: without CUDA it fails loudly instead of
    silently training on the CPU."""
    if not torch.cuda.is_available():
        raise RuntimeError("b2seg needs a CUDA (Blackwell, sm_100a) device; there is no CPU fallback")
    if int(os.environ.get("WORLD_SIZE", "1")) > 1 and "LOCAL_RANK" in os.environ:
        # one process per GPU (torchrun): this process owns cuda:LOCAL_RANK
        torch.cuda.set_device(int(os.environ["LOCAL_RANK"]) % torch.cuda.device_count())
    device = torch.device("cuda", torch.cuda.current_device())
    print("Using CUDA device")
    return device


def dice_loss(pred, target, smooth=1.0):
    """1 - (2*sum(p*t) + s) / (sum(p) + sum(t) + s) on logits `pred` (whole batch)."""
    loss, _ = lossfn.FusedSegLoss.apply(pred, target, lossfn.make_spec(w_dice=1.0, dice_smooth=smooth), None)
    return loss


def combo_loss(pred, target, alpha=0.5):
    """alpha * BCE-with-logits + (1 - alpha) * Dice."""
    spec = lossfn.make_spec(w_bce=alpha, bce_mode=0, w_dice=1.0 - alpha, dice_smooth=1.0)
    loss, _ = lossfn.FusedSegLoss.apply(pred, target, spec, None)
    return loss


def calculate_metrics(pred, target):
    """precision / recall / f1 / iou / accuracy of (pred > 0.5) against target.

    `pred` holds probabilities (the training loop passes sigmoid(outputs), train.py:53).
    Returns python floats under the reference's dict keys."""
    stats = lossfn.metrics_from_probs(pred, target)
    return lossfn.metrics_dict(stats.cpu())


def calculate_metrics_from_stats(stats):
    """Same dict from the stats block a fused loss forward left on the device
    (`criterion.last_stats`) — no second pass over the logits."""
    return lossfn.metrics_dict(stats.cpu() if torch.is_tensor(stats) else stats)


def post_process(prediction, min_size=15):
    """Binary mask (uint8) with connected components smaller than `min_size` removed."""
    from scipy import ndimage
    mask = (np.asarray(prediction) > 0.5).astype(np.uint8)
    labels, count = ndimage.label(mask)
    if count:
        areas = ndimage.sum(mask, labels, index=np.arange(1, count + 1))
        drop = np.flatnonzero(areas < min_size) + 1
        mask[np.isin(labels, drop)] = 0
    return mask


def iou_score(pred, target, apply_post_processing=False, min_size=25):
    """IoU of thresholded sigmoid(pred) against target (> 0.5)."""
    if not apply_post_processing and pred.is_cuda:
        st = lossfn.metrics_from_probs(torch.sigmoid(pred), (target > 0.5).float()).cpu()
        tp, fp, fn = float(st[10]), float(st[11]), float(st[12])
        return tp / (tp + fp + fn + 1e-7)
    prob = torch.sigmoid(pred).detach().cpu().numpy()
    truth = (target.detach().cpu().numpy() > 0.5).astype(np.uint8)
    guess = post_process(prob, min_size=min_size) if apply_post_processing else (prob > 0.5).astype(np.uint8)
    inter = np.sum(guess * truth)
    union = np.sum((guess + truth) > 0)
    return inter / (union + 1e-7)


def compute_foreground_ratio(image_type='W', img_size=(256, 256)):
    from dataset import load_original_data, CellSegmentationDataset
    data = load_original_data(image_type=image_type)
    ds = CellSegmentationDataset(data['image_paths'], data['mask_paths'], img_size=img_size)
    fg = total = 0
    for _, mask in ds:
        fg += float(mask.sum())
        total += mask.numel()
    ratio = fg / total
    print(f"Foreground (cell) pixels: {ratio*100:.2f}%")
    return ratio


def ensure_dir(directory):
    os.makedirs(directory, exist_ok=True)


class EarlyStopping:
    """Stop when the monitored score has not improved by `min_delta` for `patience` steps."""

    def __init__(self, patience=10, min_delta=0.001):
        self.patience, self.min_delta = patience, min_delta
        self.counter = 0
        self.best_score = None
        self.early_stop = False

    def step(self, current_score):
        improved = self.best_score is None or current_score > self.best_score + self.min_delta
        if improved:
            self.best_score, self.counter = current_score, 0
        else:
            self.counter += 1
            self.early_stop = self.early_stop or self.counter >= self.patience
        return self.early_stop


def save_checkpoint(path, state_dict, name):
    filename = os.path.join(path, name)
    torch.save(state_dict, filename)
    print("Saving checkpoint:", filename)
