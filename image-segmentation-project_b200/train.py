"""Epoch loops and the training driver -- API mirror of the reference `train.py`
(train_epoch :17-68, evaluate :71-112, train_model :115-244, quick_train :301-364).

Same signatures, same returned dict keys and python-float types, same per-batch weighting of the
epoch means.  What changes is where the work happens: the batch goes to the GPU once (pinned,
non-blocking), forward/backward run as planned sm_100a kernel sequences, the loss kernel
leaves the confusion counts on the device, and the host reads ONE small block per epoch instead
of five `.item()` synchronisations per batch (reference train.py:53,60 + utils.py:130-133).
"""
import copy
import os
import time
from collections import defaultdict

import torch

from b2seg import lossfn
from b2seg.optim import FusedAdam
from b2seg import _lib as L
from dataset import CellAugmenter, prepare_data
from losses import get_loss_function
from utils import EarlyStopping, calculate_metrics, get_device

_METRIC_KEYS = ("precision", "recall", "f1", "iou", "accuracy")


def _to_device(t, device):
    if t.device == device:
        return t
    if t.device.type == "cpu" and not t.is_pinned():
        t = t.pin_memory()
    return t.to(device, non_blocking=True)


class DevicePrefetcher:
    """Iterates a loader of (images, masks) host batches as device batches, one batch ahead:
    batch k+1's host->device copies run on a side stream, into one of two persistent device
    buffer sets, while step k computes -- the PCIe transfer (33 MB per 64 x 256^2 fp32 batch)
    leaves the step's critical path and no allocation happens per batch.  The reference copies
    inside the loop body on the compute stream (train.py:40-41).  A yielded batch is valid until
    the next one is requested."""

    def __init__(self, loader, device):
        self.loader, self.device = loader, torch.device(device)
        if self.device.type == "cuda" and self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())

    def _stage(self, batch, slot, side):
        if self._done[slot] is not None:
            side.wait_event(self._done[slot])  # the step that last read this slot has been enqueued
        if all(torch.is_tensor(t) and t.device == self.device for t in batch):
            return tuple(batch)  # already resident (dataset.DeviceBatchLoader): nothing to copy
        out = []
        with torch.cuda.stream(side):
            for j, t in enumerate(batch):
                if t.device.type == "cpu" and not t.is_pinned():
                    t = t.pin_memory()
                buf = self._bufs[slot].get(j)
                if buf is None or buf.dtype != t.dtype or buf.shape[1:] != t.shape[1:] or buf.shape[0] < t.shape[0]:
                    buf = torch.empty(t.shape, dtype=t.dtype, device=self.device)
                    self._bufs[slot][j] = buf
                dst = buf[:t.shape[0]]
                dst.copy_(t, non_blocking=True)
                out.append(dst)
        return tuple(out)

    def __iter__(self):
        if self.device.type != "cuda":
            raise L.B2SegError("DevicePrefetcher needs a CUDA device (there is no CPU path)")
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        self._bufs, self._done = [{}, {}], [None, None]
        events = [torch.cuda.Event(), torch.cuda.Event()]  # re-recorded: no per-batch allocation
        it = iter(self.loader)
        k = 0
        try:
            staged = self._stage(next(it), 0, side)
        except StopIteration:
            return
        while staged is not None:
            cur = torch.cuda.current_stream(self.device)
            cur.wait_stream(side)
            ready = staged
            try:
                staged = self._stage(next(it), (k + 1) & 1, side)
            except StopIteration:
                staged = None
            yield ready
            events[k & 1].record(torch.cuda.current_stream(self.device))
            self._done[k & 1] = events[k & 1]
            k += 1


def _batch_stats(criterion, outputs, masks):
    """Device stats block (f64[16]) holding this batch's confusion counts.  The fused loss
    kernel already produced it for exactly these logits; otherwise one metrics launch."""
    st = getattr(criterion, "last_stats", None)
    if st is not None and getattr(criterion, "last_logits_ref", None) is outputs:
        return st
    return lossfn.metrics_from_probs(torch.sigmoid(outputs.detach()), masks)


def train_step(model, images, masks, optimizer, criterion):
    """One optimisation step on a device batch: forward, loss, backward, optimizer.step().
    Returns (loss 0-dim tensor, stats f64[16]) without synchronising the host."""
    outputs = model(images)
    loss = criterion(outputs, masks)
    optimizer.zero_grad()
    loss.backward()
    optimizer.step()
    return loss.detach(), _batch_stats(criterion, outputs, masks)


def _finish_epoch(losses, stats, sizes):
    """Per-batch metrics weighted by batch size (reference :56-66), from one device->host copy."""
    out = defaultdict(float)
    if not sizes:
        raise ZeroDivisionError("empty data loader")
    host_stats = torch.stack(stats).cpu()
    host_loss = torch.stack([l.reshape(()).double() for l in losses]).cpu()
    total = float(sum(sizes))
    for st, lv, bs in zip(host_stats, host_loss, sizes):
        for k, v in lossfn.metrics_dict(st).items():
            out[k] += v * bs
        out["loss"] += float(lv) * bs
    for k in list(out):
        out[k] /= total
    return out


def train_epoch(model, loader, optimizer, criterion, device):
    """Train one epoch; returns {'precision','recall','f1','iou','accuracy','loss'} (floats)."""
    model.train()
    device = torch.device(device)
    losses, stats, sizes = [], [], []
    for images, masks in DevicePrefetcher(loader, device):
        loss, st = train_step(model, images, masks, optimizer, criterion)
        losses.append(loss)
        stats.append(st)
        sizes.append(images.size(0))
    return _finish_epoch(losses, stats, sizes)


def evaluate(model, loader, device, criterion):
    """Evaluate on a loader (model.eval(), no_grad); same dict as train_epoch."""
    model.eval()
    device = torch.device(device)
    losses, stats, sizes = [], [], []
    with torch.no_grad():
        for images, masks in DevicePrefetcher(loader, device):
            outputs = model(images)
            loss = criterion(outputs, masks)
            losses.append(loss.detach())
            stats.append(_batch_stats(criterion, outputs, masks))
            sizes.append(images.size(0))
    return _finish_epoch(losses, stats, sizes)


def train_model(model, train_images, train_masks, val_images, val_masks, criterion, optimizer,
                scheduler, num_epochs, device, config, augmentations_per_image=0, save_plots=True):
    """Full training run: optional offline augmentation, epochs with validation, LR schedule on
    validation IoU, best-state tracking, early stopping.  Returns the reference's result dict."""
    augmenter = CellAugmenter(augmentations_per_image)
    if augmentations_per_image > 0:
        tr_images, tr_masks = augmenter.augment_training_data(train_images, train_masks)
    else:
        tr_images, tr_masks = train_images, train_masks
    train_loader = prepare_data(tr_images, tr_masks, config['batch_size'], config['img_size'], shuffle=True)
    val_loader = prepare_data(val_images, val_masks, config['batch_size'], config['img_size'], shuffle=False)

    verbose = config.get('verbose', True)
    history = {"train": [], "val": [], "lr": []}
    best_iou, best_state, best_epoch = 0.0, None, 0
    stopper = EarlyStopping(patience=config.get('early_stopping_patience', 7),
                            min_delta=config.get('early_stopping_min_delta', 0.001))
    if verbose:
        print("Starting training...")
    t0 = time.time()
    train_metrics = val_metrics = None
    for epoch in range(num_epochs):
        train_metrics = train_epoch(model, train_loader, optimizer, criterion, device)
        val_metrics = evaluate(model, val_loader, device, criterion)
        history["train"].append(train_metrics)
        history["val"].append(val_metrics)
        if scheduler is not None:
            if isinstance(scheduler, torch.optim.lr_scheduler.ReduceLROnPlateau):
                scheduler.step(val_metrics['iou'])
            else:
                scheduler.step()
        lr = optimizer.param_groups[0]['lr']
        history["lr"].append(lr)
        if verbose:
            print(f"Epoch {epoch+1:3d}/{num_epochs} - Train Loss: {train_metrics['loss']:.4f}, "
                  f"Train IoU: {train_metrics['iou']:.4f}, Val Loss: {val_metrics['loss']:.4f}, "
                  f"Val IoU: {val_metrics['iou']:.4f}, LR: {lr:.6f}")
        if val_metrics['iou'] > best_iou:
            best_iou, best_epoch = val_metrics['iou'], epoch
            best_state = copy.deepcopy(model.state_dict())
        if stopper.step(val_metrics['iou']):
            if verbose:
                print(f"Early stopping triggered at epoch {epoch+1}")
            break
    elapsed = time.time() - t0
    if verbose:
        print(f"Training complete in {elapsed // 60:.0f}m {elapsed % 60:.0f}s")
        print(f"Best validation IoU: {best_iou:.4f} at epoch {best_epoch+1}")
    if best_state is not None:
        model.load_state_dict(best_state)
    if save_plots and config.get('save_plots', False):
        plot_training_history(history["train"], history["val"], history["lr"], config)
    augmenter.cleanup()
    return {
        'train_metrics': history["train"],
        'val_metrics': history["val"],
        'lr_history': history["lr"],
        'best_iou': best_iou,
        'best_epoch': best_epoch,
        'best_model_state': best_state,
        'training_time': elapsed,
        'final_train_metrics': train_metrics,
        'final_val_metrics': val_metrics,
    }


def plot_training_history(train_metrics, val_metrics, lr_history, config):
    """Loss / IoU / F1 / LR curves as PNG+PDF under config['save_dir'] (needs matplotlib)."""
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except ImportError:
        print("matplotlib is not installed; skipping training plots")
        return
    epochs = range(1, len(train_metrics) + 1)
    fig, axes = plt.subplots(2, 2, figsize=(15, 10))
    panels = (("loss", "Loss"), ("iou", "IoU"), ("f1", "F1 Score"))
    for ax, (key, title) in zip(axes.flat, panels):
        ax.plot(epochs, [m[key] for m in train_metrics], label="Train", marker="o", markersize=3)
        ax.plot(epochs, [m[key] for m in val_metrics], label="Validation", marker="s", markersize=3)
        ax.set_title(f"Training and Validation {title}")
        ax.set_xlabel("Epoch")
        ax.set_ylabel(title)
        ax.legend()
        ax.grid(True, alpha=0.3)
    ax = axes[1, 1]
    ax.plot(epochs, lr_history, color="red", marker="o", markersize=3)
    ax.set_title("Learning Rate")
    ax.set_xlabel("Epoch")
    ax.set_yscale("log")
    ax.grid(True, alpha=0.3)
    plt.tight_layout()
    save_dir = config.get('save_dir', 'experiments')
    os.makedirs(save_dir, exist_ok=True)
    name = config.get('name', 'model').replace(' ', '_').replace('/', '_')
    plt.savefig(os.path.join(save_dir, f'training_history_{name}.png'), dpi=300, bbox_inches='tight')
    plt.savefig(os.path.join(save_dir, f'training_history_{name}.pdf'), bbox_inches='tight')
    plt.close()


def quick_train(model, train_images, train_masks, val_images, val_masks, config, device=None,
                augmentations_per_image=0):
    """Standard setup (reference :301-364): loss from `config`, Adam(lr 1e-3, wd 1e-5),
    ReduceLROnPlateau on validation IoU (factor 0.5, threshold 0.01, min_lr 1e-6), then train_model."""
    if device is None:
        device = get_device()
    model = model.to(device)
    criterion = get_loss_function(config)
    # a torch.optim.Adam whose step() is one kernel launch (b2seg.optim / csrc/adam.cu)
    optimizer = FusedAdam(model.parameters(), lr=config.get('learning_rate', 1e-3),
                          weight_decay=config.get('weight_decay', 1e-5))
    scheduler = torch.optim.lr_scheduler.ReduceLROnPlateau(
        optimizer, mode='max', factor=0.5, patience=config.get('scheduler_patience', 3),
        threshold=0.01, min_lr=1e-6)
    return train_model(model=model, train_images=train_images, train_masks=train_masks,
                       val_images=val_images, val_masks=val_masks, criterion=criterion,
                       optimizer=optimizer, scheduler=scheduler,
                       num_epochs=config.get('num_epochs', 50), device=device, config=config,
                       augmentations_per_image=augmentations_per_image,
                       save_plots=config.get('save_plots', True))


if __name__ == "__main__":
    print("train.py: epoch loops of the B200 segmentation training step")
