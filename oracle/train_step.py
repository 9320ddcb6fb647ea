"""CPU training step of the reference path (oracle; test / baseline infrastructure only).

One step = what the reference's inner loop does per batch (train.py:43-60): forward of
UNetWithBackbone, the criterion, `loss.backward()`, `torch.optim.Adam.step()` (built as
cross_validation.py:93-97 builds it) and `calculate_metrics` -- evaluated through the
functional restatement in oracle/unet_functional.py, i.e. the same ATen fp32 CPU kernels
(MKLDNN conv, native batch norm) the reference's nn.Modules dispatch to.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` leg may
import this.
"""
import torch

from . import seg_losses, unet_functional as uf


class OracleTrainer:
    def __init__(self, state_dict, backbone="resnet34", use_attention=True, loss_name="combo",
                 lr=1e-3, weight_decay=1e-5):
        self.backbone, self.attn, self.loss_name = backbone, use_attention, loss_name
        self.sd = {}
        for k, v in state_dict.items():
            v = v.detach().clone().cpu()
            if v.is_floating_point() and "running_" not in k:
                v.requires_grad_(True)
            self.sd[k] = v
        self.leaves = [v for v in self.sd.values() if v.requires_grad]
        self.opt = torch.optim.Adam(self.leaves, lr=lr, weight_decay=weight_decay)

    def _loss(self, logits, masks):
        """fp32 arithmetic exactly as losses.py writes it for the common names (so the timed
        baseline does not pay for the float64 of the checker); float64 restatement otherwise."""
        import torch.nn.functional as F
        name = self.loss_name
        if name in ("combo", "bce", "dice"):
            bce = F.binary_cross_entropy_with_logits(logits, masks) if name != "dice" else None
            if name == "bce":
                return bce
            p, t = torch.sigmoid(logits).view(-1), masks.view(-1)  # losses.py:19-28
            dice = 1 - (2.0 * (p * t).sum() + 1.0) / (p.sum() + t.sum() + 1.0)
            return dice if name == "dice" else 0.5 * bce + 0.5 * dice
        return seg_losses.loss_value(name, logits, masks).float()

    def step(self, images, masks):
        """-> (loss float, metrics dict) after one optimiser step on the batch."""
        new_stats = {}
        logits = uf.unet_forward(self.sd, images, backbone=self.backbone, use_attention=self.attn,
                                 training=True, new_stats=new_stats)
        loss = self._loss(logits, masks)
        self.opt.zero_grad()
        loss.backward()
        self.opt.step()
        with torch.no_grad():
            for k, v in new_stats.items():
                self.sd[k].copy_(v)
            metrics = seg_losses.confusion_metrics(torch.sigmoid(logits.detach()), masks)
        return float(loss.detach()), metrics

    def evaluate_batch(self, images, masks):
        """Validation batch (train.py:87-99): eval-mode forward (running statistics), loss,
        metrics on sigmoid(outputs); no parameter or buffer changes."""
        with torch.no_grad():
            logits = uf.unet_forward(self.sd, images, backbone=self.backbone, use_attention=self.attn,
                                     training=False)
            loss = self._loss(logits, masks)
            metrics = seg_losses.confusion_metrics(torch.sigmoid(logits), masks)
        return float(loss), metrics


def epoch_means(batch_results):
    """The epoch dict train_epoch / evaluate return (train.py:54-68, :100-112): every per-batch
    metric and the per-batch loss weighted by the batch size, divided by the sample count.
    `batch_results` = [(loss float, metrics dict, batch_size)]."""
    total = sum(bs for _, _, bs in batch_results)
    out = {}
    for loss, metrics, bs in batch_results:
        for k, v in metrics.items():
            out[k] = out.get(k, 0.0) + v * bs
        out["loss"] = out.get("loss", 0.0) + loss * bs
    return {k: v / total for k, v in out.items()}
