"""Data-parallel training of one UNetWithBackbone over the GPUs of a box (north_star config 5).

One process per GPU (torch.distributed, NCCL over NVLink 5 / NVSwitch).  The plan writes every
parameter gradient into ONE flat fp32 buffer (unet_plan.UNetPlan.gflat, ~97 MB for resnet34) laid
out in the order the backward pass finishes them: decoder / gates / head, enc4, enc3, enc2,
enc1 + stem.  The backward is replayed as one CUDA graph per such bucket and each bucket's in-place
`all_reduce` is issued (async, on NCCL's stream) right behind its graph, so the exchange of the big
early buckets overlaps the encoder backward and only the last, smallest one (~1 MB) is exposed.
B2SEG_DP_BUCKETS=0 falls back to one all-reduce after the whole backward (the A/B in bench dp_c5).
BatchNorm stays per replica, as in the reference (it has no SyncBN).

The reference has no data-parallel path at all; this is the B200 addition
`north_star` asks for.  Two modes:
  * `attach(model)`: every replica evaluates its loss on its own shard and the gradients are
    AVERAGED -- exact for the shard-decomposable losses (BCE, focal) with equal shards; Dice /
    Tversky terms then normalise per replica.
  * `attach(model, criterion=crit)`: the fused loss all-reduces its eight sums and the pixel count
    (72 bytes) between its reduction and its evaluation (lossfn.loss_forward(group=...)), so the
    loss value, the confusion counts and the per-pixel gradient are those of the GLOBAL batch and
    the replica gradients are SUMMED -- the single-process semantics of every closed-form loss
    (losses.py:21-27, 99-107 normalise by whole-batch totals).
"""
import torch
import torch.distributed as dist


import os


class GradSync:
    """The gradient exchange installed on a model: `bucket(slice)` starts one async all-reduce,
    `finish(gflat)` makes the current stream wait for all of them; calling the object with the whole
    buffer is the unbucketed form."""

    def __init__(self, group, scale):
        self.group, self.scale, self.works = group, scale, []
        self.bucketed = os.environ.get("B2SEG_DP_BUCKETS", "1") != "0"

    def bucket(self, t):
        self.works.append(dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group, async_op=True))

    def finish(self, gflat):
        for w in self.works:
            w.wait()
        self.works.clear()
        if self.scale is not None:
            gflat.mul_(self.scale)

    def __call__(self, gflat):
        self.bucket(gflat)
        self.finish(gflat)


def attach(model, world_size=None, group=None, criterion=None):
    """Install the flat-gradient all-reduce on `model` and make all replicas start identical."""
    world = world_size or dist.get_world_size(group)
    if world <= 1:
        return model
    with torch.no_grad():
        for t in list(model.parameters()) + list(model.buffers()):
            dist.broadcast(t, src=0, group=group)
    global_loss = criterion is not None
    if global_loss:
        if not hasattr(criterion, "dp_group"):
            raise TypeError("criterion must be one of the fused losses of losses.py")
        criterion.dp_group = group if group is not None else dist.group.WORLD
    model._grad_hook = GradSync(group, None if global_loss else 1.0 / world)
    return model


def detach(model, criterion=None):
    model._grad_hook = None
    if criterion is not None:
        criterion.dp_group = None
    return model
