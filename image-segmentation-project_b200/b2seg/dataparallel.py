"""Data-parallel training of one UNetWithBackbone over the GPUs of a box (north_star config 5).

One process per GPU (torch.distributed, NCCL over NVLink 5 / NVSwitch).  The plan writes every
parameter gradient into ONE flat fp32 buffer (unet_plan.UNetPlan.gflat), so the whole gradient
exchange is a single in-place `all_reduce` of ~97 MB (resnet34) -- on NVSwitch the bandwidth
term is ~0.2 ms and launch latency dominates, which is why there is one bucket, not many.
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
    inv = 1.0 / world

    def hook(gflat):
        dist.all_reduce(gflat, op=dist.ReduceOp.SUM, group=group)
        if not global_loss:
            gflat.mul_(inv)

    model._grad_hook = hook
    return model


def detach(model, criterion=None):
    model._grad_hook = None
    if criterion is not None:
        criterion.dp_group = None
    return model
