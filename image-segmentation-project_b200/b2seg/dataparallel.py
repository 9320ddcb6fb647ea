"""Data-parallel training of one UNetWithBackbone over the GPUs of a box (north_star config 5).

One process per GPU (torch.distributed, NCCL over NVLink 5 / NVSwitch).  The plan writes every
parameter gradient into ONE flat fp32 buffer (unet_plan.UNetPlan.gflat), so the whole gradient
exchange is a single in-place `all_reduce` of ~97 MB (resnet34) -- on NVSwitch the bandwidth
term is ~0.2 ms and launch latency dominates, which is why there is one bucket, not many.
BatchNorm stays per replica, as in the reference (it has no SyncBN).

The reference has no data-parallel path at all; this is the B200 addition
`north_star` asks for.  Gradients are averaged, matching a mean-reduced loss over the global
batch for the shard-decomposable losses (BCE, focal); Dice/Tversky terms normalise per replica.
"""
import torch
import torch.distributed as dist


def attach(model, world_size=None, group=None):
    """Install the flat-gradient all-reduce on `model` and make all replicas start identical."""
    world = world_size or dist.get_world_size(group)
    if world <= 1:
        return model
    with torch.no_grad():
        for t in list(model.parameters()) + list(model.buffers()):
            dist.broadcast(t, src=0, group=group)
    inv = 1.0 / world

    def hook(gflat):
        dist.all_reduce(gflat, op=dist.ReduceOp.SUM, group=group)
        gflat.mul_(inv)

    model._grad_hook = hook
    return model


def detach(model):
    model._grad_hook = None
    return model
