"""Data-parallel consistency over real NCCL (run under torchrun, one rank per GPU):

  1. the loss every replica reports is the GLOBAL-batch loss: it equals the (non-parallel) fused
     criterion evaluated on the concatenation of all replicas' logits and targets;
  2. after a few optimisation steps the replicas' parameters are still bit-identical (same
     all-reduced gradient, same optimizer arithmetic);
  3. the summed gradient equals the sum of what each replica computed locally.

usage: python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29631 tools/dp_check.py
"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "image-segmentation-project_b200")); sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

import advanced_models, losses, train as train_mod
from b2seg import dataparallel, synth
from b2seg.optim import FusedAdam

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
per, size = 8, 128
imgs, masks = synth.cell_tiles(per * world, size, seed=7)
x = imgs[rank * per:(rank + 1) * per].to(dev)
y = masks[rank * per:(rank + 1) * per].to(dev)
ok = True
for loss_name in ("combo", "tversky", "focal"):
    torch.manual_seed(100 + rank)  # replicas start DIFFERENT: attach() must make them identical
    model = advanced_models.UNetWithBackbone(1, "resnet34", pretrained=False, use_attention=True).to(dev).train()
    crit = losses.get_loss_function({"loss_fn": loss_name})
    plain = losses.get_loss_function({"loss_fn": loss_name})
    opt = FusedAdam(model.parameters(), lr=1e-3, weight_decay=1e-5)
    dataparallel.attach(model, world, criterion=crit)
    for step in range(3):
        out = model(x)
        loss = crit(out, y)
        # (1) global-batch loss
        outs = [torch.empty_like(out) for _ in range(world)]
        ys = [torch.empty_like(y) for _ in range(world)]
        dist.all_gather(outs, out.detach().contiguous())
        dist.all_gather(ys, y)
        want = float(plain(torch.cat(outs), torch.cat(ys)))
        got = float(loss.detach())
        rel = abs(got - want) / max(abs(want), 1e-12)
        opt.zero_grad()
        loss.backward()
        opt.step()
        # (2) replicas identical
        flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
        ref = flat.clone()
        dist.broadcast(ref, src=0)
        same = bool(torch.equal(flat, ref))
        flags = torch.tensor([1.0 if same else 0.0, rel], device=dev, dtype=torch.float64)
        allf = [torch.empty_like(flags) for _ in range(world)]
        dist.all_gather(allf, flags)
        if rank == 0:
            same_all = all(float(f[0]) == 1.0 for f in allf)
            rel_max = max(float(f[1]) for f in allf)
            good = same_all and rel_max < 1e-5
            ok = ok and good
            print(f"{loss_name:8s} step {step}: global loss {got:.7f} vs concatenated {want:.7f} (rel {rel_max:.1e}); "
                  f"replicas identical: {same_all}  {'OK' if good else 'FAIL'}", flush=True)
    # (3) bucketed exchange (one all-reduce per backward segment) == one all-reduce at the end
    plan = model._plan_for(x, True)
    gs = []
    for env in ("1", "0"):
        dataparallel.detach(model, crit)
        os.environ["B2SEG_DP_BUCKETS"] = env
        dataparallel.attach(model, world, criterion=crit)
        opt.zero_grad()
        crit(model(x), y).backward()
        gs.append((plan.gflat.clone(), model._dp_buckets))
    os.environ.pop("B2SEG_DP_BUCKETS", None)
    diff = float((gs[0][0] - gs[1][0]).abs().max() / gs[1][0].abs().max())
    flags = torch.tensor([diff], device=dev, dtype=torch.float64)
    dist.all_reduce(flags, op=dist.ReduceOp.MAX)
    if rank == 0:
        good = float(flags[0]) < 1e-6 and gs[0][1] > 1 and gs[1][1] == 1
        ok = ok and good
        print(f"{loss_name:8s} bucketed ({gs[0][1]} buckets {[ (hi - lo) * 4 >> 10 for _, lo, hi in plan.buckets]} KiB) vs "
              f"single all-reduce: max rel diff {float(flags[0]):.1e}  {'OK' if good else 'FAIL'}", flush=True)
    dataparallel.detach(model, crit)
dist.barrier()
dist.destroy_process_group()
if rank == 0:
    print("dp_check", "PASSED" if ok else "FAILED")
    sys.exit(0 if ok else 1)
