"""In-graph cost of each op kind: rebuild the forward / backward sequences of the plan WITHOUT one
kind of op, capture and replay them, and report the time the graph loses.  Unlike tools/op_times.py
(every op isolated between two events) this shows what a kind costs on the dependent chain with
programmatic dependent launch and the side branch active.  Timing only: a sequence with ops
removed computes garbage.  usage: graph_costs.py [batch] [size]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "image-segmentation-project_b200")); sys.path.insert(0, ROOT)
import torch
import advanced_models, losses, train as train_mod
from b2seg import synth, convops as co, _lib as L

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 64
size = int(sys.argv[2]) if len(sys.argv) > 2 else 256
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = advanced_models.UNetWithBackbone(1, "resnet34", pretrained=False, use_attention=True).to(dev).train()
crit = losses.get_loss_function({"loss_fn": "combo"})
opt = torch.optim.Adam(model.parameters(), lr=1e-3, weight_decay=1e-5)
imgs, masks = synth.cell_tiles(min(batch, 16), size, seed=1)
idx = torch.arange(batch) % imgs.shape[0]
imgs, masks = imgs[idx].to(dev), masks[idx].to(dev)
for _ in range(3):
    train_mod.train_step(model, imgs, masks, opt, crit)
plan = model._plan_for(imgs, True)
EW = {v: k[3:].lower() for k, v in vars(L).items() if k.startswith("EW_") and isinstance(v, int) and k != "EW_MAX_PARTIAL_ROWS"}


def label(kind):
    return EW.get(int(kind[2:]), kind) if kind.startswith("ew") and kind[2:].isdigit() else kind


def timed(seq, iters=30):
    for _ in range(3):
        seq.run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        seq.run()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def study(name, ops, side_kinds, join_before):
    base = timed(co.Seq(ops, side_kinds=side_kinds, join_before_kinds=join_before))
    print(f"== {name}: {len(ops)} ops, {base:.3f} ms")
    kinds = {}
    for op in ops:
        kinds.setdefault(op.kind, 0)
        kinds[op.kind] += 1
    rows = []
    for kind, n in kinds.items():
        kept = [op for op in ops if op.kind != kind]
        if not kept:
            continue
        t = timed(co.Seq(kept, side_kinds=side_kinds, join_before_kinds=join_before))
        rows.append((base - t, kind, n))
    for d, kind, n in sorted(rows, reverse=True):
        print(f"   {label(kind):22s} x{n:3d}  {d * 1e3:8.1f} us on the chain  ({d / n * 1e3:6.1f} us each)")
    return base


study("fwd", plan.fwd_ops, (f"ew{L.EW_PACK}",), ("conv",))
study("bwd", plan.bwd_ops, ("wgrad",), ())
# groups
groups = {
    "all BatchNorm bwd (reduce+finalize+apply)": [f"ew{L.EW_BN_BWD_REDUCE}", f"ew{L.EW_BN_BWD_FINALIZE}", f"ew{L.EW_BN_BWD_APPLY}"],
    "all wgrad (side branch)": ["wgrad"],
    "all dgrad3x3": ["dgrad3x3"],
}
base = timed(co.Seq(plan.bwd_ops, side_kinds=("wgrad",)))
for gname, prefixes in groups.items():
    kept = [op for op in plan.bwd_ops if not any(op.kind.startswith(p) for p in prefixes)]
    t = timed(co.Seq(kept, side_kinds=("wgrad",)))
    print(f"bwd without {gname:45s} {t:.3f} ms  (-{(base - t) * 1e3:.0f} us)")
