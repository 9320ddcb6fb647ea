"""Per-op event timings of one planned U-Net step (debug / profiling aid; run on the GPU box).
usage: op_times.py [batch] [size] [backbone] [attn 0/1]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "image-segmentation-project_b200")); sys.path.insert(0, ROOT)
import torch
import advanced_models, losses
from b2seg import synth

batch = int(sys.argv[1]) if len(sys.argv) > 1 else 64
size = int(sys.argv[2]) if len(sys.argv) > 2 else 256
backbone = sys.argv[3] if len(sys.argv) > 3 else "resnet34"
attn = bool(int(sys.argv[4])) if len(sys.argv) > 4 else True
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = advanced_models.UNetWithBackbone(1, backbone, pretrained=False, use_attention=attn).to(dev).train()
crit = losses.get_loss_function({"loss_fn": "combo"})
imgs, masks = synth.cell_tiles(min(batch, 16), size, seed=1)
idx = torch.arange(batch) % imgs.shape[0]
imgs, masks = imgs[idx].to(dev), masks[idx].to(dev)
out = model(imgs); loss = crit(out, masks); loss.backward()
plan = model._plan_for(imgs, True)
torch.cuda.synchronize()
NAMES = {1: "bn_finalize", 2: "bn_apply", 3: "bn_bwd_reduce", 4: "bn_bwd_finalize", 5: "bn_bwd_apply",
         6: "maxpool_fwd", 7: "maxpool_bwd", 8: "stem_fwd", 9: "stem_wgrad", 10: "dot1_fwd", 11: "dot1_bwd",
         12: "colsum_finalize", 13: "colsum", 14: "pack", 15: "gate_apply", 16: "gate_bwd", 17: "se_pool",
         18: "se_mlp", 19: "se_scale", 20: "se_bwd_reduce", 21: "se_mlp_bwd", 22: "se_bwd_apply", 23: "fill"}
for tag, seq in (("fwd", plan.fwd), ("bwd", plan.bwd)):
    acc = [0.0] * len(seq.ops)
    R = 5
    seq.run_timed()
    for _ in range(R):
        for i, t in enumerate(seq.run_timed()):
            acc[i] += t / R
    print(f"== {tag}: {len(seq.ops)} ops, {sum(acc):.3f} ms")
    for i, (op, t) in enumerate(zip(seq.ops, acc)):
        kind = op.kind
        if kind.startswith("ew"):
            kind = NAMES.get(int(kind[2:]), kind)
        extra = f" {op.flops / t / 1e9:7.1f} TF/s" if op.flops else ""
        print(f"{tag}{i:4d} {kind:16s} {t * 1e3:9.1f} us{extra}  {op.desc}")
