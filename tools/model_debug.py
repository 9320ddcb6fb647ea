"""Layer-by-layer error report of the B200 U-Net plan against the fp32 CPU oracle (debug aid;
run on the GPU box).  usage: model_debug.py [backbone] [attn 0/1] [size] [batch]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "image-segmentation-project_b200")); sys.path.insert(0, ROOT)
import torch
from oracle import unet_functional as uf, state_layout, seg_losses
import advanced_models, losses

backbone = sys.argv[1] if len(sys.argv) > 1 else "resnet34"
attn = bool(int(sys.argv[2])) if len(sys.argv) > 2 else True
size = int(sys.argv[3]) if len(sys.argv) > 3 else 64
batch = int(sys.argv[4]) if len(sys.argv) > 4 else 2
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = advanced_models.UNetWithBackbone(1, backbone, pretrained=False, use_attention=attn)
sd = uf.synth_state(model.state_dict(), seed=3)
model.load_state_dict(sd)
model = model.to(dev).train()
g = torch.Generator().manual_seed(11)
x = torch.rand(batch, 1, size, size, generator=g)
t = (torch.rand(batch, 1, size, size, generator=g) > 0.8).float()

def rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp(min=1e-12))

taps = {}
new_stats = {}
ref_logits = uf.unet_forward(sd, x, backbone=backbone, use_attention=attn, training=True, new_stats=new_stats, taps=taps)
crit_ref = lambda lo, ta: seg_losses.torch_loss("combo", lo, ta) if hasattr(seg_losses, "torch_loss") else None
logits = model(x.to(dev))
torch.cuda.synchronize()
plan = next(iter(model._plans.values()))
print(f"plan: {len(plan.fwd_ops)} fwd ops / {plan.launches_fwd} launches, {len(plan.bwd_ops)} bwd ops / {plan.launches_bwd} launches, {plan.nbytes/2**20:.1f} MiB")
for name, act in plan.named.items():
    got = act.v.torch().float().cpu().permute(0, 3, 1, 2)
    print(f"  {name:4s} rel err {rel(got, taps[name]):.3e}  max|ref| {float(taps[name].abs().max()):.3f}")
print(f"  logits rel err {rel(logits.cpu(), ref_logits):.3e}")
for k in ("bn1.running_mean", "bn1.running_var", "enc4.2.bn2.running_var"):
    print(f"  {k} rel err {rel(model.state_dict()[k].cpu(), new_stats[k]):.3e}")
# backward with the combo loss on both sides
crit = losses.get_loss_function({"loss_fn": "combo"})
loss = crit(logits, t.to(dev))
loss.backward()
torch.cuda.synchronize()
import torch.nn.functional as F
def combo(lo, ta):
    p = torch.sigmoid(lo).flatten(); tt = ta.flatten()
    dice = 1 - (2 * (p * tt).sum() + 1) / (p.sum() + tt.sum() + 1)
    return 0.5 * F.binary_cross_entropy_with_logits(lo, ta) + 0.5 * dice
_, ref_loss, ref_grads = uf.forward_backward(sd, x, t, combo, backbone=backbone, use_attention=attn, training=True)
print(f"loss {float(loss):.6f} ref {ref_loss:.6f}")
worst = []
for n, p in model.named_parameters():
    rg = ref_grads[n]
    if rg is None:
        continue
    e = rel(p.grad.cpu(), rg)
    worst.append((e, n, float(rg.norm())))
worst.sort(reverse=True)
for e, n, nr in worst[:25]:
    print(f"  grad {n:40s} rel err {e:.3e}  |ref| {nr:.3e}")
print("median grad rel err", sorted(w[0] for w in worst)[len(worst) // 2])
