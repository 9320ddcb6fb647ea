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
taps_q = {}
with torch.no_grad():
    ref_logits_q = uf.unet_forward(sd, x, backbone=backbone, use_attention=attn, training=True, taps=taps_q, storage_round=uf.bf16_round)
crit_ref = lambda lo, ta: seg_losses.torch_loss("combo", lo, ta) if hasattr(seg_losses, "torch_loss") else None
logits = model(x.to(dev))
torch.cuda.synchronize()
plan = next(iter(model._plans.values()))
print(f"plan: {len(plan.fwd_ops)} fwd ops / {plan.launches_fwd} launches, {len(plan.bwd_ops)} bwd ops / {plan.launches_bwd} launches, {plan.nbytes/2**20:.1f} MiB")
for name, act in plan.named.items():
    got = act.v.torch().float().cpu().permute(0, 3, 1, 2)
    print(f"  {name:4s} rel err vs fp32 {rel(got, taps[name]):.3e}  vs bf16-storage oracle {rel(got, taps_q[name]):.3e}  (oracle bf16 vs fp32 {rel(taps_q[name], taps[name]):.3e})")
print(f"  logits rel err vs fp32 {rel(logits.cpu(), ref_logits):.3e} vs bf16-storage {rel(logits.cpu(), ref_logits_q):.3e} (oracle bf16 vs fp32 {rel(ref_logits_q, ref_logits):.3e})")
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
_, ref_loss_q, ref_grads_q = uf.forward_backward(sd, x, t, combo, backbone=backbone, use_attention=attn, training=True, storage_round=uf.bf16_round)
print(f"loss {float(loss):.6f} ref {ref_loss:.6f}")
worst = []
for n, p in model.named_parameters():
    rg = ref_grads[n]
    if rg is None:
        continue
    e = rel(p.grad.cpu(), rg)
    worst.append((e, n, float(rg.norm()), rel(p.grad.cpu(), ref_grads_q[n]), rel(ref_grads_q[n], rg)))
worst.sort(reverse=True)
for e, n, nr, eq, qq in worst[:25]:
    print(f"  grad {n:36s} vs fp32 {e:.3e} vs bf16-storage {eq:.3e} (oracle bf16 vs fp32 {qq:.3e}) |ref| {nr:.3e}")
print("median vs bf16-storage", sorted(w[3] for w in worst)[len(worst) // 2], "max", max(w[3] for w in worst))
print("median grad rel err", sorted(w[0] for w in worst)[len(worst) // 2])
