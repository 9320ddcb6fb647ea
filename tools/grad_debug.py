"""Gradient report of the B200 U-Net plan against the fp32 CPU oracle, in network order (debug
aid; run on the GPU box).  usage: grad_debug.py [backbone] [attn 0/1] [size] [batch] [train 0/1]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "image-segmentation-project_b200")); sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F
from oracle import unet_functional as uf
import advanced_models, losses

backbone = sys.argv[1] if len(sys.argv) > 1 else "resnet34"
attn = bool(int(sys.argv[2])) if len(sys.argv) > 2 else True
size = int(sys.argv[3]) if len(sys.argv) > 3 else 64
batch = int(sys.argv[4]) if len(sys.argv) > 4 else 2
train = bool(int(sys.argv[5])) if len(sys.argv) > 5 else True
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = advanced_models.UNetWithBackbone(1, backbone, pretrained=False, use_attention=attn)
sd = uf.synth_state(model.state_dict(), seed=3)
model.load_state_dict(sd)
model = model.to(dev)
model.train(train)
g = torch.Generator().manual_seed(11)
x = torch.rand(batch, 1, size, size, generator=g)
t = (torch.rand(batch, 1, size, size, generator=g) > 0.8).float()


def rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp(min=1e-30))


def cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a @ b) / (a.norm() * b.norm()).clamp(min=1e-30))


def combo(lo, ta):
    p = torch.sigmoid(lo).flatten(); tt = ta.flatten()
    dice = 1 - (2 * (p * tt).sum() + 1) / (p.sum() + tt.sum() + 1)
    return 0.5 * F.binary_cross_entropy_with_logits(lo, ta) + 0.5 * dice


logits = model(x.to(dev))
crit = losses.get_loss_function({"loss_fn": "combo"})
loss = crit(logits, t.to(dev))
loss.backward()
torch.cuda.synchronize()
ref_logits, ref_loss, ref_grads = uf.forward_backward(sd, x, t, combo, backbone=backbone, use_attention=attn, training=train)
_, _, ref_q = uf.forward_backward(sd, x, t, combo, backbone=backbone, use_attention=attn, training=train,
                                  storage_round=uf.bf16_round)
print(f"{backbone} attn={attn} {size}x{size} b{batch} train={train}: loss {float(loss):.6f} ref {ref_loss:.6f}; "
      f"logits rel {rel(logits.detach().cpu(), ref_logits):.3e}")
names = [n for n, _ in model.named_parameters()]
gmax = max(float(v.norm()) for v in ref_grads.values() if v is not None)
bad = 0
for n in reversed(names):
    rg = ref_grads[n]
    if rg is None:
        continue
    pg = dict(model.named_parameters())[n].grad
    if pg is None:
        print(f"  {n:40s} NO GRAD"); bad += 1; continue
    pg = pg.cpu()
    tiny = float(rg.norm()) < 1e-6 * gmax
    c = cos(pg, rg)
    flag = "" if (tiny or c > 0.98) else "  <<<<"
    bad += bool(flag)
    print(f"  {n:40s} cos {c:+.4f} rel {rel(pg, rg):.2e} | oracle-bf16 cos {cos(ref_q[n], rg):+.4f} rel {rel(ref_q[n], rg):.2e} "
          f"|ref| {float(rg.norm()):.2e} |got| {float(pg.norm()):.2e}{' (tiny)' if tiny else ''}{flag}")
print("flagged", bad)
