"""Teacher-forced layer parity of the B200 U-Net plan: the fp32 oracle is fed the plan's own
stored activations at every storage point, so each layer's forward output and every parameter
gradient can be compared tightly (debug aid; run on the GPU box).
usage: teacher_debug.py [backbone] [attn 0/1] [size] [batch] [train 0/1] [loss]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "image-segmentation-project_b200")); sys.path.insert(0, ROOT)
import torch
from oracle import unet_functional as uf, seg_losses
from b2seg import synth
import advanced_models, losses

backbone = sys.argv[1] if len(sys.argv) > 1 else "resnet34"
attn = bool(int(sys.argv[2])) if len(sys.argv) > 2 else True
size = int(sys.argv[3]) if len(sys.argv) > 3 else 64
batch = int(sys.argv[4]) if len(sys.argv) > 4 else 2
train = bool(int(sys.argv[5])) if len(sys.argv) > 5 else True
loss_name = sys.argv[6] if len(sys.argv) > 6 else "combo"
dev = torch.device("cuda:0")
torch.manual_seed(5)
model = advanced_models.UNetWithBackbone(1, backbone, pretrained=False, use_attention=attn)
sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
model = model.to(dev)
model.train(train)
x, t = synth.cell_tiles(batch, size, seed=77)


def rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp(min=1e-30))


def cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a @ b) / (a.norm() * b.norm()).clamp(min=1e-30))


logits = model(x.to(dev))
crit = losses.get_loss_function({"loss_fn": loss_name})
loss = crit(logits, t.to(dev))
loss.backward()
torch.cuda.synchronize()
plan = next(iter(model._plans.values()))
ours = plan.stored_nchw()
computed = {}
loss_fn = lambda lg, tt: seg_losses.loss_value(loss_name, lg, tt).float()
act_grads = {}
ref_logits, ref_loss, ref_grads = uf.forward_backward(sd, x, t, loss_fn, backbone=backbone, use_attention=attn,
                                                      training=train, store=uf.teacher_store(ours, computed, grads=act_grads))
ours_g = plan.stored_nchw(grads=True)
_, _, twin_grads = uf.forward_backward(sd, x, t, loss_fn, backbone=backbone, use_attention=attn, training=train,
                                       store=uf.teacher_store(ours, None, grad_round=uf.bf16_round))
print(f"{backbone} attn={attn} {size}x{size} b{batch} train={train}: loss {float(loss):.6f} teacher-forced ref {ref_loss:.6f}")
worst = 0.0
for name, v in ours.items():
    e = rel(v, computed[name])
    worst = max(worst, e)
    if e > 6e-3:
        print(f"  fwd {name:24s} rel {e:.3e}  <<<<")
print(f"forward: {len(ours)} storage points, worst rel err {worst:.3e}")
print("activation gradients (backward order):")
for name in reversed(list(ours.keys())):
    if name in ours_g and name in act_grads:
        e = rel(ours_g[name], act_grads[name])
        if e > 8e-3 or os.environ.get("VERBOSE"):
            print(f"  dL/d {name:22s} rel {e:.3e} cos {cos(ours_g[name], act_grads[name]):+.5f}")
gmax = max(float(v.norm()) for v in ref_grads.values() if v is not None)
bad = 0
params = dict(model.named_parameters())
for n in reversed(list(params)):
    rg = ref_grads[n]
    if rg is None:
        continue
    pg = params[n].grad.cpu()
    tiny = float(rg.norm()) < 1e-5 * gmax
    c, e = cos(pg, rg), rel(pg, rg)
    et = rel(twin_grads[n], rg)
    flag = not tiny and (c < 0.995 or e > max(2.5 * et, 1e-2))
    bad += flag
    if flag or os.environ.get("VERBOSE"):
        print(f"  grad {n:40s} cos {c:+.5f} rel {e:.2e} twin-rel {et:.2e} |ref| {float(rg.norm()):.2e} |got| {float(pg.norm()):.2e}{'  <<<<' if flag else ''}")
errs = sorted(rel(params[n].grad.cpu(), ref_grads[n]) for n in params if ref_grads[n] is not None and float(ref_grads[n].norm()) >= 1e-5 * gmax)
terrs = sorted(rel(twin_grads[n], ref_grads[n]) for n in params if ref_grads[n] is not None and float(ref_grads[n].norm()) >= 1e-5 * gmax)
print(f"bf16-gradient twin of the oracle: median rel {terrs[len(terrs)//2]:.3e}, max {terrs[-1]:.3e}")
print(f"backward: {len(errs)} gradients, median rel {errs[len(errs)//2]:.3e}, max {errs[-1]:.3e}, flagged {bad}")
# free-running comparison for context
with torch.no_grad():
    free = uf.unet_forward(sd, x, backbone=backbone, use_attention=attn, training=train)
print(f"free-running fp32 oracle: logits rel {rel(logits.detach().cpu(), free):.3e}")

# isolate BatchNorm backward of decoder{k}.4 from upstream error: recompute dyb in fp64 from OUR b.g
if train:
    for k in (4, 3, 2, 1):
        bg, b, yb = ours_g[f"dec{k}.b"].double(), ours[f"dec{k}.b"].double(), ours[f"dec{k}.yb"].double()
        gamma = sd[f"decoder{k}.4.weight"].double().view(1, -1, 1, 1)
        dz = bg * (b > 0)
        mean = yb.mean((0, 2, 3), keepdim=True)
        var = yb.var((0, 2, 3), unbiased=False, keepdim=True)
        yh = (yb - mean) / (var + 1e-5).sqrt()
        dy = gamma / (var + 1e-5).sqrt() * (dz - dz.mean((0, 2, 3), keepdim=True) - yh * (dz * yh).mean((0, 2, 3), keepdim=True))
        print(f"decoder{k}.4 BN backward recomputed from our own inputs: rel {rel(ours_g[f'dec{k}.yb'], dy):.3e}; "
              f"|dz| {float(dz.norm()):.3e} |dy/c1| {float((dy / (gamma / (var + 1e-5).sqrt())).norm()):.3e}")
        og = act_grads[f"dec{k}.b"].double()
        print(f"    oracle: |dL/db| {float(og.norm()):.3e}  ours-vs-oracle dL/db rel {rel(bg, og):.3e}; per-channel-mean part "
              f"{float(og.mean((0,2,3)).norm() * (og[0,0].numel()*og.shape[0])**0.5):.3e}")
    for k in (4, 3):
        m_ours = ours[f"dec{k}.b"] > 0
        m_orc = computed[f"dec{k}.b"] > 0
        og = act_grads[f"dec{k}.b"].double()
        mism = (m_ours != m_orc)
        print(f"decoder{k}.4 ReLU mask mismatches: {int(mism.sum())} of {mism.numel()}; gradient energy at them "
              f"{float((og[mism] ** 2).sum() / (og ** 2).sum()):.3e}; max |oracle pre-act| there "
              f"{float(computed[f'dec{k}.b'][mism].abs().max()) if mism.any() else 0:.3e}, max |ours| {float(ours[f'dec{k}.b'][mism].abs().max()) if mism.any() else 0:.3e}")
