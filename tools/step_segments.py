"""Where one training step's time goes outside / inside the planned graphs (run on the GPU box).
Each segment is event-timed over `iters` repetitions in isolation (device time; launch-bound
segments show their CPU-bound enqueue time).  usage: step_segments.py [batch] [size]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "image-segmentation-project_b200")); sys.path.insert(0, ROOT)
import torch
import advanced_models, losses, train as train_mod
from b2seg import synth

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
iters = 20


def timed(fn):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


out = model(imgs)
seg = {}
seg["full train_step"] = timed(lambda: train_mod.train_step(model, imgs, masks, opt, crit))
seg["fwd graph (plan.fwd.run)"] = timed(plan.fwd.run)
seg["bwd graph (plan.bwd.run)"] = timed(plan.bwd.run)
seg["model(imgs) incl. autograd glue"] = timed(lambda: model(imgs))
seg["loss fwd (criterion)"] = timed(lambda: crit(out.detach(), masks))


def fwd_bwd():
    o = model(imgs); l = crit(o, masks); opt.zero_grad(); l.backward()


seg["fwd + loss + backward (no optimizer)"] = timed(fwd_bwd)
seg["optimizer.step (torch Adam)"] = timed(opt.step)
seg["optimizer.zero_grad"] = timed(opt.zero_grad)
for k, v in seg.items():
    print(f"{k:42s} {v:8.3f} ms")
print("graph?", plan.fwd.handle is not None, "launches", plan.launches_fwd, plan.launches_bwd)
