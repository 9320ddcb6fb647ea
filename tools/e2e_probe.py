"""A/B of the end-to-end loop variants (run on the GPU box): where do host batches cost time?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "image-segmentation-project_b200")); sys.path.insert(0, ROOT)
import torch
import advanced_models, losses, train as T
from b2seg import synth
from b2seg.optim import FusedAdam

dev = torch.device("cuda:0")
B, S, POOL, STEPS = 64, 256, 4, 30
torch.manual_seed(0)
model = advanced_models.UNetWithBackbone(1, "resnet34", pretrained=False, use_attention=True).to(dev).train()
crit = losses.get_loss_function({"loss_fn": "combo"})
opt = FusedAdam(model.parameters(), lr=1e-3, weight_decay=1e-5)
bi, bm = synth.cell_tiles(64, S, seed=1)
idx = torch.arange(B * POOL) % 64
hi = bi[idx].view(POOL, B, 1, S, S).contiguous().pin_memory()
hm = bm[idx].view(POOL, B, 1, S, S).contiguous().pin_memory()
di, dm = hi.to(dev), hm.to(dev)
rl = torch.empty(8, dtype=torch.float32).pin_memory()
rs = torch.empty(8, 16, dtype=torch.float64).pin_memory()


def gen(n):
    for i in range(n):
        yield hi[i % POOL], hm[i % POOL]


def resident(n):
    for i in range(n):
        T.train_step(model, di[i % POOL], dm[i % POOL], opt, crit)


def direct_sync(n):
    for i in range(n):
        l, s = T.train_step(model, hi[i % POOL].to(dev, non_blocking=True), hm[i % POOL].to(dev, non_blocking=True), opt, crit)
        l.cpu(); s.cpu()


def direct_async(n):
    for i in range(n):
        l, s = T.train_step(model, hi[i % POOL].to(dev, non_blocking=True), hm[i % POOL].to(dev, non_blocking=True), opt, crit)
        rl[i % 8].copy_(l, non_blocking=True); rs[i % 8].copy_(s, non_blocking=True)


def prefetch_async(n):
    for i, (im, mk) in enumerate(T.DevicePrefetcher(gen(n), dev)):
        l, s = T.train_step(model, im, mk, opt, crit)
        rl[i % 8].copy_(l, non_blocking=True); rs[i % 8].copy_(s, non_blocking=True)


def prefetch_sync(n):
    for i, (im, mk) in enumerate(T.DevicePrefetcher(gen(n), dev)):
        l, s = T.train_step(model, im, mk, opt, crit)
        l.cpu(); s.cpu()


for name, fn in (("resident", resident), ("direct_sync", direct_sync), ("direct_async", direct_async),
                 ("prefetch_async", prefetch_async), ("prefetch_sync", prefetch_sync), ("resident", resident)):
    fn(4); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(STEPS); torch.cuda.synchronize(); e1.record(); torch.cuda.synchronize()
    print(f"{name:16s} {e0.elapsed_time(e1) / STEPS:8.3f} ms/step")
