"""Event-timed micro-benchmark of single conv shapes (fprop+stats / dgrad / wgrad) on the GPU box.
usage: conv_bench.py [case ...]   case = n,h,w,cin,cout[,k[,stride]]   (default: the U-Net's big layers)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "image-segmentation-project_b200")); sys.path.insert(0, ROOT)
import torch
from b2seg import convops as co

dev = torch.device("cuda:0")
cases = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]] or [
    (64, 64, 64, 64, 64), (64, 32, 32, 128, 128), (64, 16, 16, 256, 256), (64, 8, 8, 512, 512),
    (64, 128, 128, 96, 32), (64, 128, 128, 32, 32), (64, 64, 64, 128, 64)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def timed(op, reps=int(os.environ.get('REPS', '10'))):
    op.run(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        if not os.environ.get('NOFLUSH'): flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); op.run(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]

for c in cases:
    n, h, w, cin, cout = c[:5]
    k = c[5] if len(c) > 5 else 3
    stride = c[6] if len(c) > 6 else 1
    ho, wo = h // stride, w // stride
    x = torch.randn(n, h, w, cin, device=dev).to(torch.bfloat16)
    dy = torch.randn(n, ho, wo, cout, device=dev).to(torch.bfloat16)
    wt = torch.randn(cout, cin, k, k, device=dev) / (cin * k * k) ** 0.5
    y = torch.zeros(n, ho, wo, cout, dtype=torch.bfloat16, device=dev)
    dx = torch.zeros(n, h, w, cin, dtype=torch.bfloat16, device=dev)
    stats = torch.zeros(co.stat_rows(n, ho, wo), 2, cout, device=dev)
    gw = torch.zeros(cout, cin, k, k, device=dev)
    ws = torch.empty(co.wgrad_workspace_bytes(cin, cout, k * k, n, ho, wo) // 4 + 4, device=dev)
    ops = {"fprop": co.conv_fprop(co.View(x), co.pack_conv_fprop(wt), k, stride, co.View(y), stats=stats),
           "dgrad": co.conv_dgrad(co.View(dy), co.pack_conv_dgrad(wt), k, stride, co.View(dx)),
           "dgrad+acc": co.conv_dgrad(co.View(dy), co.pack_conv_dgrad(wt), k, stride, co.View(dx), accumulate=True),
           "wgrad": co.conv_wgrad(co.View(x), co.View(dy), k, stride, gw, ws)}
    fl = 2.0 * n * ho * wo * cin * cout * k * k
    print(f"case n{n} {h}x{w} cin{cin} cout{cout} k{k} s{stride}  ({fl / 1e9:.1f} GFLOP)")
    for name, op in ops.items():
        us = timed(op)
        print(f"   {name:10s} {us:8.1f} us  {fl / us / 1e6:7.1f} TF/s")
