"""Per-case error report for the tcgen05 conv kernels (debug aid; run on the GPU box)."""
import os, sys, traceback
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "image-segmentation-project_b200")); sys.path.insert(0, ROOT)
import torch, torch.nn.functional as F
from b2seg import convops as co

dev = torch.device("cuda:0")
which = sys.argv[1] if len(sys.argv) > 1 else "all"
rb = lambda t: t.to(torch.bfloat16).float()
nhwc = lambda t: t.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)
nchw = lambda t: t.float().cpu().permute(0, 3, 1, 2).contiguous()

def report(tag, got, ref):
    err = (got - ref).abs()
    rel = float(err.max() / ref.abs().max().clamp(min=1e-6))
    nbad = int((err > 2.0 ** -7 * ref.abs() + 2e-3 * ref.abs().max()).sum())
    print(f"  {tag:8s} maxerr {float(err.max()):.4e} refmax {float(ref.abs().max()):.3e} rel {rel:.3e} bad {nbad}/{err.numel()}", flush=True)
    return nbad

CASES = [(2, 16, 16, 64, 64, 3, 1), (1, 8, 16, 64, 32, 1, 1), (4, 8, 8, 128, 256, 3, 1), (1, 32, 48, 96, 32, 3, 1),
         (2, 12, 20, 32, 32, 3, 1), (3, 3, 3, 512, 512, 3, 1), (2, 16, 16, 64, 128, 3, 2), (2, 32, 32, 64, 128, 1, 2),
         (8, 64, 64, 64, 64, 3, 1), (4, 32, 32, 128, 128, 3, 1)]
for (n, h, w, cin, cout, k, stride) in CASES:
    print(f"case n{n} {h}x{w} cin{cin} cout{cout} k{k} s{stride}", flush=True)
    g = torch.Generator().manual_seed(1)
    x = rb(torch.randn(n, cin, h, w, generator=g)); wt = rb(torch.randn(cout, cin, k, k, generator=g) / (cin*k*k) ** 0.5)
    bias = torch.randn(cout, generator=g); pad = k // 2; ho, wo = h // stride, w // stride
    xd = nhwc(x).to(dev)
    dy = rb(torch.randn(n, cout, ho, wo, generator=g)); dyd = nhwc(dy).to(dev)
    try:
        if which in ("all", "fprop"):
            yd = torch.zeros(n, ho, wo, cout, dtype=torch.bfloat16, device=dev)
            stats = torch.zeros(co.stat_rows(n, ho, wo), 2, cout, device=dev)
            co.conv_fprop(co.View(xd), co.pack_conv_fprop(wt.to(dev)), k, stride, co.View(yd), bias=bias.to(dev), stats=stats).run()
            torch.cuda.synchronize()
            y_ref = F.conv2d(x, wt, bias, stride, pad)
            report("fprop", nchw(yd), y_ref)
            s = stats.double().sum(0).cpu()
            print("  stats sum err", float((s[0] - y_ref.double().sum((0,2,3))).abs().max()), "sumsq relerr",
                  float(((s[1] - (y_ref.double()**2).sum((0,2,3))).abs() / (y_ref.double()**2).sum((0,2,3))).max()), flush=True)
        if which in ("all", "dgrad"):
            dxd = torch.zeros(n, h, w, cin, dtype=torch.bfloat16, device=dev)
            co.conv_dgrad(co.View(dyd), co.pack_conv_dgrad(wt.to(dev)), k, stride, co.View(dxd)).run()
            torch.cuda.synchronize()
            report("dgrad", nchw(dxd), F.conv_transpose2d(dy, wt, None, stride, pad, output_padding=stride - 1))
        if which in ("all", "wgrad"):
            gw = torch.zeros(cout, cin, k, k, device=dev)
            ws = torch.empty(co.wgrad_workspace_bytes(cin, cout, k*k, n, ho, wo) // 4 + 4, device=dev)
            co.conv_wgrad(co.View(xd), co.View(dyd), k, stride, gw, ws).run()
            torch.cuda.synchronize()
            report("wgrad", gw.cpu(), torch.nn.grad.conv2d_weight(x, wt.shape, dy, stride, pad))
    except Exception:
        traceback.print_exc(); print("ABORT", flush=True); sys.exit(1)
print("conv_debug finished")
