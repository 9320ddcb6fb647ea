#!/usr/bin/env python
"""Achieved HBM GB/s of the memory-bound kernels: loss forward (8 B/px), loss
backward (12 B/px), calculate_metrics on probabilities (8 B/px), augment permute / warp (2 B/px
per plane pair -> 4 B/px image+mask).  CUDA events on the launching stream, inputs larger than
L2 (64 x 1024 x 1024 fp32 = 268 MB per tensor), rotating over 2 buffer sets.

    python tools/hbm_bench.py [--n 64 --size 1024 --iters 20] > profiles/rNN_hbm_kernels.json
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "image-segmentation-project_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402


def timed(fn, iters, warm=3):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=64)
    ap.add_argument("--size", type=int, default=1024)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--only", default="", help="'aug': skip the loss / metric kernels")
    a = ap.parse_args()
    from b2seg import lossfn, augment
    import losses
    dev = torch.device("cuda:0")
    peak = 6650.0
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peak = json.load(open(pk)).get("hbm_gbs", peak)
    g = torch.Generator(device="cpu").manual_seed(1234)
    shape = (a.n, 1, a.size, a.size)
    npx = a.n * a.size * a.size
    xs = [(torch.randn(shape, generator=g) * 2).to(dev) for _ in range(2)]
    ts = [(torch.rand(shape, generator=g) > 0.9).float().to(dev) for _ in range(2)]
    out = {"shape": list(shape), "pixels": npx, "hbm_peak_gbs": peak, "kernels": {}}

    def rec(name, ms, bytes_px, px=npx):
        gbs = bytes_px * px / (ms * 1e-3) / 1e9
        out["kernels"][name] = {"us": round(ms * 1e3, 2), "bytes_per_px": bytes_px,
                                "gbs": round(gbs, 1), "frac_of_peak": round(gbs / peak, 4)}

    if a.only != "aug":
        for name in ("combo", "dice", "bce", "focal", "tversky"):
            crit = losses.get_loss_function({"loss_fn": name})
            spec = crit._spec()
            saved = {}

            def fwd(i, spec=spec):
                saved["r"] = lossfn.loss_forward(xs[i & 1], ts[i & 1], spec)

            rec(f"seg_loss_fwd[{name}]", timed(fwd, a.iters), 8)
            _, _, coef, sv = saved["r"]

            def bwd(i, coef=coef, sv=sv):
                lossfn.loss_backward(sv, coef, None, shape)

            rec(f"seg_loss_bwd[{name}]", timed(bwd, a.iters), 12)

        ps = [torch.sigmoid(x) for x in xs]
        rec("seg_metrics_from_probs", timed(lambda i: lossfn.metrics_from_probs(ps[i & 1], ts[i & 1]), a.iters), 8)
        del ps

    # augmentation: image + mask planes, uint8
    planes = [torch.randint(0, 256, (2 * a.n, a.size, a.size), generator=g, dtype=torch.uint8).to(dev)
              for _ in range(2)]
    ppx = a.n * a.size * a.size  # pixels per (image, mask) pair -> 4 B/px
    rec("aug_hflip", timed(lambda i: augment.hflip(planes[i & 1]), a.iters), 4, ppx)
    rec("aug_vflip", timed(lambda i: augment.vflip(planes[i & 1]), a.iters), 4, ppx)
    rec("aug_rot90", timed(lambda i: augment.rot90(planes[i & 1], 1), a.iters), 4, ppx)
    rec("aug_transpose", timed(lambda i: augment.transpose(planes[i & 1]), a.iters), 4, ppx)
    m = augment.affine_matrix(a.size, a.size, 1.03, (0.02, -0.03), 11.0, (3.0, -2.0))
    rec("aug_warp_affine_linear_reflect",
        timed(lambda i: augment.warp_affine(planes[i & 1], m, augment.LINEAR, augment.REFLECT), a.iters), 4, ppx)
    rec("aug_warp_affine_nearest_const",
        timed(lambda i: augment.warp_affine(planes[i & 1], m, augment.NEAREST, augment.CONSTANT), a.iters), 4, ppx)
    yy, xx = torch.meshgrid(torch.arange(a.size, dtype=torch.float32), torch.arange(a.size, dtype=torch.float32),
                            indexing="ij")
    mx = (xx + 3.0 * torch.sin(yy / 17.0)).to(dev)
    my = (yy + 3.0 * torch.cos(xx / 13.0)).to(dev)
    rec("aug_remap_linear_reflect",
        timed(lambda i: augment.remap(planes[i & 1], mx, my), a.iters), 4, ppx)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
