#!/usr/bin/env python
"""How far do free-running gradients move when NOTHING but rounding noise changes?  Control for
tests/test_model_gpu.py::test_fixed_point_batchnorm_statistics_match_finalize_kernel: the same model and batch
(a) with BatchNorm statistics through the finalize kernels (B2SEG_BN_FUSED_FWD=0) against the fixed-point
    accumulators (=2, the default), and
(b) finalize kernels against finalize kernels with the input perturbed by 2^-9 relative (one bf16 ulp).
Prints the per-tensor relative gradient differences of both, worst first.

    python tools/bn_noise_check.py [size=192] [batch=4]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "image-segmentation-project_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)
import torch  # noqa: E402


def run(mode, x, t):
    import advanced_models
    import losses
    os.environ["B2SEG_BN_FUSED_FWD"] = mode
    torch.manual_seed(23)
    model = advanced_models.UNetWithBackbone(1, "resnet34", pretrained=False, use_attention=True).cuda().train()
    crit = losses.get_loss_function({"loss_fn": "combo"})
    logits = model(x.cuda())
    crit(logits, t.cuda()).backward()
    torch.cuda.synchronize()
    return logits.detach().float().cpu(), {n: p.grad.detach().double().cpu() for n, p in model.named_parameters()}


def main():
    from b2seg import synth
    size = int(sys.argv[1]) if len(sys.argv) > 1 else 192
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    x, t = synth.cell_tiles(batch, size, seed=17)
    l0, g0 = run("0", x, t)
    l2, g2 = run("2", x, t)
    xp = x * (1.0 + 2.0 ** -9 * torch.sign(torch.randn(x.shape, generator=torch.Generator().manual_seed(1))))
    lp, gp = run("0", xp, t)
    gmax = max(float(v.norm()) for v in g0.values())

    def rel(a, b):
        return {n: float((a[n] - b[n]).norm()) / max(float(b[n].norm()), 1e-2 * gmax) for n in b}

    r2, rp = rel(g2, g0), rel(gp, g0)
    print(f"logits: fixed-point vs finalize {float((l2 - l0).norm() / l0.norm()):.2e}, "
          f"perturbed input vs finalize {float((lp - l0).norm() / l0.norm()):.2e}")
    print(f"{'parameter':40s} {'fixed-point':>12s} {'1-ulp input':>12s}")
    for n in sorted(r2, key=lambda k: -max(r2[k], rp[k]))[:12]:
        print(f"{n:40s} {r2[n]:12.3f} {rp[n]:12.3f}")
    print(f"worst: fixed-point {max(r2.values()):.3f}, 1-ulp input {max(rp.values()):.3f}")


if __name__ == "__main__":
    main()
