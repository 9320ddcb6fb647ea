"""tcgen05 convolution kernels (tap-gather GEMM + weight-gradient GEMM) vs a torch-CPU fp32
oracle evaluated on the same bf16-rounded operands.

Tolerance: products of bf16 values are exact in fp32 and both sides accumulate in fp32, so the
only differences are summation order and the final bf16 rounding of the output:
|err| <= 2^-8 * |ref| + 1e-3 * sqrt(K)-scaled absolute slack.
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _nhwc(t):  # NCHW fp32 cpu -> NHWC bf16
    return t.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16)


def _nchw(t):  # NHWC (any) -> NCHW fp32 cpu
    return t.float().cpu().permute(0, 3, 1, 2).contiguous()


def _rb(t):  # round to bf16, keep fp32
    return t.to(torch.bfloat16).float()


def _close(got, ref, k):
    tol = 2.0 ** -7 * ref.abs() + 2e-3 * ref.abs().max().clamp(min=1e-3)
    bad = (got - ref).abs() > tol
    assert not bad.any(), (int(bad.sum()), float((got - ref).abs().max()), float(ref.abs().max()), k)


CASES = [
    # n, h, w, cin, cout, k, stride
    (2, 16, 16, 64, 64, 3, 1),
    (4, 8, 8, 128, 256, 3, 1),
    (1, 32, 48, 96, 32, 3, 1),     # decoder1: cin not a multiple of 64, narrow N
    (2, 12, 20, 32, 32, 3, 1),     # ragged tiles, cin < 64
    (3, 3, 3, 512, 512, 3, 1),     # 96-px input at enc4: tiny odd grid
    (2, 16, 16, 64, 128, 3, 2),
    (2, 32, 32, 64, 128, 1, 2),    # downsample 1x1 s2
    (2, 16, 16, 256, 128, 1, 1),   # attention W_g
    (8, 64, 64, 64, 64, 3, 1),     # many tiles -> BLOCK_N path with >= SM-count CTAs
    (4, 32, 32, 128, 128, 3, 1),
]


@pytest.mark.parametrize("n,h,w,cin,cout,k,stride", CASES)
def test_conv_fprop_dgrad_wgrad(cuda, n, h, w, cin, cout, k, stride):
    from b2seg import convops as co
    g = torch.Generator().manual_seed(n * 1000 + h * 10 + cin + cout + k + stride)
    x = _rb(torch.randn(n, cin, h, w, generator=g))
    wt = _rb(torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5)
    bias = torch.randn(cout, generator=g)
    pad = k // 2
    ho, wo = h // stride, w // stride

    # ---------------- fprop (+bias, + BatchNorm partial statistics) ----------------
    y_ref = F.conv2d(x, wt, bias, stride, pad)
    xd = _nhwc(x).to(cuda)
    yd = torch.zeros(n, ho, wo, cout, dtype=torch.bfloat16, device=cuda)
    rows = co.stat_rows(n, ho, wo)
    stats = torch.zeros(rows, 2, cout, dtype=torch.float32, device=cuda)
    wp = co.pack_conv_fprop(wt.to(cuda))
    op = co.conv_fprop(co.View(xd), wp, k, stride, co.View(yd), bias=bias.to(cuda), stats=stats)
    op.run()
    torch.cuda.synchronize()
    _close(_nchw(yd), y_ref, "fprop")
    s = stats.double().sum(0).cpu()
    np.testing.assert_allclose(s[0].numpy(), y_ref.double().sum((0, 2, 3)).numpy(), rtol=1e-3,
                               atol=1e-2 * float(y_ref.abs().max()) * (n * ho * wo) ** 0.5)
    np.testing.assert_allclose(s[1].numpy(), (y_ref.double() ** 2).sum((0, 2, 3)).numpy(), rtol=2e-3)
    # fixed-point accumulators (stats_fixed): the same per-CTA partial sums, added as integers with atomics into
    # one int64 row -> equal to the fp64 sum of the partial rows up to the 2^-20 / 2^-16 quantum per CTA, and
    # bit-identical from run to run (integer addition commutes)
    acc = torch.zeros(2, cout, dtype=torch.int64, device=cuda)
    opf = co.conv_fprop(co.View(xd), wp, k, stride, co.View(yd), bias=bias.to(cuda), stats=acc)
    opf.run()
    torch.cuda.synchronize()
    first = acc.clone()
    sf = torch.stack([first[0].double() / 2.0 ** 20, first[1].double() / 2.0 ** 16]).cpu()
    assert float((sf[0] - s[0]).abs().max()) <= rows * 2.0 ** -21 + 1e-9, "fixed-point sum"
    assert float((sf[1] - s[1]).abs().max()) <= rows * 2.0 ** -17 + 1e-9, "fixed-point sumsq"
    acc.zero_()
    opf.run()
    torch.cuda.synchronize()
    assert torch.equal(acc, first)

    # ---------------- dgrad ----------------
    dy = _rb(torch.randn(n, cout, ho, wo, generator=g))
    dx_ref = F.conv_transpose2d(dy, wt, None, stride, pad, output_padding=stride - 1)
    dyd = _nhwc(dy).to(cuda)
    dxd = torch.zeros(n, h, w, cin, dtype=torch.bfloat16, device=cuda)
    wd = co.pack_conv_dgrad(wt.to(cuda))
    co.conv_dgrad(co.View(dyd), wd, k, stride, co.View(dxd)).run()
    torch.cuda.synchronize()
    _close(_nchw(dxd), dx_ref, "dgrad")
    # accumulate mode: running it again on top doubles the result
    co.conv_dgrad(co.View(dyd), wd, k, stride, co.View(dxd), accumulate=True).run()
    torch.cuda.synchronize()
    got2 = _nchw(dxd)
    ref2 = _rb(dx_ref) * 2
    assert (got2 - ref2).abs().max() <= 2.0 ** -6 * ref2.abs().max() + 1e-3

    # ---------------- wgrad ----------------
    gw_ref = torch.nn.grad.conv2d_weight(x, wt.shape, dy, stride, pad)
    gw = torch.full((cout, cin, k, k), 7.0, dtype=torch.float32, device=cuda)
    ws = torch.empty(co.wgrad_workspace_bytes(cin, cout, k * k, n, ho, wo) // 4 + 4, dtype=torch.float32,
                     device=cuda)
    co.conv_wgrad(co.View(xd), co.View(dyd), k, stride, gw, ws).run()
    torch.cuda.synchronize()
    scale = float(gw_ref.abs().max())
    np.testing.assert_allclose(gw.cpu().numpy(), gw_ref.numpy(), rtol=2e-3, atol=2e-4 * scale)


@pytest.mark.parametrize("mode", ["0", "2"])  # 0 = generic kernel only, 2 = halo kernel wherever it fits
@pytest.mark.parametrize("n,h,w,cin,cout", [(2, 16, 16, 64, 64), (3, 32, 24, 128, 192), (2, 20, 40, 96, 32),
                                            (1, 64, 64, 256, 128)])
def test_wgrad3x3_halo_and_generic_agree_with_reference(cuda, monkeypatch, mode, n, h, w, cin, cout):
    """Both 3x3 stride-1 weight-gradient kernels (wgrad.cu: per-tap loads vs one halo brick per
    chunk) against torch.nn.grad.conv2d_weight, including accumulate mode."""
    from b2seg import convops as co
    monkeypatch.setenv("B2SEG_WGRAD_HALO", mode)
    g = torch.Generator().manual_seed(n + h + w + cin + cout)
    x = _rb(torch.randn(n, cin, h, w, generator=g))
    dy = _rb(torch.randn(n, cout, h, w, generator=g))
    gw_ref = torch.nn.grad.conv2d_weight(x, (cout, cin, 3, 3), dy, 1, 1)
    xd, dyd = _nhwc(x).to(cuda), _nhwc(dy).to(cuda)
    gw = torch.full((cout, cin, 3, 3), 7.0, dtype=torch.float32, device=cuda)
    ws = torch.empty(co.wgrad_workspace_bytes(cin, cout, 9, n, h, w) // 4 + 4, dtype=torch.float32, device=cuda)
    co.conv_wgrad(co.View(xd), co.View(dyd), 3, 1, gw, ws).run()
    torch.cuda.synchronize()
    scale = float(gw_ref.abs().max())
    np.testing.assert_allclose(gw.cpu().numpy(), gw_ref.numpy(), rtol=2e-3, atol=2e-4 * scale)
    co.conv_wgrad(co.View(xd), co.View(dyd), 3, 1, gw, ws, accumulate=True).run()
    torch.cuda.synchronize()
    np.testing.assert_allclose(gw.cpu().numpy(), 2 * gw_ref.numpy(), rtol=2e-3, atol=4e-4 * scale)


@pytest.mark.parametrize("n,h,w,cin,cout,skip", [(2, 8, 8, 128, 64, 64), (1, 16, 24, 32, 16, 0),
                                                  (4, 4, 4, 512, 256, 256), (2, 64, 64, 64, 32, 64)])
def test_conv_transpose_into_concat_slice(cuda, n, h, w, cin, cout, skip):
    """ConvTranspose2d(k2,s2) written straight into channels [skip, skip+cout) of the concat
    buffer (advanced_models.py:284-333: torch.cat((skip, up), 1) fused away), + its gradients."""
    from b2seg import convops as co
    g = torch.Generator().manual_seed(cin + cout + h)
    x = _rb(torch.randn(n, cin, h, w, generator=g))
    wt = _rb(torch.randn(cin, cout, 2, 2, generator=g) / cin ** 0.5)
    bias = torch.randn(cout, generator=g)
    y_ref = F.conv_transpose2d(x, wt, bias, 2)
    xd = _nhwc(x).to(cuda)
    cat = torch.full((n, 2 * h, 2 * w, skip + cout), 3.0, dtype=torch.bfloat16, device=cuda)
    co.convT_fprop(co.View(xd), co.pack_convT_fprop(wt.to(cuda)), co.View(cat, skip, cout),
                   bias=bias.to(cuda)).run()
    torch.cuda.synchronize()
    _close(_nchw(cat[..., skip:]), y_ref, "convT")
    if skip:
        assert bool((cat[..., :skip] == 3.0).all())   # the skip slice is untouched

    dcat = _rb(torch.randn(n, skip + cout, 2 * h, 2 * w, generator=g))
    dy = dcat[:, skip:]
    xr = x.clone().requires_grad_(True)
    wr = wt.clone().requires_grad_(True)
    F.conv_transpose2d(xr, wr, bias, 2).backward(dy)
    dcd = _nhwc(dcat).to(cuda)
    dxd = torch.zeros(n, h, w, cin, dtype=torch.bfloat16, device=cuda)
    co.convT_dgrad(co.View(dcd, skip, cout), co.pack_convT_dgrad(wt.to(cuda)), co.View(dxd)).run()
    gw = torch.zeros(cin, cout, 2, 2, dtype=torch.float32, device=cuda)
    ws = torch.empty(co.wgrad_workspace_bytes(cout, cin, 4, n, h, w) // 4 + 4, dtype=torch.float32, device=cuda)
    co.convT_wgrad(co.View(xd), co.View(dcd, skip, cout), gw, ws).run()
    torch.cuda.synchronize()
    _close(_nchw(dxd), xr.grad, "convT_dgrad")
    np.testing.assert_allclose(gw.cpu().numpy(), wr.grad.numpy(), rtol=2e-3,
                               atol=2e-4 * float(wr.grad.abs().max()))


def test_seq_replays_ops(cuda):
    from b2seg import convops as co
    x = torch.randn(2, 16, 16, 64, device=cuda).to(torch.bfloat16)
    w1 = co.pack_conv_fprop(torch.randn(64, 64, 3, 3, device=cuda) / 24)
    y1 = torch.zeros_like(x)
    y2 = torch.zeros_like(x)
    seq = co.Seq([co.conv_fprop(co.View(x), w1, 3, 1, co.View(y1)), co.conv_fprop(co.View(y1), w1, 3, 1, co.View(y2))])
    assert seq.launches == 2
    seq.run()
    torch.cuda.synchronize()
    ref = F.conv2d(F.conv2d(x.float().permute(0, 3, 1, 2), w1.float().reshape(3, 3, 64, 64).permute(2, 3, 0, 1), None, 1, 1)
                   .to(torch.bfloat16).float(), w1.float().reshape(3, 3, 64, 64).permute(2, 3, 0, 1), None, 1, 1)
    _close(y2.float().permute(0, 3, 1, 2).cpu(), ref.cpu(), "seq")


# ---- round 2: the TMA-store / reduce-add epilogue and the second MMA issuer against the plain paths ----
VARIANT_CASES = [
    # n, h, w, cin, cout, k, stride
    (8, 64, 64, 64, 64, 3, 1),     # resident weights, two issuers, TMA epilogue
    (2, 40, 24, 96, 32, 3, 1),     # ragged tiles (TMA clips), partial last K chunk, N = 32
    (4, 16, 16, 256, 256, 3, 1),   # streamed weights (one issuer), N = 128
    (2, 32, 32, 64, 32, 1, 1),     # tapgemm: 1x1, N = 32 (register statistics)
    (2, 16, 16, 128, 64, 1, 1),    # tapgemm: N = 64 (single staging slot)
]


@pytest.mark.parametrize("n,h,w,cin,cout,k,stride", VARIANT_CASES)
def test_epilogue_and_issuer_variants_agree(cuda, monkeypatch, n, h, w, cin, cout, k, stride):
    """B2SEG_HALO_TMA_OUT / B2SEG_TG_TMA_OUT / B2SEG_HALO_ISSUERS are read when an op is created: the same
    problem planned with the TMA epilogue + two issuers (default) and with direct stores + one issuer must give
    bit-identical outputs and statistics; in accumulate mode the TMA reduce-add rounds the new term to bf16
    before the L2 adds it (one extra rounding), so it is compared within bf16 resolution."""
    from b2seg import convops as co
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(7 * n + h + cin + cout + k)
    x = torch.randn(n, h, w, cin, generator=g).to(torch.bfloat16).to(dev)
    dy = torch.randn(n, h // stride, w // stride, cout, generator=g).to(torch.bfloat16).to(dev)
    wt = (torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).to(dev)
    base = torch.randn(n, h, w, cin, generator=g).to(torch.bfloat16).to(dev)
    ho, wo = h // stride, w // stride

    def run(plain):
        for name in ("B2SEG_HALO_TMA_OUT", "B2SEG_TG_TMA_OUT"):
            if plain:
                monkeypatch.setenv(name, "0")
            else:
                monkeypatch.delenv(name, raising=False)
        if plain:
            monkeypatch.setenv("B2SEG_HALO_ISSUERS", "1")
        else:
            monkeypatch.delenv("B2SEG_HALO_ISSUERS", raising=False)
        y = torch.zeros(n, ho, wo, cout, dtype=torch.bfloat16, device=dev)
        stats = torch.zeros(co.stat_rows(n, ho, wo), 2, cout, device=dev)
        dx = torch.zeros(n, h, w, cin, dtype=torch.bfloat16, device=dev)
        dxa = base.clone()
        co.conv_fprop(co.View(x), co.pack_conv_fprop(wt), k, stride, co.View(y), stats=stats).run()
        co.conv_dgrad(co.View(dy), co.pack_conv_dgrad(wt), k, stride, co.View(dx)).run()
        co.conv_dgrad(co.View(dy), co.pack_conv_dgrad(wt), k, stride, co.View(dxa), accumulate=True).run()
        torch.cuda.synchronize()
        return y, stats.sum(0), dx, dxa

    y0, s0, dx0, dxa0 = run(plain=True)
    y1, s1, dx1, dxa1 = run(plain=False)
    assert torch.equal(y0, y1)
    assert torch.equal(dx0, dx1)
    assert torch.allclose(s0, s1, rtol=1e-5, atol=1e-3)  # rows are dealt to CTA groups differently, sums agree
    d = (dxa0.float() - dxa1.float()).abs()
    tol = 2.0 ** -7 * dxa0.float().abs() + 2.0 ** -8 * (dxa0.float() - base.float()).abs() + 1e-6
    assert not (d > tol).any(), float(d.max())
