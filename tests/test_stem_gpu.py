"""The encoder stem (7x7 stride-2 conv of the 1-channel image, advanced_models.py:76,268) on the
tensor cores against torch's fp32 convolution of the same bf16-rounded operands, with its
BatchNorm partial sums and its weight gradient; ragged sizes leave partly empty 8x16 bricks."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,h,w", [(2, 64, 64), (1, 34, 50), (3, 96, 160), (2, 256, 256)])
def test_stem_forward_stats_and_wgrad(cuda, n, h, w):
    from b2seg import ewops as eo, _lib as L
    g = torch.Generator().manual_seed(n * 1000 + h + w)
    x = torch.rand(n, h, w, generator=g).to(cuda)
    wt = (torch.randn(64, 1, 7, 7, generator=g) * 0.2).to(cuda)
    oh, ow = h // 2, w // 2
    y = torch.full((n, oh, ow, 64), float("nan"), dtype=torch.bfloat16, device=cuda)
    rows = eo.stem_stat_rows(n, h, w)
    st = torch.full((rows, 2, 64), float("nan"), dtype=torch.float32, device=cuda)
    eo.ew(L.EW_STEM_FWD, p=[x, wt, y, st], i=[n, h, w]).run()
    # reference: fp32 conv of the operands as the tensor cores see them (bf16-rounded)
    xr, wr = x.bfloat16().float(), wt.bfloat16().float()
    ref = F.conv2d(xr[:, None], wr, stride=2, padding=3).permute(0, 2, 3, 1)
    got = y.float()
    assert torch.isfinite(got).all()
    err = (got - ref).norm() / ref.norm()
    assert float(err) < 3e-3, float(err)          # bf16 output rounding only
    # against the exact fp32 convolution (what the reference computes): bf16 operand rounding
    exact = F.conv2d(x[:, None], wt, stride=2, padding=3).permute(0, 2, 3, 1)
    assert float((got - exact).norm() / exact.norm()) < 6e-3
    # BatchNorm partials: fixed-order fp32 sums of the UNROUNDED accumulators
    s = st.double().sum(0)
    torch.testing.assert_close(s[0], ref.double().sum((0, 1, 2)), rtol=2e-3, atol=2e-2)
    torch.testing.assert_close(s[1], (ref.double() ** 2).sum((0, 1, 2)), rtol=2e-3, atol=2e-2)
    # weight gradient of the stem from a random dy
    dy = (torch.randn(n, oh, ow, 64, generator=g) * 0.5).to(cuda).bfloat16()
    wrows = eo.stem_wgrad_rows(n, h, w)
    part = torch.empty(wrows * 3136 + 64, dtype=torch.float32, device=cuda)
    dw = torch.zeros(64 * 49, dtype=torch.float32, device=cuda)
    eo.ew(L.EW_STEM_WGRAD, p=[x, dy, part], i=[n, h, w]).run()
    eo.ew(L.EW_COLSUM_FINALIZE, p=[part, dw], i=[wrows, 3136, 0, 3136]).run()
    xg = x[:, None].clone().requires_grad_(False)
    wg = wt.clone().requires_grad_(True)
    F.conv2d(xg, wg, stride=2, padding=3).backward(dy.float().permute(0, 3, 1, 2))
    refw = wg.grad.reshape(-1)
    assert float((dw - refw).norm() / refw.norm()) < 6e-3
