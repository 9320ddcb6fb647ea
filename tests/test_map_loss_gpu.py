"""Boundary-style losses as CUDA kernels (csrc/map_loss.cu, the exp-log term of csrc/seg_loss.cu;
SURVEY 8f-4): the device distance transform against scipy (bit-exact), the boundary weight map
against the torch formula of the reference (float sum order: 2e-6), the three losses and their
gradients against the oracle and the reference-generated fixtures."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import seg_losses

pytestmark = pytest.mark.gpu


def _blobs(rng, b, h, w):
    from scipy.ndimage import gaussian_filter
    return np.stack([(gaussian_filter(rng.random((h, w)), 4) > 0.5).astype(np.float32) for _ in range(b)])[:, None]


@pytest.mark.parametrize("b,h,w", [(4, 256, 256), (3, 37, 53), (2, 128, 300), (1, 1, 17), (1, 23, 1), (2, 1024, 1024)])
def test_distance_transform_is_scipy_exact(cuda, b, h, w):
    from b2seg import lossfn
    rng = np.random.default_rng(b * h + w)
    t = _blobs(rng, b, h, w)
    t[0, 0] = 1.0 if b > 1 else t[0, 0]            # a plane without any zero: scipy's (-1, 0) convention
    if b > 2:
        t[1, 0] = 0.0                               # an empty plane
        t[2, 0] = 1.0
        t[2, 0, h // 2, w // 3] = 0.0               # a single zero: the longest searches
    dev = torch.from_numpy(t).to(cuda)
    got = lossfn.distance_transform(dev, mode=0).cpu().numpy()
    for i in range(b):
        ref = seg_losses.edt(t[i, 0])
        assert np.array_equal(got[i, 0], ref.astype(np.float32)), i
    # soft planes (Hausdorff feeds sigmoid outputs): every non-zero value is foreground
    soft = (rng.random((2, 1, min(h, 64), min(w, 64))).astype(np.float32) + 0.01)
    soft[1, 0, 0, 0] = 0.0
    got = lossfn.distance_transform(torch.from_numpy(soft).to(cuda), mode=0).cpu().numpy()
    for i in range(2):
        assert np.array_equal(got[i, 0], seg_losses.edt(soft[i, 0]).astype(np.float32))
    # weight map of DistanceMapPenalizedCrossEntropy: 1 + alpha * d / (max + 1e-7), float32, bit-exact
    for alpha in (1.0, 0.37):
        got = lossfn.distance_transform(dev, mode=1, alpha=alpha).cpu()
        ref = 1 + alpha * seg_losses.distance_map(torch.from_numpy(t))
        assert torch.equal(got.view(ref.shape), ref)


@pytest.mark.parametrize("shape,theta0,theta", [((4, 1, 256, 256), 3, 5), ((2, 1, 37, 53), 3, 5), ((2, 2, 64, 40), 5, 7),
                                                ((1, 1, 8, 8), 7, 15), ((3, 1, 33, 31), 1, 1)])
def test_boundary_weight_map(cuda, shape, theta0, theta):
    from b2seg import lossfn
    rng = np.random.default_rng(theta0 * 100 + theta)
    t = _blobs(rng, shape[0] * shape[1], shape[2], shape[3]).reshape(shape)
    tt = torch.from_numpy(t)
    ref = seg_losses.boundary_weight_map(tt, theta0, theta)
    k = seg_losses.gaussian_kernel(theta, theta / 3.0)
    got = lossfn.boundary_weight_map(tt.to(cuda), theta0, theta, k, 10.0).view(shape).cpu()
    assert float((got - ref).abs().max()) <= 2e-6 * float(ref.abs().max())
    soft = torch.rand(shape)                       # soft targets: plain max / min arithmetic
    ref = seg_losses.boundary_weight_map(soft, theta0, theta)
    got = lossfn.boundary_weight_map(soft.to(cuda), theta0, theta, k, 10.0).view(shape).cpu()
    assert float((got - ref).abs().max()) <= 2e-6 * float(ref.abs().max())
    with pytest.raises(Exception):
        lossfn.boundary_weight_map(tt.to(cuda), 4, 5, k, 10.0)


@pytest.mark.parametrize("tag,shape", [("kat", (4, 1, 256, 256)), ("small", (2, 1, 32, 48))])
def test_map_losses_match_oracle_and_reference_fixture(cuda, golden_dir, tag, shape):
    import losses
    with open(os.path.join(golden_dir, "loss_kat.json")) as f:
        kat = json.load(f)
    logits, target = seg_losses.kat_inputs(shape)
    for name in seg_losses.MAP_LOSSES:
        crit = losses.get_loss_function({"loss_fn": name})
        x = logits.to(cuda).requires_grad_(True)
        val = crit(x, target.to(cuda))
        val.backward()
        g = x.grad.cpu().double()
        ref = kat[tag][name]
        oval, ograd = seg_losses.loss_and_grad(name, logits, target)
        assert float(val) == pytest.approx(ref["loss"], rel=2e-5), name
        assert float(val) == pytest.approx(oval, rel=2e-5), name
        assert float(g.abs().sum()) == pytest.approx(ref["grad_l1"], rel=1e-3), name
        if name == "exponential_logarithmic":
            # pointwise where the reference's own float32 rounding of (1 - ce) is not the signal:
            # pixels with ce > 1e-3; everything together through the L1 norm above
            p = torch.sigmoid(logits.double())
            ce = -(target * torch.log(p) + (1 - target) * torch.log1p(-p))
            m = (ce > 1e-3).reshape(g.shape)
            np.testing.assert_allclose(g[m].numpy(), ograd.reshape(g.shape)[m].numpy(), rtol=2e-3, atol=1e-10)
            assert crit.last_stats is not None and float(crit.last_stats[8]) == logits.numel()
        else:
            np.testing.assert_allclose(g.numpy(), ograd.reshape(g.shape).numpy(), rtol=2e-4, atol=1e-10, err_msg=name)


def test_explog_extremes_and_hausdorff(cuda):
    """Saturated logits: the clamped branch (1 - ce < 1e-7) has value (-log 1e-7)^g and zero slope;
    p == t exactly contributes 0 (the reference yields NaN there: stated in DESIGN).  Hausdorff's
    distance maps now come from the device transform; the value equals the oracle's formula."""
    import losses
    x = torch.tensor([40.0, -40.0, 40.0, -40.0, 0.3, -2.0, 25.0, -25.0], device=cuda).view(1, 1, 2, 4).requires_grad_(True)
    t = torch.tensor([1.0, 0.0, 0.0, 1.0, 1.0, 0.0, 1.0, 0.0], device=cuda).view(1, 1, 2, 4)
    crit = losses.ExponentialLogarithmicLoss()
    val = crit(x, t)
    val.backward()
    assert torch.isfinite(val) and torch.isfinite(x.grad).all()
    assert abs(float(x.grad.view(-1)[0])) < 1e-12 and abs(float(x.grad.view(-1)[1])) < 1e-12  # only the Dice part is left
    ov, _ = seg_losses.loss_and_grad("exponential_logarithmic", x.detach().cpu(), t.cpu())
    assert float(val) == pytest.approx(ov, rel=1e-5)
    logits, target = seg_losses.kat_inputs((2, 1, 32, 48))
    hd = losses.HausdorffDistanceLoss()
    got = float(hd(logits.to(cuda), target.to(cuda)))
    prob = torch.sigmoid(logits)
    dist = seg_losses.hausdorff_dtm(prob) ** 2.0 + seg_losses.hausdorff_dtm(target) ** 2.0
    assert got == pytest.approx(float(((prob - target) ** 2 * dist).mean()), rel=1e-5)
