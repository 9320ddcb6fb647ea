"""CUDA loss / metric kernels vs the oracle and the reference-generated fixtures (through the C ABI)."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import seg_losses

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def loss_kat(golden_dir):
    with open(os.path.join(golden_dir, "loss_kat.json")) as f:
        return json.load(f)


def _run(name, logits, target, cfg, dev):
    import losses
    crit = losses.get_loss_function(dict(cfg, loss_fn=name))
    x = logits.to(dev).requires_grad_(True)
    val = crit(x, target.to(dev))
    val.backward()
    return float(val), x.grad.cpu(), crit


@pytest.mark.parametrize("tag,shape", [("kat", (4, 1, 256, 256)), ("small", (2, 1, 32, 48))])
def test_losses_match_reference_fixture(cuda, loss_kat, tag, shape):
    logits, target = seg_losses.kat_inputs(shape)
    for cfg_tag, cfg in (("", {}), ("@custom", loss_kat[tag]["custom_cfg"])):
        for name in seg_losses.CLOSED_FORM:
            ref = loss_kat[tag][name + cfg_tag]
            val, grad, _ = _run(name, logits, target, cfg, cuda)
            # north star: loss within 1e-3 relative; we hold 2e-5
            assert val == pytest.approx(ref["loss"], rel=2e-5), (name, cfg_tag)
            assert float(grad.abs().sum()) == pytest.approx(ref["grad_l1"], rel=2e-4), (name, cfg_tag)
    for name in ("distance_map_bce", "boundary", "exponential_logarithmic"):
        ref = loss_kat[tag][name]
        val, grad, _ = _run(name, logits, target, {}, cuda)
        assert val == pytest.approx(ref["loss"], rel=1e-4), name
        assert float(grad.abs().sum()) == pytest.approx(ref["grad_l1"], rel=1e-3), name


def test_pointwise_gradients(cuda, golden_dir):
    grads = np.load(os.path.join(golden_dir, "loss_small_grads.npz"))
    logits, target = seg_losses.kat_inputs((2, 1, 32, 48))
    for name in seg_losses.CLOSED_FORM:
        _, g, _ = _run(name, logits, target, {}, cuda)
        _, go = seg_losses.loss_and_grad(name, logits, target)
        np.testing.assert_allclose(g.numpy(), go.numpy(), rtol=2e-4, atol=1e-9, err_msg=name)
        np.testing.assert_allclose(g.numpy(), grads[f"grad_{name}"], rtol=2e-3, atol=1e-9, err_msg=name)


@pytest.mark.parametrize("n", [1, 3, 4, 5, 1023, 4097, 65536 + 7])
def test_ragged_sizes_and_misalignment(cuda, n):
    import losses
    g = torch.Generator().manual_seed(n)
    x = torch.randn(n + 1, generator=g) * 3
    t = (torch.rand(n + 1, generator=g) > 0.7).float()
    for off in (0, 1):  # off=1 -> pointers not 16-byte aligned -> scalar path
        xs, ts = x[off:off + n], t[off:off + n]
        xd = x.to(cuda)[off:off + n].detach().requires_grad_(True)
        td = t.to(cuda)[off:off + n]
        for name in ("combo", "focal", "tversky_recall", "balanced_bce"):
            crit = losses.get_loss_function({"loss_fn": name})
            val = crit(xd, td)
            (gd,) = torch.autograd.grad(val, xd)
            vo, go = seg_losses.loss_and_grad(name, xs, ts)
            assert float(val) == pytest.approx(vo, rel=2e-5, abs=1e-7), (name, n, off)
            np.testing.assert_allclose(gd.cpu().numpy(), go.numpy(), rtol=3e-4, atol=1e-8)


def test_soft_targets(cuda):
    """Targets strictly between 0 and 1 (label smoothing / resized masks): the focal term takes
    its general exp(-bce) branch instead of the hard-target shortcut."""
    import losses
    g = torch.Generator().manual_seed(11)
    x = torch.randn(2, 1, 40, 52, generator=g) * 3
    t = torch.rand(2, 1, 40, 52, generator=g)
    t[0, 0, :4] = 0.0
    t[0, 0, 4:8] = 1.0
    for name in ("focal", "triple_combo", "combo", "dice"):
        crit = losses.get_loss_function({"loss_fn": name})
        xd = x.to(cuda).requires_grad_(True)
        val = crit(xd, t.to(cuda))
        (gd,) = torch.autograd.grad(val, xd)
        vo, go = seg_losses.loss_and_grad(name, x, t)
        assert float(val) == pytest.approx(vo, rel=2e-5, abs=1e-7), name
        np.testing.assert_allclose(gd.cpu().numpy(), go.numpy(), rtol=3e-4, atol=1e-8, err_msg=name)


def test_global_batch_loss_across_replicas(cuda):
    """Data-parallel exactness: each replica reduces its shard, the eight sums and
    the pixel count are added (what the NCCL all-reduce does) and b2seg_seg_loss_finish evaluates
    the loss on the totals -- value, confusion counts and per-pixel gradient must equal the
    single-process loss on the whole batch, for losses whose ratios do not decompose over shards."""
    import ctypes as C
    import losses
    from b2seg import lossfn, _lib as L
    logits, target = seg_losses.kat_inputs((4, 1, 64, 96))
    x, t = logits.to(cuda), target.to(cuda)
    shards = [(x[:1], t[:1]), (x[1:], t[1:])]  # unequal shards: the pixel count is summed too
    for name in ("dice", "tversky_recall", "combo", "balanced_bce", "focal", "sensitivity_specificity"):
        spec = losses.get_loss_function({"loss_fn": name})._spec()
        full_loss, full_stats, full_coef, saved = lossfn.loss_forward(x, t, spec)
        g_full = lossfn.loss_backward(saved, full_coef, None, x.shape)
        parts = [lossfn.loss_forward(xs, ts, spec) for xs, ts in shards]
        tot = torch.zeros(L.STATS_LEN, dtype=torch.float64, device=cuda)
        tot[:L.S_N + 1] = sum(p[1][:L.S_N + 1] for p in parts)
        small = torch.empty(1 + L.COEF_LEN, dtype=torch.float32, device=cuda)
        L.check(L.lib().b2seg_seg_loss_finish(C.byref(spec), 0, tot.data_ptr(), small.data_ptr(),
                                              small.data_ptr() + 4, L.stream_ptr()))
        assert float(small[0]) == pytest.approx(float(full_loss), rel=1e-6), name
        assert float(tot[L.S_N]) == x.numel()
        for i in (L.S_TP, L.S_FP, L.S_FN, L.S_TN):
            assert float(tot[i]) == float(full_stats[i]), (name, i)
        g_parts = torch.cat([lossfn.loss_backward(p[3], small[1:], None, xs.shape)
                             for p, (xs, _) in zip(parts, shards)])
        torch.testing.assert_close(g_parts, g_full, rtol=1e-5, atol=1e-10)
        # the same loss evaluated per shard and averaged is NOT the global loss for ratio losses
        if name == "dice":
            per_shard = sum(float(p[0]) for p in parts) / 2
            assert abs(per_shard - float(full_loss)) > 1e-4


def test_metrics_match_fixture_and_counts_are_exact(cuda, loss_kat):
    import utils
    logits, target = seg_losses.kat_inputs((4, 1, 256, 256))
    m = utils.calculate_metrics(torch.sigmoid(logits).to(cuda), target.to(cuda))
    for k, v in loss_kat["kat"]["metrics"].items():
        assert m[k] == pytest.approx(v, abs=1e-6), k     # north star: Dice/IoU within 1e-3 absolute
    # fused path: counts from the loss forward equal the oracle's integer counts exactly
    import losses
    crit = losses.get_loss_function({"loss_fn": "combo"})
    crit(logits.to(cuda), target.to(cuda))
    st = crit.last_stats.cpu().numpy()
    p = torch.sigmoid(logits)
    pb = (p > 0.5).double()
    tt = target.double()
    assert st[10] == float((pb * tt).sum()) and st[11] == float((pb * (1 - tt)).sum())
    assert st[12] == float(((1 - pb) * tt).sum()) and st[13] == float(((1 - pb) * (1 - tt)).sum())


def test_large_shape_properties(cuda):
    """BASELINE roofline shape [64,1,1024,1024]: size-independent properties (the oracle would
    take minutes): linearity of the sums under batch concatenation and exact counts."""
    from b2seg import lossfn
    g = torch.Generator(device="cuda").manual_seed(5)
    n = 64 * 1024 * 1024
    x = torch.randn(n, device=cuda, generator=g) * 2
    t = (torch.rand(n, device=cuda, generator=g) > 0.9).float()
    spec = lossfn.make_spec(w_bce=0.5, w_dice=0.5, w_focal=0.1)
    _, s_all, _, _ = lossfn.loss_forward(x, t, spec)
    h = n // 2
    _, s_a, _, _ = lossfn.loss_forward(x[:h], t[:h], spec)
    _, s_b, _, _ = lossfn.loss_forward(x[h:], t[h:], spec)
    s_all, s_a, s_b = s_all.cpu().numpy(), s_a.cpu().numpy(), s_b.cpu().numpy()
    for i in range(6):
        assert s_all[i] == pytest.approx(s_a[i] + s_b[i], rel=1e-6)
    for i in (1, 6, 7, 8):  # integer-valued sums are exact well above 2^24
        assert s_all[i] == s_a[i] + s_b[i]
    assert s_all[1] == float(t.double().sum())
    assert s_all[6] == float((torch.sigmoid(x) > 0.5).double().sum())
    # determinism: same launch twice -> bitwise identical block
    _, s2, _, _ = lossfn.loss_forward(x, t, spec)
    assert np.array_equal(s2.cpu().numpy(), s_all)


def test_error_behaviour(cuda):
    import losses
    crit = losses.get_loss_function({"loss_fn": "dice"})
    with pytest.raises(Exception):
        crit(torch.zeros(4, 1, 8, 8), torch.zeros(4, 1, 8, 8))  # CPU tensors: no fallback
    with pytest.raises(ValueError):
        crit(torch.zeros(4, 1, 8, 8, device=cuda), torch.zeros(4, 1, 8, 4, device=cuda))
    assert type(losses.get_loss_function({"loss_fn": "nope"})).__name__ == "ComboLoss"
