"""The oracle restatements are pinned against fixtures produced by the reference's own code
(tests/golden/make_golden.py) and, for the OpenCV arithmetic, against cv2 itself."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import augment_cv as A
from oracle import seg_losses, unet_functional


@pytest.fixture(scope="module")
def loss_kat(golden_dir):
    with open(os.path.join(golden_dir, "loss_kat.json")) as f:
        return json.load(f)


@pytest.mark.parametrize("tag,shape", [("kat", (4, 1, 256, 256)), ("small", (2, 1, 32, 48))])
def test_loss_oracle_matches_reference(loss_kat, tag, shape):
    logits, target = seg_losses.kat_inputs(shape)
    assert abs(float(logits.double().sum()) - loss_kat[tag]["logits_sum"]) < 1e-6
    assert float(target.sum()) == loss_kat[tag]["target_sum"]
    for cfg_tag, cfg in (("", {}), ("@custom", loss_kat[tag]["custom_cfg"])):
        for name in seg_losses.CLOSED_FORM:
            ref = loss_kat[tag][name + cfg_tag]
            val, grad = seg_losses.loss_and_grad(name, logits, target, cfg)
            # the fp32 reference vs a float64 restatement: 2e-5 relative
            assert val == pytest.approx(ref["loss"], rel=2e-5), (name, cfg_tag)
            assert float(grad.abs().sum()) == pytest.approx(ref["grad_l1"], rel=2e-4), (name, cfg_tag)


def test_loss_oracle_pointwise_grads(golden_dir):
    grads = np.load(os.path.join(golden_dir, "loss_small_grads.npz"))
    logits, target = seg_losses.kat_inputs((2, 1, 32, 48))
    for name in seg_losses.CLOSED_FORM:
        _, g = seg_losses.loss_and_grad(name, logits, target)
        ref = grads[f"grad_{name}"]
        np.testing.assert_allclose(g.numpy(), ref, rtol=2e-3, atol=1e-9, err_msg=name)


def test_metrics_oracle(loss_kat):
    for tag, shape in (("kat", (4, 1, 256, 256)), ("small", (2, 1, 32, 48))):
        logits, target = seg_losses.kat_inputs(shape)
        m = seg_losses.confusion_metrics(torch.sigmoid(logits), target)
        for k, v in loss_kat[tag]["metrics"].items():
            assert m[k] == pytest.approx(v, abs=1e-9), (tag, k)
        assert m["iou"] == pytest.approx(loss_kat[tag]["iou_score"], abs=1e-7)


def _ref_like_state(backbone, attn):
    """Shapes of the reference state dict, rebuilt from torchvision + the decoder spec."""
    from oracle.state_layout import reference_state_shapes
    return reference_state_shapes(backbone, attn)


@pytest.mark.parametrize("backbone,attn,size,batch", [("resnet34", True, 64, 2), ("resnet34", False, 64, 2),
                                                      ("resnet50", False, 64, 2), ("resnet34", True, 96, 1)])
def test_unet_oracle_matches_reference(golden_dir, backbone, attn, size, batch):
    tag = f"{backbone}_{'attn' if attn else 'noattn'}_{size}"
    with open(os.path.join(golden_dir, "model_kat.json")) as f:
        meta = json.load(f)[tag]
    arr = np.load(os.path.join(golden_dir, "model_kat.npz"))
    shapes = _ref_like_state(backbone, attn)
    assert len(shapes) == meta["n_state"]
    sd = unet_functional.synth_state(shapes, seed=3)
    g = torch.Generator().manual_seed(11)
    x = torch.rand(batch, 1, size, size, generator=g)
    t = (torch.rand(batch, 1, size, size, generator=g) > 0.8).float()
    new_stats = {}
    loss_fn = lambda lg, tt: seg_losses.loss_value("combo", lg, tt).float()
    logits, loss, grads = unet_functional.forward_backward(
        sd, x, t, loss_fn, backbone=backbone, use_attention=attn, training=True, new_stats=new_stats)
    np.testing.assert_allclose(logits.numpy(), arr[tag + "_logits_train"], rtol=1e-3, atol=2e-4)
    assert loss == pytest.approx(meta["loss"], rel=1e-4)
    np.testing.assert_allclose(new_stats["bn1.running_mean"].numpy(), arr[tag + "_bn1_running_mean"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(new_stats["bn1.running_var"].numpy(), arr[tag + "_bn1_running_var"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(grads["conv_final.weight"].numpy(), arr[tag + "_grad_conv_final"], rtol=2e-3, atol=1e-6)
    np.testing.assert_allclose(grads["input_conv.weight"].numpy(), arr[tag + "_grad_input_conv"], rtol=5e-3, atol=1e-5)
    for k, ref in meta["grad_norms"].items():
        got = float(grads[k].double().norm())
        assert got == pytest.approx(ref, rel=5e-3, abs=1e-6), k
    with torch.no_grad():
        ev = unet_functional.unet_forward(dict(sd, **new_stats), x, backbone=backbone, use_attention=attn,
                                          training=False)
    np.testing.assert_allclose(ev.numpy(), arr[tag + "_logits_eval"], rtol=2e-3, atol=2e-3)  # fp32 conv-order noise


def test_augment_oracle_matches_cv2():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(0)
    flags = {A.LINEAR: cv2.INTER_LINEAR, A.NEAREST: cv2.INTER_NEAREST}
    borders = {A.CONSTANT: cv2.BORDER_CONSTANT, A.REFLECT: cv2.BORDER_REFLECT, A.REFLECT101: cv2.BORDER_REFLECT_101}
    for (h, w) in [(64, 64), (97, 131), (256, 256)]:
        img = cv2.GaussianBlur(rng.integers(0, 256, (h, w), dtype=np.uint8), (5, 5), 1.5)
        assert np.array_equal(A.hflip(img), cv2.flip(img, 1))
        assert np.array_equal(A.vflip(img), cv2.flip(img, 0))
        assert np.array_equal(A.transpose(img), cv2.transpose(img))
        for k in range(4):
            assert np.array_equal(A.rot90(img, k), np.rot90(img, k))
        for _ in range(3):
            M = A.affine_matrix(h, w, rng.uniform(0.9, 1.1), (rng.uniform(-.05, .05), rng.uniform(-.05, .05)),
                                rng.uniform(-15, 15), (rng.uniform(-5, 5), rng.uniform(-5, 5)))
            for it, cf in flags.items():
                for bt, cb in borders.items():
                    ref = cv2.warpAffine(img, M, (w, h), flags=cf, borderMode=cb, borderValue=0)
                    assert np.array_equal(A.warp_affine_u8(img, M, it, bt, 0), ref)
        mx, my = A.elastic_maps(h, w, 8.0, 4.0, rng)
        for it, cf in flags.items():
            for bt, cb in borders.items():
                ref = cv2.remap(img, mx, my, cf, borderMode=cb, borderValue=0)
                assert np.array_equal(A.remap_u8(img, mx, my, it, bt, 0), ref)


def _smooth_u8(rng, h, w):
    cv2 = pytest.importorskip("cv2")
    a = cv2.GaussianBlur(rng.integers(0, 256, (h, w), dtype=np.uint8), (0, 0), 3)
    return cv2.normalize(a, None, 0, 255, cv2.NORM_MINMAX).astype(np.uint8)


@pytest.mark.parametrize("h,w,dh,dw", [(96, 96, 64, 64), (512, 512, 256, 256), (520, 696, 256, 256), (300, 300, 128, 128),
                                       (100, 130, 256, 256), (97, 131, 64, 64), (64, 64, 96, 96), (64, 64, 64, 64)])
def test_preproc_oracle_resize_matches_cv2(h, w, dh, dw):
    """oracle/preproc_cv.py against cv2.resize itself (dataset.py:52-53): AREA (all three of OpenCV's
    code paths: integer-factor, fractional, enlarging) and NEAREST bit-exact; LINEAR bit-exact when
    shrinking, within 1 LSB on < 1 % of the pixels when enlarging (cv2's own SIMD path)."""
    cv2 = pytest.importorskip("cv2")
    from oracle import preproc_cv as P
    rng = np.random.default_rng(h * 7 + dw)
    for img in (_smooth_u8(rng, h, w), rng.integers(0, 256, (h, w), dtype=np.uint8)):
        assert np.array_equal(P.resize_area_u8(img, dh, dw), cv2.resize(img, (dw, dh), interpolation=cv2.INTER_AREA))
        assert np.array_equal(P.resize_nearest_u8(img, dh, dw), cv2.resize(img, (dw, dh), interpolation=cv2.INTER_NEAREST))
        d = np.abs(P.resize_linear_u8(img, dh, dw).astype(int) - cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR).astype(int))
        assert d.max() <= (1 if (dh > h or dw > w) else 0) and (d > 0).mean() < 0.01


@pytest.mark.parametrize("h,w", [(256, 256), (128, 128), (64, 64), (250, 250), (100, 130)])
def test_preproc_oracle_normalize_matches_reference_arithmetic(h, w):
    """normalize_microscopy_image (dataset.py:30-43) on the host with np.percentile and cv2's CLAHE
    vs the restatement: clip, CLAHE and the final float image all bit-exact (incl. sizes the 8x8
    tile grid does not divide, where OpenCV pads by reflection, and a constant image)."""
    cv2 = pytest.importorskip("cv2")
    from oracle import preproc_cv as P
    rng = np.random.default_rng(h + 3 * w)
    for img in (_smooth_u8(rng, h, w), rng.integers(0, 256, (h, w), dtype=np.uint8),
                (rng.integers(0, 40, (h, w)) + 100).astype(np.uint8), np.full((h, w), 77, np.uint8)):
        lo, hi = np.percentile(img, [2, 98])
        clipped = np.clip(img, lo, hi).astype(np.uint8)
        eq = cv2.createCLAHE(clipLimit=2.0, tileGridSize=(8, 8)).apply(clipped)
        ref = (eq - eq.min()) / (eq.max() - eq.min() + 1e-8)
        got, pct = P.percentile_clip_u8(img)
        assert np.array_equal(got, clipped) and pct == [lo, hi]
        assert np.array_equal(P.clahe_u8(clipped), eq)
        assert np.array_equal(P.normalize_microscopy_image(img), ref.astype(np.float32))


def test_augmentation_map_oracles_match_their_formulas():
    """oracle/augment_cv.py's grid / optical map builders: identity when undistorted, monotone,
    and the maps drive cv2.remap to the same picture as the oracle's own remap."""
    cv2 = pytest.importorskip("cv2")
    h, w = 90, 120
    mx, my = A.grid_distortion_maps(h, w, [1.0] * 6, [1.0] * 6)
    xs, ys = np.meshgrid(np.arange(w, dtype=np.float32), np.arange(h, dtype=np.float32))
    # albumentations' linspace per cell includes both end points: identity to within a pixel
    assert np.abs(mx - xs).max() <= 1.5 and np.abs(my - ys).max() <= 1.5
    mx, my = A.grid_distortion_maps(h, w, [1.2, 0.8, 1.1, 0.9, 1.0, 1.3], [0.7, 1.3, 1.0, 1.1, 0.9, 1.0])
    assert (np.diff(mx, axis=1) >= 0).all() and (np.diff(my, axis=0) >= 0).all()
    ox, oy = A.optical_distortion_maps(h, w, 0.0, 0, 0)
    assert np.allclose(ox, xs, atol=1e-3) and np.allclose(oy, ys, atol=1e-3)
    ox, oy = A.optical_distortion_maps(h, w, 0.4, 3, -2)
    img = _smooth_u8(np.random.default_rng(0), h, w)
    assert np.array_equal(A.remap_u8(img, ox, oy, A.LINEAR, A.REFLECT, 0),
                          cv2.remap(img, ox, oy, cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT))


def test_map_loss_oracle_matches_reference_fixture(loss_kat):
    """The boundary-style losses of the oracle against the values the reference itself produced
    (tests/golden/make_golden.py), and its distance transform against the definition."""
    for tag, shape in (("kat", (4, 1, 256, 256)), ("small", (2, 1, 32, 48))):
        logits, target = seg_losses.kat_inputs(shape)
        for name in seg_losses.MAP_LOSSES:
            ref = loss_kat[tag][name]
            val, grad = seg_losses.loss_and_grad(name, logits, target)
            assert val == pytest.approx(ref["loss"], rel=2e-5), name
            assert float(grad.abs().sum()) == pytest.approx(ref["grad_l1"], rel=2e-4), name
    rng = np.random.default_rng(0)
    for plane in ((rng.random((23, 31)) > 0.3).astype(np.float32), np.ones((9, 14), np.float32),
                  np.zeros((5, 5), np.float32), rng.random((12, 12)).astype(np.float32) + 0.1):
        assert np.array_equal(seg_losses.edt(plane), seg_losses.edt_brute(plane))
