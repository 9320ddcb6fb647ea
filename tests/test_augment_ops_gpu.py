"""Non-geometric augmentation kernels (csrc/augment_ops.cu), the batched warps and
the `albumentations` stand-in, against cv2 / numpy / scipy on the same inputs.  Integer work is
bit-exact; filter2D (float accumulation) within 1 LSB on < 0.1 % of the pixels; the Gaussian of the
elastic field within 1e-5; noise by its statistics and its determinism per seed."""
import random

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
cv2 = pytest.importorskip("cv2")


def _pair(rng, h, w):
    img = cv2.GaussianBlur(rng.integers(0, 256, (h, w), dtype=np.uint8), (0, 0), 2)
    msk = (rng.random((h, w)) > 0.8).astype(np.uint8) * 255
    return img, msk


def test_lut_brightness_contrast_and_rects(cuda):
    from b2seg import augment as ag
    rng = np.random.default_rng(1)
    img, msk = _pair(rng, 100, 131)
    planes = torch.from_numpy(np.stack([img, msk])).to(cuda)
    table = rng.integers(0, 256, 256, dtype=np.uint8)
    out = ag.lut_image(planes, table).cpu().numpy()
    assert np.array_equal(out[0], table[img]) and np.array_equal(out[1], msk)
    for cf in (False, True):
        got = ag.brightness_contrast(planes, 1.17, 0.83, contrast_first=cf).cpu().numpy()

        def bright(a):
            return cv2.LUT(a, np.clip(np.arange(0, 256) * 1.17, 0, 255).astype(np.uint8))

        def contr(a):
            lut = np.arange(0, 256) * 0.83 + a.mean() * (1 - 0.83)
            return cv2.LUT(a, np.clip(lut, 0, 255).astype(np.uint8))

        ref = bright(contr(img)) if cf else contr(bright(img))
        assert np.array_equal(got[0], ref) and np.array_equal(got[1], msk)
    rects = [(3, 5, 11, 13), (90, 120, 100, 131), (0, 0, 1, 1), (50, 60, 50, 70)]
    got = ag.fill_rects(planes, rects, 7, mask_value=None).cpu().numpy()
    ref = img.copy()
    for y1, x1, y2, x2 in rects:
        ref[y1:y2, x1:x2] = 7
    assert np.array_equal(got[0], ref) and np.array_equal(got[1], msk)


def test_filter2d_pad_and_gaussian(cuda):
    from b2seg import augment as ag
    from scipy.ndimage import gaussian_filter
    rng = np.random.default_rng(2)
    img, msk = _pair(rng, 97, 140)
    planes = torch.from_numpy(np.stack([img, msk])).to(cuda)
    for k in (3, 5, 7):
        kern = rng.random((k, k)).astype(np.float32)
        kern /= kern.sum()
        got = ag.filter2d_image(planes, kern).cpu().numpy()
        ref = cv2.filter2D(img, -1, kern)
        d = np.abs(got[0].astype(int) - ref.astype(int))
        assert d.max() <= 1 and (d > 0).mean() < 1e-3, (k, d.max(), (d > 0).mean())
        assert np.array_equal(got[1], msk)
    for border, flag in ((ag.REFLECT, cv2.BORDER_REFLECT), (ag.REFLECT101, cv2.BORDER_REFLECT_101), (ag.CONSTANT, cv2.BORDER_CONSTANT)):
        got = ag.pad(planes, 5, 9, 120, 3, border, 31).cpu().numpy()
        for p, src in enumerate((img, msk)):
            assert np.array_equal(got[p], cv2.copyMakeBorder(src, 5, 9, 120, 3, flag, value=31))
    field = rng.uniform(-1, 1, (2, 64, 80)).astype(np.float32)
    got = ag.gaussian_filter_f32(torch.from_numpy(field).to(cuda), 4.0, gain=3.0, add_coord=(1, 2)).cpu().numpy()
    xs, ys = np.meshgrid(np.arange(80), np.arange(64))
    assert np.abs(got[0] - (xs + 3.0 * gaussian_filter(field[0].astype(np.float64), 4.0))).max() < 1e-4
    assert np.abs(got[1] - (ys + 3.0 * gaussian_filter(field[1].astype(np.float64), 4.0))).max() < 1e-4


def test_gauss_noise_statistics_and_determinism(cuda):
    from b2seg import augment as ag
    planes = torch.full((2, 512, 512), 128, dtype=torch.uint8, device=cuda)
    a = ag.gauss_noise(planes, 0.0, 5.0, seed=11)
    b = ag.gauss_noise(planes, 0.0, 5.0, seed=11)
    c = ag.gauss_noise(planes, 0.0, 5.0, seed=12)
    assert torch.equal(a, b) and not torch.equal(a, c) and torch.equal(a[1], planes[1])
    d = a[0].float().cpu().numpy() - 128.0
    # truncating cast of 128 + N(0, 25): mean shifts by -0.5, variance grows by 1/12
    assert abs(d.mean() + 0.5) < 0.05 and abs(d.std() - np.sqrt(25 + 1 / 12)) < 0.05
    u = ag.uniform_field((4, 256, 256), 3, cuda).cpu().numpy()
    assert u.min() > -1 and u.max() < 1 and abs(u.mean()) < 0.01 and abs(u.var() - 1 / 3) < 0.01


def test_batched_warps_equal_single_plane_warps_and_cv2(cuda):
    from b2seg import augment as ag
    rng = np.random.default_rng(3)
    h, w = 120, 136
    pairs = [_pair(rng, h, w) for _ in range(5)]
    src = torch.from_numpy(np.stack([np.stack(p) for p in pairs])).to(cuda)
    mats = [ag.affine_matrix(h, w, scale=(1 + 0.03 * i, 1 - 0.02 * i), translate=(0.01 * i, -0.02), rotate_deg=7.0 * i - 10,
                             shear_deg=(i, -i)) for i in range(5)]
    for border, flag in ((ag.CONSTANT, cv2.BORDER_CONSTANT), (ag.REFLECT, cv2.BORDER_REFLECT)):
        got = ag.warp_affine_batch(src, mats, (ag.LINEAR, ag.NEAREST), border, (0, 0)).cpu().numpy()
        for i, (img, msk) in enumerate(pairs):
            m = np.array(mats[i], dtype=np.float64)
            assert np.array_equal(got[i, 0], cv2.warpAffine(img, m, (w, h), flags=cv2.INTER_LINEAR, borderMode=flag, borderValue=0))
            assert np.array_equal(got[i, 1], cv2.warpAffine(msk, m, (w, h), flags=cv2.INTER_NEAREST, borderMode=flag, borderValue=0))
            one = ag.warp_affine(src[i, 0:1], mats[i], ag.LINEAR, border).cpu().numpy()[0]
            assert np.array_equal(one, got[i, 0])
    mx, my = ag.optical_distortion_maps(h, w, 0.31, 2, -1, cuda)
    got = ag.remap_batch(src, mx, my, (ag.LINEAR, ag.NEAREST), ag.REFLECT).cpu().numpy()
    for i, (img, msk) in enumerate(pairs):
        assert np.array_equal(got[i, 0], cv2.remap(img, mx.cpu().numpy(), my.cpu().numpy(), cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT))
        assert np.array_equal(got[i, 1], cv2.remap(msk, mx.cpu().numpy(), my.cpu().numpy(), cv2.INTER_NEAREST, borderMode=cv2.BORDER_REFLECT))


def test_device_map_builders_match_oracle(cuda):
    from b2seg import augment as ag
    from oracle import augment_cv
    h, w = 96, 130
    xs, ys = [1.1, 0.8, 1.25, 0.9, 1.0, 1.2], [0.75, 1.3, 1.0, 0.85, 1.1, 0.95]
    gx, gy = ag.grid_distortion_maps(h, w, xs, ys, cuda)
    ox, oy = augment_cv.grid_distortion_maps(h, w, xs, ys)
    assert np.array_equal(gx.cpu().numpy(), ox) and np.array_equal(gy.cpu().numpy(), oy)
    gx, gy = ag.optical_distortion_maps(h, w, -0.37, 3, -2, cuda)
    ox, oy = augment_cv.optical_distortion_maps(h, w, -0.37, 3, -2)
    assert np.abs(gx.cpu().numpy() - ox).max() <= 2e-5 and np.abs(gy.cpu().numpy() - oy).max() <= 2e-5
    ex, ey = ag.elastic_maps(h, w, 30.0, 5.0, seed=9, device=cuda)
    xx, yy = np.meshgrid(np.arange(w), np.arange(h))
    dx, dy = ex.cpu().numpy() - xx, ey.cpu().numpy() - yy
    # a smooth field of the right scale: alpha * sigma_noise / (2 sqrt(pi) sigma) ~ 0.98
    assert 0.4 < dx.std() < 2.0 and 0.4 < dy.std() < 2.0 and np.abs(np.diff(dx, axis=1)).max() < 0.5
    assert not np.allclose(dx, dy)


def test_albumentations_stand_in_against_cv2(cuda):
    """Every geometric transform of the stand-in, with its own sampled parameters, equals cv2 applied
    with those parameters; the non-geometric ones leave the mask untouched; Compose composes."""
    import copy
    import albumentations as A
    from b2seg import augment as ag
    rng = np.random.default_rng(4)
    img, msk = _pair(rng, 128, 128)
    random.seed(123)
    cases = {
        "hflip": (A.HorizontalFlip(p=1.0), lambda a, p: cv2.flip(a, 1)),
        "vflip": (A.VerticalFlip(p=1.0), lambda a, p: cv2.flip(a, 0)),
        "transpose": (A.Transpose(p=1.0), lambda a, p: a.T),
        "rot90": (A.RandomRotate90(p=1.0), lambda a, p: np.rot90(a, p["factor"])),
    }
    for name, (t, ref) in cases.items():
        params = t.get_params(128, 128)
        out = t.apply_planes(torch.from_numpy(np.stack([img, msk])).to(cuda), **params).cpu().numpy()
        assert np.array_equal(out[0], ref(img, params)) and np.array_equal(out[1], ref(msk, params)), name
    t = A.Affine(scale=(0.95, 1.05), translate_percent=(-0.05, 0.05), rotate=(-15, 15), shear=(-5, 5),
                 interpolation=cv2.INTER_LINEAR, border_mode=cv2.BORDER_REFLECT, p=1.0)
    params = t.get_params(128, 128)
    out = t.apply_planes(torch.from_numpy(np.stack([img, msk])).to(cuda), **params).cpu().numpy()
    m = np.array(params["matrix"])
    assert np.array_equal(out[0], cv2.warpAffine(img, m, (128, 128), flags=cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT))
    assert np.array_equal(out[1], cv2.warpAffine(msk, m, (128, 128), flags=cv2.INTER_NEAREST, borderMode=cv2.BORDER_REFLECT))
    for t in (A.GridDistortion(num_steps=5, distort_limit=0.3, border_mode=cv2.BORDER_REFLECT, p=1.0),
              A.OpticalDistortion(distort_limit=0.5, shift_limit=0.5, border_mode=cv2.BORDER_REFLECT, p=1.0)):
        params = t.get_params(128, 128)
        out = t.apply_planes(torch.from_numpy(np.stack([img, msk])).to(cuda), **params).cpu().numpy()
        if isinstance(t, A.GridDistortion):
            mx, my = ag.grid_distortion_maps(128, 128, params["xsteps"], params["ysteps"], cuda)
        else:
            mx, my = ag.optical_distortion_maps(128, 128, params["k"], params["dx"], params["dy"], cuda)
        mx, my = mx.cpu().numpy(), my.cpu().numpy()
        assert np.array_equal(out[0], cv2.remap(img, mx, my, cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT))
        assert np.array_equal(out[1], cv2.remap(msk, mx, my, cv2.INTER_NEAREST, borderMode=cv2.BORDER_REFLECT))
    # every candidate of forward_selection_integration.py:62-181 constructs, deep-copies, takes p, and runs
    cands = [A.GaussNoise(var_limit=(10.0, 50.0), mean=0, per_channel=True, p=1.0), A.InvertImg(p=1.0),
             A.Solarize(threshold=128, p=1.0), A.RandomGamma(gamma_limit=(80, 120), p=1.0),
             A.ColorJitter(brightness=0.2, contrast=0.2, saturation=0, hue=0, p=1.0),
             A.AdvancedBlur(blur_limit=(3, 7), sigmaX_limit=(0.2, 1.0), sigmaY_limit=(0.2, 1.0), rotate_limit=90,
                            beta_limit=(0.5, 8.0), noise_limit=(0.9, 1.1), p=1.0),
             A.Downscale(scale_min=0.5, scale_max=0.75, interpolation=cv2.INTER_LINEAR, p=1.0),
             A.CoarseDropout(max_holes=8, max_height=8, max_width=8, min_holes=4, min_height=4, min_width=4, fill_value=0, p=1.0),
             A.GridDropout(ratio=0.1, unit_size_min=4, unit_size_max=8, random_offset=True, p=0.5),
             A.ElasticTransform(alpha=20, sigma=4, p=1.0)]
    for t in cands:
        t2 = copy.deepcopy(t)
        t2.p = 0.7
        res = A.Compose([t2])(image=img, mask=msk, force_apply=True) if False else t2(image=img, mask=msk, force_apply=True)
        assert res["image"].shape == img.shape and res["image"].dtype == np.uint8, type(t).__name__
        if not t2.geometric:
            assert np.array_equal(res["mask"], msk), type(t).__name__
    assert np.array_equal(A.InvertImg(p=1.0)(image=img, mask=msk)["image"], 255 - img)
    sol = A.Solarize(threshold=128, p=1.0)(image=img, mask=msk)["image"]
    assert np.array_equal(sol, np.where(img >= 128, 255 - img, img))
    crop = A.RandomCrop(height=96, width=96, p=1.0)(image=img, mask=msk)
    assert crop["image"].shape == (96, 96) and crop["mask"].shape == (96, 96)
    cp = A.CropAndPad(percent=(-0.1, 0.1), pad_mode=cv2.BORDER_REFLECT, p=1.0)(image=img, mask=msk)
    assert cp["image"].shape == img.shape and set(np.unique(cp["mask"])) <= {0, 255}
    pipe = A.Compose([A.RandomRotate90(p=0.5), A.VerticalFlip(p=0.5)])
    random.seed(7)
    r1 = pipe(image=img, mask=msk)
    random.seed(7)
    r2 = pipe(image=img, mask=msk)
    assert np.array_equal(r1["image"], r2["image"]) and np.array_equal(r1["mask"], r2["mask"])
    assert A.Compose([])(image=img, mask=msk)["image"] is not None


def test_cell_augmenter_keeps_copies_on_the_device(cuda, tmp_path, monkeypatch):
    import dataset as ds
    rng = np.random.default_rng(6)
    monkeypatch.chdir(tmp_path)
    pi, pm = [], []
    for i in range(3):
        img, msk = _pair(rng, 96, 96)
        a, b = str(tmp_path / f"c_{i}.tif"), str(tmp_path / f"c_{i}_GT.tif")
        assert cv2.imwrite(a, img) and cv2.imwrite(b, msk)
        pi.append(a)
        pm.append(b)
    writes = []
    real_write = cv2.imwrite
    monkeypatch.setattr(cv2, "imwrite", lambda *a, **k: writes.append(a[0]) or real_write(*a, **k))
    aug = ds.CellAugmenter(2, seed=5)
    images, masks = aug.augment_training_data(pi, pm)
    assert images[:3] == pi and len(images) == 9 and len(masks) == 9 and not writes
    assert all(p.startswith("b2seg-mem://") for p in images[3:])
    data = ds.CellSegmentationDataset(images, masks, img_size=(64, 64))
    x, y = data[5]
    assert x.shape == (1, 64, 64) and float(x.min()) >= 0 and float(x.max()) <= 1 and set(y.unique().tolist()) <= {0.0, 1.0}
    aug.cleanup()
    with pytest.raises(FileNotFoundError):
        ds.CellSegmentationDataset(images, masks, img_size=(64, 64))[4]
    # same seed, same copies
    a1 = ds.CellAugmenter(1, seed=9)
    i1, _ = a1.augment_training_data(pi, pm)
    p1 = [ds._read_plane(p).clone() for p in i1[3:]]
    a2 = ds.CellAugmenter(1, seed=9)
    i2, _ = a2.augment_training_data(pi, pm)
    assert all(torch.equal(a, ds._read_plane(p)) for a, p in zip(p1, i2[3:]))
    a1.cleanup(); a2.cleanup()
