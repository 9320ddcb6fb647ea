"""GPU input pipeline (csrc/preproc.cu) against OpenCV / numpy themselves: the
arithmetic of CellSegmentationDataset.__getitem__ (reference dataset.py:30-66).  Integer / LUT work:
bit-exact.  The one exception is plain INTER_LINEAR when ENLARGING (used only by the Downscale /
CropAndPad augmentation candidates): cv2 takes a different code path there and differs from its
own documented fixed-point scheme by 1 LSB on < 1 % of the pixels (same for oracle/preproc_cv.py)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
cv2 = pytest.importorskip("cv2")


def _smooth(rng, h, w):
    a = rng.integers(0, 256, (h, w), dtype=np.uint8)
    a = cv2.GaussianBlur(a, (0, 0), 3)
    return cv2.normalize(a, None, 0, 255, cv2.NORM_MINMAX).astype(np.uint8)


SHAPES = [(96, 96, 64, 64), (512, 512, 256, 256), (520, 696, 256, 256), (1024, 1024, 256, 256),
          (300, 300, 128, 128), (100, 130, 256, 256), (256, 256, 256, 256), (97, 131, 64, 64), (64, 64, 96, 96)]


@pytest.mark.parametrize("h,w,dh,dw", SHAPES)
def test_resize_matches_cv2(cuda, h, w, dh, dw):
    from b2seg import preproc
    rng = np.random.default_rng(h * 1000 + w)
    planes = np.stack([_smooth(rng, h, w), rng.integers(0, 256, (h, w), dtype=np.uint8),
                       (rng.random((h, w)) > 0.8).astype(np.uint8) * 255])
    dev = torch.from_numpy(planes).to(cuda)
    for code, flag, tol in ((preproc.AREA, cv2.INTER_AREA, 0), (preproc.NEAREST, cv2.INTER_NEAREST, 0),
                            (preproc.LINEAR, cv2.INTER_LINEAR, 1 if (dh > h or dw > w) else 0)):
        got = preproc.resize(dev, (dw, dh), code).cpu().numpy()
        for p in range(planes.shape[0]):
            ref = cv2.resize(planes[p], (dw, dh), interpolation=flag)
            d = np.abs(ref.astype(int) - got[p].astype(int))
            assert d.max() <= tol, (flag, p, d.max())
            assert (d > 0).mean() <= 0.01
    mk = preproc.resize(dev[2:3], (dw, dh), preproc.NEAREST, binary_f32=True).cpu().numpy()[0]
    assert np.array_equal(mk, (cv2.resize(planes[2], (dw, dh), interpolation=cv2.INTER_NEAREST) > 0).astype(np.float32))


def _reference_normalize(image):
    """dataset.py:30-43, verbatim arithmetic on the host (numpy + cv2)."""
    p_low, p_high = np.percentile(image, [2, 98])
    clipped = np.clip(image, p_low, p_high)
    eq = cv2.createCLAHE(clipLimit=2.0, tileGridSize=(8, 8)).apply(clipped.astype(np.uint8))
    return (eq - eq.min()) / (eq.max() - eq.min() + 1e-8), clipped.astype(np.uint8), eq, (p_low, p_high)


@pytest.mark.parametrize("h,w", [(256, 256), (128, 128), (64, 64), (250, 250), (100, 130)])
def test_normalize_microscopy_matches_reference_arithmetic(cuda, h, w):
    from b2seg import preproc
    rng = np.random.default_rng(h + w)
    imgs = [_smooth(rng, h, w), rng.integers(0, 256, (h, w), dtype=np.uint8),
            (rng.integers(0, 40, (h, w)) + 100).astype(np.uint8), np.full((h, w), 77, np.uint8)]
    out, st = preproc.normalize_microscopy(torch.from_numpy(np.stack(imgs)).to(cuda), return_stages=True)
    out = out.cpu().numpy()
    for i, im in enumerate(imgs):
        ref, clipped, eq, pct = _reference_normalize(im)
        assert tuple(st["percentiles"][i].cpu().numpy()) == pct
        assert np.array_equal(st["clipped"][i].cpu().numpy(), clipped)
        assert np.array_equal(st["clahe"][i].cpu().numpy(), eq)
        assert np.array_equal(out[i], ref.astype(np.float32))


def test_dataset_on_device_equals_host_recipe(cuda, tmp_path, monkeypatch):
    """CellSegmentationDataset (GPU-preprocessed, device-resident) against the reference's
    __getitem__ arithmetic on the host, on a small .tif tree of two source sizes; after the first
    pass nothing is read from disk again."""
    import dataset as ds
    from oracle import preproc_cv
    rng = np.random.default_rng(5)
    paths_i, paths_m = [], []
    for i, (h, w) in enumerate([(96, 96), (96, 96), (150, 120), (96, 96), (150, 120)]):
        im, mk = _smooth(rng, h, w), (rng.random((h, w)) > 0.7).astype(np.uint8) * 255
        pi, pm = str(tmp_path / f"img{i}.tif"), str(tmp_path / f"msk{i}.tif")
        assert cv2.imwrite(pi, im) and cv2.imwrite(pm, mk)
        paths_i.append(pi)
        paths_m.append(pm)
    data = ds.CellSegmentationDataset(paths_i, paths_m, img_size=(64, 64))
    for i in range(len(data)):
        gi, gm = data[i]
        assert gi.is_cuda and gi.shape == (1, 64, 64) and gi.dtype == torch.float32
        im = cv2.imread(paths_i[i], cv2.IMREAD_GRAYSCALE)
        mk = cv2.imread(paths_m[i], cv2.IMREAD_GRAYSCALE)
        im = cv2.resize(im, (64, 64), interpolation=cv2.INTER_AREA)
        mk = cv2.resize(mk, (64, 64), interpolation=cv2.INTER_NEAREST)
        ref = _reference_normalize(im)[0].astype(np.float32)
        assert np.array_equal(gi[0].cpu().numpy(), ref)
        assert np.array_equal(gm[0].cpu().numpy(), (mk > 0).astype(np.float32))
        oi, om = preproc_cv.dataset_item(cv2.imread(paths_i[i], 0), cv2.imread(paths_m[i], 0), (64, 64))
        assert np.array_equal(oi, ref) and np.array_equal(om, (mk > 0).astype(np.float32))
    # second epoch: no file access at all
    calls = []
    monkeypatch.setattr(cv2, "imread", lambda *a, **k: calls.append(a) or None)
    loader = ds.prepare_data(paths_i, paths_m, batch_size=2, img_size=(64, 64), shuffle=True)
    shapes = [tuple(b[0].shape) for b in loader] + [tuple(b[0].shape) for b in loader]
    assert shapes == [(2, 1, 64, 64), (2, 1, 64, 64), (1, 1, 64, 64)] * 2 and len(loader) == 3
    assert not calls
    data2 = ds.CellSegmentationDataset(paths_i, paths_m, img_size=(64, 64), normalize=False)
    im = cv2.resize(ds._read_plane(paths_i[0]), (64, 64), interpolation=cv2.INTER_AREA)
    assert np.array_equal(data2[0][0][0].cpu().numpy(), im.astype(np.float32) / 255.0)
