"""CUDA augmentation gathers vs cv2 (the arithmetic the reference's albumentations calls bottom
out in) and vs the numpy oracle: bit-exact, including the bilinear warps."""
import numpy as np
import pytest
import torch

from oracle import augment_cv as A

pytestmark = pytest.mark.gpu


def _img(rng, h, w):
    cv2 = pytest.importorskip("cv2")
    return cv2.GaussianBlur(rng.integers(0, 256, (h, w), dtype=np.uint8), (5, 5), 1.5)


@pytest.mark.parametrize("h,w", [(1, 1), (5, 37), (64, 64), (97, 131), (256, 256), (192, 384), (128, 64), (1024, 1024)])
def test_permutations_bit_exact(cuda, h, w):
    from b2seg import augment
    rng = np.random.default_rng(h * 1000 + w)
    planes = rng.integers(0, 256, (2, h, w), dtype=np.uint8)
    d = torch.from_numpy(planes).to(cuda)
    assert np.array_equal(augment.hflip(d).cpu().numpy(), A.hflip(planes))
    assert np.array_equal(augment.vflip(d).cpu().numpy(), A.vflip(planes))
    assert np.array_equal(augment.transpose(d).cpu().numpy(), A.transpose(planes))
    for k in range(-1, 5):
        assert np.array_equal(augment.rot90(d, k).cpu().numpy(), A.rot90(planes, k)), k
    # idempotence / inverse properties
    assert torch.equal(augment.hflip(augment.hflip(d)), d)
    assert torch.equal(augment.rot90(augment.rot90(d, 1), 3), d)
    assert torch.equal(augment.transpose(augment.transpose(d)), d)


@pytest.mark.parametrize("h,w", [(64, 64), (97, 131), (256, 256), (1024, 1024)])
def test_warp_affine_matches_cv2(cuda, h, w):
    cv2 = pytest.importorskip("cv2")
    from b2seg import augment
    rng = np.random.default_rng(h + w)
    img = _img(rng, h, w)
    mask = ((rng.random((h, w)) > 0.8) * 255).astype(np.uint8)
    d = torch.from_numpy(np.stack([img, mask])).to(cuda)
    flags = {augment.LINEAR: cv2.INTER_LINEAR, augment.NEAREST: cv2.INTER_NEAREST}
    borders = {augment.CONSTANT: cv2.BORDER_CONSTANT, augment.REFLECT: cv2.BORDER_REFLECT,
               augment.REFLECT101: cv2.BORDER_REFLECT_101}
    for trial in range(4):
        M = augment.affine_matrix(h, w, rng.uniform(0.95, 1.05), (rng.uniform(-.05, .05), rng.uniform(-.05, .05)),
                                  rng.uniform(-15, 15), (rng.uniform(-5, 5), rng.uniform(-5, 5)))
        if trial == 3:  # a wild transform: footprint does not fit the staged tile -> global path
            M = augment.affine_matrix(h, w, 0.2, (0.3, -0.2), 77.0, (20, -10))
        Mn = np.array(M)
        for it, cf in flags.items():
            for bt, cb in borders.items():
                got = augment.warp_affine(d, M, it, bt, 0).cpu().numpy()
                for p in range(2):
                    ref = cv2.warpAffine(d[p].cpu().numpy(), Mn, (w, h), flags=cf, borderMode=cb, borderValue=0)
                    assert np.array_equal(got[p], ref), (trial, it, bt, p, np.abs(got[p].astype(int) - ref).max())
        assert np.array_equal(augment.warp_affine(d, M, augment.LINEAR, augment.CONSTANT, 0).cpu().numpy()[0],
                              A.warp_affine_u8(img, Mn, A.LINEAR, A.CONSTANT, 0))


@pytest.mark.parametrize("h,w", [(64, 64), (97, 131), (256, 256)])
def test_remap_matches_cv2(cuda, h, w):
    cv2 = pytest.importorskip("cv2")
    from b2seg import augment
    rng = np.random.default_rng(h * 3 + w)
    img = _img(rng, h, w)
    d = torch.from_numpy(img[None]).to(cuda)
    maps = [A.elastic_maps(h, w, 8.0, 4.0, rng),
            A.grid_distortion_maps(h, w, 1 + rng.uniform(-.3, .3, 5), 1 + rng.uniform(-.3, .3, 5)),
            A.optical_distortion_maps(h, w, 0.4, 3.0, -2.0)]
    for mx, my in maps:
        mxd, myd = torch.from_numpy(mx).to(cuda), torch.from_numpy(my).to(cuda)
        for it, cf in ((augment.LINEAR, cv2.INTER_LINEAR), (augment.NEAREST, cv2.INTER_NEAREST)):
            for bt, cb in ((augment.CONSTANT, cv2.BORDER_CONSTANT), (augment.REFLECT, cv2.BORDER_REFLECT),
                           (augment.REFLECT101, cv2.BORDER_REFLECT_101)):
                ref = cv2.remap(img, mx, my, cf, borderMode=cb, borderValue=0)
                got = augment.remap(d, mxd, myd, it, bt, 0).cpu().numpy()[0]
                assert np.array_equal(got, ref), (it, bt, np.abs(got.astype(int) - ref).max())


@pytest.mark.parametrize("h,w,planes", [(512, 512, 40), (200, 328, 24), (1024, 1024, 12)])
def test_warp_many_planes_large_rotations(cuda, h, w, planes):
    """A CTA walks several planes with one set of source coordinates (the next plane's footprint is fetched while
    the current one is gathered), and rotations up to 90 degrees give tall / wide footprints."""
    cv2 = pytest.importorskip("cv2")
    from b2seg import augment
    rng = np.random.default_rng(planes)
    host = rng.integers(0, 256, (planes, h, w), dtype=np.uint8)
    d = torch.from_numpy(host).to(cuda)
    for angle, scale in ((8.0, 1.02), (31.0, 0.97), (45.0, 1.05), (90.0, 1.0), (-63.0, 0.9)):
        M = augment.affine_matrix(h, w, scale, (0.01, -0.02), angle, (2.5, -1.5))
        Mn = np.array(M)
        for it, cf in ((augment.LINEAR, cv2.INTER_LINEAR), (augment.NEAREST, cv2.INTER_NEAREST)):
            for bt, cb in ((augment.REFLECT, cv2.BORDER_REFLECT), (augment.CONSTANT, cv2.BORDER_CONSTANT)):
                got = augment.warp_affine(d, M, it, bt, 7).cpu().numpy()
                for p in range(planes):
                    ref = cv2.warpAffine(host[p], Mn, (w, h), flags=cf, borderMode=cb, borderValue=7)
                    assert np.array_equal(got[p], ref), (angle, it, bt, p)
    yy, xx = np.meshgrid(np.arange(h, dtype=np.float32), np.arange(w, dtype=np.float32), indexing="ij")
    mx = (xx + 3.0 * np.sin(yy / 17.0)).astype(np.float32)
    my = (yy + 3.0 * np.cos(xx / 13.0)).astype(np.float32)
    got = augment.remap(d, torch.from_numpy(mx).to(cuda), torch.from_numpy(my).to(cuda)).cpu().numpy()
    for p in range(planes):
        ref = cv2.remap(host[p], mx, my, cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT)
        assert np.array_equal(got[p], ref), p


def test_errors(cuda):
    from b2seg import augment
    with pytest.raises(TypeError):
        augment.hflip(torch.zeros(1, 4, 4, device=cuda))
    with pytest.raises(Exception):
        augment.hflip(torch.zeros(1, 4, 4, dtype=torch.uint8))
    with pytest.raises(ValueError):
        augment.hflip(torch.zeros(0, 4, 4, dtype=torch.uint8, device=cuda))
