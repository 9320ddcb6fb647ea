"""The import surface a reference user keeps: `albumentations` names used by
forward_selection_integration.py:62-181 and dataset.py:88-124, and `cell_segmentation.X` aliases
(main.py:5, loss_function_cross_validation.py:4-5).  No GPU work here."""
import copy
import importlib
import random

import pytest


def test_albumentations_surface():
    import albumentations as A
    assert "image-segmentation-project_b200" in A.__file__
    names = ["Compose", "HorizontalFlip", "VerticalFlip", "Transpose", "RandomRotate90", "Affine", "GridDistortion",
             "OpticalDistortion", "ElasticTransform", "RandomCrop", "CropAndPad", "InvertImg", "Solarize", "RandomGamma",
             "ColorJitter", "GaussNoise", "CoarseDropout", "GridDropout", "AdvancedBlur", "Downscale"]
    for n in names:
        assert hasattr(A, n), n
    t = A.Affine(scale=(0.9, 1.1), translate_percent=(-0.1, 0.1), rotate=(-30, 30), shear=(-10, 10), p=0.5)
    t2 = copy.deepcopy(t)
    t2.p = 0.7
    assert t.p == 0.5 and t2.p == 0.7
    random.seed(1)
    a = t.get_params(64, 64)
    random.seed(1)
    b = t.get_params(64, 64)
    assert a == b and len(a["matrix"]) == 2 and len(a["matrix"][0]) == 3
    pipe = A.Compose([A.HorizontalFlip(p=0.5), t2])
    assert len(pipe.transforms) == 2 and pipe.transforms[1] is t2
    with pytest.raises(KeyError):
        A.HorizontalFlip(p=1.0)(None)  # albumentations: data goes in as named arguments


def test_cell_segmentation_aliases():
    import cell_segmentation
    import advanced_models
    import cross_validation
    import losses
    assert importlib.import_module("cell_segmentation.cross_validation") is cross_validation
    assert importlib.import_module("cell_segmentation.advanced_models") is advanced_models
    from cell_segmentation.losses import get_loss_function
    assert get_loss_function is losses.get_loss_function
    from cell_segmentation.advanced_models import UNetWithBackbone
    assert UNetWithBackbone is advanced_models.UNetWithBackbone
    with pytest.raises(ImportError):
        importlib.import_module("cell_segmentation.no_such_module")


def test_product_modules_do_not_import_the_oracle():
    """The oracle is test infrastructure: nothing under the package may import it."""
    import os
    import re
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "image-segmentation-project_b200")
    bad = []
    for d, _, files in os.walk(root):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(d, f)).read()
                if re.search(r"^\s*(from|import)\s+oracle\b", src, re.M):
                    bad.append(os.path.join(d, f))
    assert not bad, bad
