"""The reference's OWN experiment scripts on the B200 modules, with only the import path changed:
`image-segmentation-project_b200/` first on sys.path (advanced_models, losses, utils, train, dataset,
cross_validation, albumentations, cell_segmentation), then oracle/_ref (the unmodified reference
files staged by oracle/make_ref.py) for the scripts themselves.  matplotlib / seaborn (plotting,
out of scope, not installed) are stubbed at import.  `pretrained` is switched off in the config
dictionaries passed in (no network for the ImageNet weights); nothing in the scripts is edited."""
import os
import sys
import types

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
cv2 = pytest.importorskip("cv2")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")


@pytest.fixture(scope="module")
def reference_scripts():
    if not os.path.exists(os.path.join(REF, "forward_selection_integration.py")):
        pytest.skip("oracle/_ref not staged (python oracle/make_ref.py in the build container)")
    for name in ("matplotlib", "matplotlib.pyplot", "seaborn"):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__getattr__ = lambda attr: (_ for _ in ()).throw(AttributeError(attr)) if attr.startswith("__") else (lambda *a, **k: None)
            sys.modules[name] = m
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    if REF not in sys.path:
        sys.path.append(REF)  # AFTER the B200 package: only the experiment scripts come from here
    yield REF
    sys.path.remove(REF)


def _tree(root, n=12, size=96):
    from b2seg import synth
    imgs, masks = synth.cell_tiles(n, size, seed=31)
    os.makedirs(os.path.join(root, "Labelled_images"))
    os.makedirs(os.path.join(root, "GT_masks"))
    for i in range(n):
        stem = f"cell_{i:03d}_syn_1W_0"
        cv2.imwrite(os.path.join(root, "Labelled_images", stem + ".tif"), (imgs[i, 0].numpy() * 255).astype(np.uint8))
        cv2.imwrite(os.path.join(root, "GT_masks", stem + "_GT.tif"), (masks[i, 0].numpy() * 255).astype(np.uint8))


def test_forward_selection_script_runs_on_gpu_augment_kernels(cuda, reference_scripts, tmp_path, monkeypatch):
    """BASELINE config 3: the reference's AugmentationSelector, unchanged, evaluating one candidate."""
    monkeypatch.chdir(tmp_path)
    _tree(str(tmp_path / "data"))
    import forward_selection_integration as fsi
    assert os.path.dirname(os.path.abspath(fsi.__file__)) == reference_scripts
    import albumentations
    assert "image-segmentation-project_b200" in albumentations.__file__ and fsi.A is albumentations
    import cell_segmentation
    assert cell_segmentation.forward_selection_integration is fsi  # main.py:5 resolves to the same module
    base = {"backbone": "resnet34", "pretrained": False, "use_attention": False, "loss_fn": "combo",
            "learning_rate": 1e-3, "weight_decay": 1e-5, "num_epochs": 3, "batch_size": 4, "img_size": (64, 64),
            "image_type": "W"}
    sel = fsi.AugmentationSelector(base, str(tmp_path / "data"), cv_folds=2, quick_evaluation=True)
    assert len(sel.augmentation_candidates) == 18 and len(sel.image_paths) == 12
    torch.manual_seed(0)
    score = sel._evaluate_augmentation_set(["horizontal_flip", "affine"])
    assert 0.0 <= score <= 1.0 and np.isfinite(score)
    pipe = sel._create_augmentation_pipeline(["grid_distortion", "optical_distortion", "random_rotate_90", "transpose"])
    assert all(t.p == 0.7 for t in pipe.transforms)
    imgs, msks = sel._create_augmented_dataset(sel.image_paths[:2], sel.mask_paths[:2], pipe, augmentations_per_image=2)
    assert len(imgs) == 6 and all(os.path.exists(p) for p in imgs[2:])
    sel._cleanup_temp_files()


def test_loss_comparison_script_configs_through_model_comparator(cuda, reference_scripts, tmp_path, monkeypatch):
    """BASELINE config 2: loss_function_cross_validation.get_loss_configurations() through ModelComparator."""
    monkeypatch.chdir(tmp_path)
    _tree(str(tmp_path / "data"))
    import loss_function_cross_validation as lcv
    import cross_validation
    assert lcv.ModelComparator is cross_validation.ModelComparator
    assert "image-segmentation-project_b200" in cross_validation.__file__
    configs = [dict(c, pretrained=False, num_epochs=2, img_size=(64, 64), verbose=False)
               for c in lcv.get_loss_configurations() if c["loss_fn"] in ("dice", "focal")]
    assert len(configs) == 2
    comp = lcv.ModelComparator(data_dir=str(tmp_path / "data"), image_type="W", n_splits=2, augmentations_per_image=1,
                               verbose=False)
    res = comp.run_cv_comparison([(lcv.UNetWithBackbone, c) for c in configs])
    assert set(res["comparison_summary"]) == {"Dice Loss", "Focal Loss"} and res["cv_folds_used"] == 2
    for summary in res["comparison_summary"].values():
        assert 0.0 <= summary["iou_mean"] <= 1.0 and np.isfinite(summary["loss_mean"])
