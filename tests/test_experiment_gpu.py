"""End to end through the reference's driver API on a synthetic data tree (config 2's path):
load_original_data -> CellAugmenter (GPU gathers) -> prepare_data -> train_model (planned U-Net
step, fused loss, FusedAdam) -> evaluate -> CrossValidator result structure."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _make_tree(root, n, size=96):
    cv2 = pytest.importorskip("cv2")
    from b2seg import synth
    imgs, masks = synth.cell_tiles(n, size, seed=21)
    os.makedirs(os.path.join(root, "Labelled_images"))
    os.makedirs(os.path.join(root, "GT_masks"))
    for i in range(n):
        stem = f"cell_{i:03d}_syn_1W_0"
        cv2.imwrite(os.path.join(root, "Labelled_images", stem + ".tif"),
                    (imgs[i, 0].numpy() * 255).astype(np.uint8))
        cv2.imwrite(os.path.join(root, "GT_masks", stem + "_GT.tif"),
                    (masks[i, 0].numpy() * 255).astype(np.uint8))
    # a file of the other image type must be ignored by image_type='W'
    cv2.imwrite(os.path.join(root, "GT_masks", "cell_999_syn_1B_0_GT.tif"), np.zeros((size, size), np.uint8))


def test_cross_validation_end_to_end(cuda, tmp_path, monkeypatch):
    import advanced_models
    import cross_validation as cvm
    monkeypatch.chdir(tmp_path)  # the augmenter writes ./temp_augmentation
    _make_tree(str(tmp_path / "data"), 10)
    config = {"name": "r34_dice", "backbone": "resnet34", "pretrained": False, "use_attention": False,
              "loss_fn": "dice", "learning_rate": 1e-3, "weight_decay": 1e-5, "num_epochs": 3,
              "batch_size": 4, "img_size": (64, 64)}
    torch.manual_seed(0)
    cv = cvm.CrossValidator(str(tmp_path / "data"), "W", n_splits=2, augmentations_per_image=1, verbose=False)
    assert len(cv.image_paths) == 10
    res = cv.cross_validate_single_model(advanced_models.UNetWithBackbone, config)
    assert res["config"] is config and len(res["fold_results"]) == 2
    for fr in res["fold_results"]:
        assert set(fr) == {"precision", "recall", "f1", "iou", "accuracy", "loss", "training_time"}
        assert all(np.isfinite(v) for v in fr.values())
        assert 0.0 <= fr["iou"] <= 1.0 and 0.0 < fr["loss"] < 1.0
    assert set(res["cv_summary"]) == {f"{m}_{s}" for m in ("precision", "recall", "f1", "iou", "accuracy", "loss")
                                      for s in ("mean", "std")}
    assert not os.path.exists("temp_augmentation")  # CellAugmenter.cleanup ran


def test_training_reduces_loss_and_learns_masks(cuda):
    """30 optimisation steps on synthetic cell tiles through train.train_step: the BCE+Dice loss
    falls and the IoU of the fused confusion counts rises (the training step is wired end to end)."""
    import advanced_models
    import losses
    import train as train_mod
    from b2seg import synth, lossfn
    from b2seg.optim import FusedAdam
    torch.manual_seed(1)
    model = advanced_models.UNetWithBackbone(1, "resnet34", pretrained=False, use_attention=True).to(cuda).train()
    crit = losses.get_loss_function({"loss_fn": "combo"})
    opt = FusedAdam(model.parameters(), lr=1e-3, weight_decay=1e-5)
    imgs, masks = synth.cell_tiles(8, 64, seed=3)
    imgs, masks = imgs.to(cuda), masks.to(cuda)
    hist = []
    for _ in range(30):
        loss, stats = train_mod.train_step(model, imgs, masks, opt, crit)
        hist.append((float(loss), lossfn.metrics_dict(stats.cpu())["iou"]))
    assert all(np.isfinite(l) for l, _ in hist)
    assert hist[-1][0] < 0.9 * hist[0][0], hist[::5]
    assert hist[-1][1] > max(0.5, hist[0][1] + 0.3), hist[::5]
