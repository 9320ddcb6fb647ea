"""Dataset, path discovery, loaders and the offline augmenter -- API mirror of the reference
`dataset.py` (CellSegmentationDataset :17-66, load_original_data :69-118, prepare_data :121-138,
CellAugmenter :140-210).

The per-sample preprocessing (INTER_AREA resize, 2/98-percentile clip, CLAHE, min-max) is the
reference's OpenCV arithmetic on the host: it is the step *before* the hot path.
The augmenter's geometric transforms (rot90, affine warp, vertical flip) run on the GPU gather
kernel (b2seg.augment / csrc/augment.cu), image and mask planes in one launch; parameter
sampling mirrors albumentations' parameterisation with numpy's RNG (albumentations itself is an
un-vendored dependency of the reference: dataset.py:12, requirements.txt:7).
"""
import os
import shutil

import cv2
import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset


class CellSegmentationDataset(Dataset):
    """(image [1,H,W] fp32 in [0,1], mask [1,H,W] fp32 in {0,1}) pairs from file paths."""

    def __init__(self, image_paths, mask_paths, img_size=(256, 256), normalize=True):
        self.image_paths = image_paths
        self.mask_paths = mask_paths
        self.img_size = img_size
        self.normalize = normalize

    def __len__(self):
        return len(self.image_paths)

    def normalize_microscopy_image(self, image):
        """2..98 percentile clip, CLAHE(2.0, 8x8) on uint8, min-max to [0,1]."""
        lo, hi = np.percentile(image, [2, 98])
        clipped = np.clip(image, lo, hi).astype(np.uint8)
        eq = cv2.createCLAHE(clipLimit=2.0, tileGridSize=(8, 8)).apply(clipped)
        return (eq - eq.min()) / (eq.max() - eq.min() + 1e-8)

    def __getitem__(self, idx):
        image = cv2.imread(self.image_paths[idx], cv2.IMREAD_GRAYSCALE)
        mask = cv2.imread(self.mask_paths[idx], cv2.IMREAD_GRAYSCALE)
        if image is None or mask is None:
            raise FileNotFoundError(f"cannot read {self.image_paths[idx]} / {self.mask_paths[idx]}")
        image = cv2.resize(image, self.img_size, interpolation=cv2.INTER_AREA)
        mask = cv2.resize(mask, self.img_size, interpolation=cv2.INTER_NEAREST)
        if self.normalize:
            image = self.normalize_microscopy_image(image)
        else:
            image = image.astype(np.float32) / 255.0
        image = torch.from_numpy(np.ascontiguousarray(image, dtype=np.float32)).unsqueeze(0)
        mask = torch.from_numpy((mask > 0).astype(np.float32)).unsqueeze(0)
        return image, mask


def load_original_data(data_dir="manual_labels", image_type='W'):
    """Paths of the original (non-augmented) pairs: masks `<stem>_GT.tif` in GT_masks/, images
    `<stem>.tif` in Labelled_images/; the 4th '_' field of the name carries the type (1B / 1W)."""
    images_dir = os.path.join(data_dir, "Labelled_images")
    masks_dir = os.path.join(data_dir, "GT_masks")
    if not os.path.exists(images_dir) or not os.path.exists(masks_dir):
        raise FileNotFoundError(f"Data directories not found in {data_dir}")
    images = set(os.listdir(images_dir))
    image_paths, mask_paths = [], []
    for mask_file in sorted(os.listdir(masks_dir)):
        if not mask_file.endswith('GT.tif'):
            continue
        if mask_file.split('_')[3][1] != image_type:
            continue
        original = mask_file[:-len('_GT.tif')] + '.tif'
        if original in images:
            image_paths.append(os.path.join(images_dir, original))
            mask_paths.append(os.path.join(masks_dir, mask_file))
    print(f"Loaded {len(image_paths)} original {image_type} images")
    return {'image_paths': image_paths, 'mask_paths': mask_paths}


def prepare_data(image_paths, mask_paths, batch_size=2, img_size=(256, 256), shuffle=True):
    """DataLoader over CellSegmentationDataset; batches are pinned so the H2D copy is async."""
    dataset = CellSegmentationDataset(image_paths, mask_paths, img_size=img_size)
    return DataLoader(dataset, batch_size=batch_size, shuffle=shuffle,
                      pin_memory=torch.cuda.is_available())


class CellAugmenter:
    """Offline augmenter with the reference's pipeline (dataset.py:148-154):
    RandomRotate90(p=.5) -> Affine(scale .95-1.05, translate +-5 %, rotate +-15 deg, shear +-5 deg,
    p=.3; bilinear image / nearest mask, constant-0 border) -> VerticalFlip(p=.5) ->
    AdvancedBlur(3..7, p=.3).  The geometric stages run on the GPU gather kernel."""

    def __init__(self, augmentations_per_image=3, seed=None):
        self.augmentations_per_image = augmentations_per_image
        # one directory per process: folds / grid configs run one per GPU (b2seg.shard) and each
        # rank writes and cleans up its own augmented copies
        rank = os.environ.get("RANK") if int(os.environ.get("WORLD_SIZE", "1")) > 1 else None
        self.temp_dir = "temp_augmentation" if rank is None else f"temp_augmentation_rank{rank}"
        self.rng = np.random.default_rng(seed)

    # -- one augmented (image, mask) pair, uint8 numpy in / out
    def pipeline(self, image, mask):
        from b2seg import augment as ag
        dev = torch.device("cuda", torch.cuda.current_device())
        planes = torch.from_numpy(np.stack([image, mask])).to(dev)
        rng = self.rng
        if rng.random() < 0.5:
            planes = ag.rot90(planes, int(rng.integers(0, 4)))
        if rng.random() < 0.3:
            h, w = planes.shape[1:]
            s = rng.uniform(0.95, 1.05)
            m = ag.affine_matrix(h, w, scale=s,
                                 translate=(rng.uniform(-0.05, 0.05), rng.uniform(-0.05, 0.05)),
                                 rotate_deg=rng.uniform(-15, 15),
                                 shear_deg=(rng.uniform(-5, 5), rng.uniform(-5, 5)))
            img = ag.warp_affine(planes[0:1], m, interp=ag.LINEAR, border=ag.CONSTANT, cval=0)
            msk = ag.warp_affine(planes[1:2], m, interp=ag.NEAREST, border=ag.CONSTANT, cval=0)
            planes = torch.cat([img, msk])
        if rng.random() < 0.5:
            planes = ag.vflip(planes)
        out = planes.cpu().numpy()
        img, msk = out[0], out[1]
        if rng.random() < 0.3:  # AdvancedBlur: non-geometric, stays on the host
            k = int(rng.choice([3, 5, 7]))
            sx, sy = rng.uniform(0.2, 1.0, size=2)
            img = cv2.GaussianBlur(img, (k, k), sigmaX=float(sx), sigmaY=float(sy))
        return {"image": img, "mask": msk}

    def augment_training_data(self, train_images, train_masks):
        """Returns (images, masks): the originals first, then `augmentations_per_image` augmented
        copies per readable pair, written as .tif under `temp_dir`."""
        if self.augmentations_per_image == 0:
            return train_images, train_masks
        os.makedirs(self.temp_dir, exist_ok=True)
        all_images, all_masks = list(train_images), list(train_masks)
        for img_path, mask_path in zip(train_images, train_masks):
            image = cv2.imread(img_path, cv2.IMREAD_GRAYSCALE)
            mask = cv2.imread(mask_path, cv2.IMREAD_GRAYSCALE)
            if image is None or mask is None:
                print(f"Warning: Could not load {img_path} or {mask_path}")
                continue
            mask = (mask > 0).astype(np.uint8) * 255
            stem = os.path.splitext(os.path.basename(img_path))[0]
            for i in range(self.augmentations_per_image):
                res = self.pipeline(image, mask)
                a_img = os.path.join(self.temp_dir, f"{stem}_aug{i}.tif")
                a_msk = os.path.join(self.temp_dir, f"{stem}_mask{i}.tif")
                if cv2.imwrite(a_img, res["image"]) and cv2.imwrite(a_msk, res["mask"]):
                    all_images.append(a_img)
                    all_masks.append(a_msk)
                else:
                    print(f"Warning: Failed to save augmented files for {stem}")
        print(f"Training data: {len(train_images)} original + {len(all_images) - len(train_images)} "
              f"augmented = {len(all_images)} total")
        return all_images, all_masks

    def cleanup(self):
        if os.path.exists(self.temp_dir):
            shutil.rmtree(self.temp_dir)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.cleanup()
        return False
