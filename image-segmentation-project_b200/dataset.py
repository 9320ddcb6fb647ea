"""Dataset, path discovery, loaders and the offline augmenter -- API mirror of the reference
`dataset.py` (CellSegmentationDataset :17-66, load_original_data :69-118, prepare_data :121-138,
CellAugmenter :140-210).

Where the work happens:
  * files are decoded ONCE per process (cv2.imread on the host; decoded planes are cached);
  * the per-sample preprocessing of `__getitem__` (INTER_AREA / INTER_NEAREST resize, 2..98
    percentile clip, CLAHE(2.0, 8x8), min-max) runs on the GPU for the whole dataset at first
    access (b2seg.preproc -> csrc/preproc.cu, bit-exact vs cv2); `__getitem__` then returns views
    of device-resident fp32 tensors, so an epoch touches neither the disk nor the host;
  * `CellAugmenter` runs its pipeline (`albumentations.Compose` of this package: GPU gather
    kernels) on the device and keeps the augmented copies IN MEMORY: the "paths" it returns for
    them are `b2seg-mem://` names that `CellSegmentationDataset` resolves to device planes -- no
    `.tif` round trip (reference dataset.py:186-193).  `B2SEG_AUG_TO_DISK=1` restores the files.
"""
import itertools
import os
import random
import shutil

import albumentations as A
import cv2
import numpy as np
import torch
from torch.utils.data import Dataset

from b2seg import _lib as L
from b2seg import preproc

_MEM_PREFIX = "b2seg-mem://"
_MEM_PLANES = {}      # virtual path -> uint8 CUDA plane [H,W]
_DECODED = {}         # real path -> uint8 numpy plane (decoded once per process)
_mem_ids = itertools.count()


def _read_plane(path):
    """uint8 plane of `path`: a device tensor for in-memory augmented copies, else a numpy array."""
    if path.startswith(_MEM_PREFIX):
        plane = _MEM_PLANES.get(path)
        if plane is None:
            raise FileNotFoundError(f"{path}: the augmenter that owned this copy was cleaned up")
        return plane
    img = _DECODED.get(path)
    if img is None:
        img = cv2.imread(path, cv2.IMREAD_GRAYSCALE)
        if img is None:
            raise FileNotFoundError(f"cannot read {path}")
        if len(_DECODED) > int(os.environ.get("B2SEG_DECODE_CACHE", "20000")):
            _DECODED.clear()
        _DECODED[path] = img
    return img


class CellSegmentationDataset(Dataset):
    """(image [1,H,W] fp32 in [0,1], mask [1,H,W] fp32 in {0,1}) pairs from file paths.  Samples
    are CUDA tensors: the whole dataset is preprocessed on the GPU at first access."""

    def __init__(self, image_paths, mask_paths, img_size=(256, 256), normalize=True):
        self.image_paths = image_paths
        self.mask_paths = mask_paths
        self.img_size = img_size
        self.normalize = normalize
        self._images = self._masks = None

    def __len__(self):
        return len(self.image_paths)

    def normalize_microscopy_image(self, image):
        """2..98 percentile clip, CLAHE(2.0, 8x8) on uint8, min-max to [0,1] of ONE uint8 plane
        (numpy in, numpy float32 out; runs on the GPU)."""
        dev = torch.device("cuda", torch.cuda.current_device())
        plane = torch.from_numpy(np.ascontiguousarray(image, dtype=np.uint8)).to(dev)
        return preproc.normalize_microscopy(plane.unsqueeze(0))[0].cpu().numpy()

    def materialize(self, device=None):
        """Decode (first use only), upload and preprocess every sample; stacks of equal source
        size go through the kernels together."""
        if self._images is not None:
            return self
        if not torch.cuda.is_available():
            raise L.B2SegError("CellSegmentationDataset preprocesses on a Blackwell GPU (no CPU fallback)")
        dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        n = len(self.image_paths)
        dw, dh = self.img_size
        images = torch.empty((n, 1, dh, dw), dtype=torch.float32, device=dev)
        masks = torch.empty((n, 1, dh, dw), dtype=torch.float32, device=dev)
        groups = {}
        for i, (ip, mp) in enumerate(zip(self.image_paths, self.mask_paths)):
            im, mk = _read_plane(ip), _read_plane(mp)
            groups.setdefault((tuple(im.shape), tuple(mk.shape)), []).append((i, im, mk))

        def stack(planes):
            if torch.is_tensor(planes[0]):
                return torch.stack([p.to(dev) for p in planes])
            return torch.from_numpy(np.stack(planes)).to(dev)

        for items in groups.values():
            idx = torch.tensor([i for i, _, _ in items], device=dev)
            for lo in range(0, len(items), 1024):
                part = items[lo:lo + 1024]
                # a group can mix file-backed (numpy) and in-memory (device) planes
                im = torch.stack([p if torch.is_tensor(p) else torch.from_numpy(p).to(dev) for _, p, _ in part])
                mk = torch.stack([p if torch.is_tensor(p) else torch.from_numpy(p).to(dev) for _, _, p in part])
                fi, fm = preproc.dataset_items(im, mk, self.img_size, self.normalize)
                images[idx[lo:lo + 1024]] = fi
                masks[idx[lo:lo + 1024]] = fm
        self._images, self._masks = images, masks
        return self

    def __getitem__(self, idx):
        self.materialize()
        return self._images[idx], self._masks[idx]


class DeviceBatchLoader:
    """What `prepare_data` returns: iterates (images [B,1,H,W], masks [B,1,H,W]) batches of a
    materialised CellSegmentationDataset straight from device memory.  Same surface the training
    loops use of torch's DataLoader (`len`, iteration, `.dataset`, `.batch_size`); like it, the last
    batch may be smaller (no drop_last, reference dataset.py:138) and `shuffle=True` draws a fresh
    permutation from torch's global RNG every epoch."""

    def __init__(self, dataset, batch_size=2, shuffle=True):
        self.dataset, self.batch_size, self.shuffle = dataset, int(batch_size), bool(shuffle)

    def __len__(self):
        return (len(self.dataset) + self.batch_size - 1) // self.batch_size

    def __iter__(self):
        ds = self.dataset.materialize()
        n = len(ds)
        order = torch.randperm(n) if self.shuffle else torch.arange(n)
        order = order.to(ds._images.device)
        for lo in range(0, n, self.batch_size):
            sel = order[lo:lo + self.batch_size]
            yield ds._images[sel], ds._masks[sel]


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
    """Loader over CellSegmentationDataset (reference dataset.py:121-138): device-resident batches."""
    dataset = CellSegmentationDataset(image_paths, mask_paths, img_size=img_size)
    return DeviceBatchLoader(dataset, batch_size=batch_size, shuffle=shuffle)


class CellAugmenter:
    """Offline augmenter with the reference's pipeline (dataset.py:148-154):
    RandomRotate90(p=.5) -> Affine(scale .95-1.05, translate +-5 %, rotate +-15 deg, shear +-5 deg,
    p=.3; bilinear image / nearest mask, constant-0 border) -> VerticalFlip(p=.5) ->
    AdvancedBlur(3..7, p=.3), built from this package's `albumentations` (GPU kernels).
    Parameters are drawn from Python's `random` module, so `random.seed(n)` in an experiment
    script makes the augmented copies repeatable; `seed=` seeds it here."""

    def __init__(self, augmentations_per_image=3, seed=None):
        self.augmentations_per_image = augmentations_per_image
        # one directory per process: folds / grid configs run one per GPU (b2seg.shard) and each
        # rank writes and cleans up its own augmented copies
        rank = os.environ.get("RANK") if int(os.environ.get("WORLD_SIZE", "1")) > 1 else None
        self.temp_dir = "temp_augmentation" if rank is None else f"temp_augmentation_rank{rank}"
        if seed is not None:
            random.seed(seed)
        self.pipeline = A.Compose([
            A.RandomRotate90(p=0.5),
            A.Affine(scale=(0.95, 1.05), translate_percent=(-0.05, 0.05),
                     rotate=(-15, 15), shear=(-5, 5), p=0.3),
            A.VerticalFlip(p=0.5),
            A.AdvancedBlur(blur_limit=(3, 7), p=0.3),
        ])
        self._mem = []
        self._to_disk = os.environ.get("B2SEG_AUG_TO_DISK", "0") == "1"

    def augment_training_data(self, train_images, train_masks):
        """Returns (images, masks): the originals first, then `augmentations_per_image` augmented
        copies per readable pair.  The copies stay on the device under `b2seg-mem://` names (or are
        written as .tif under `temp_dir` with B2SEG_AUG_TO_DISK=1, as the reference does)."""
        if self.augmentations_per_image == 0:
            return train_images, train_masks
        dev = torch.device("cuda", torch.cuda.current_device())
        if self._to_disk:
            os.makedirs(self.temp_dir, exist_ok=True)
        tag = next(_mem_ids)
        all_images, all_masks = list(train_images), list(train_masks)
        for img_path, mask_path in zip(train_images, train_masks):
            try:
                image, mask = _read_plane(img_path), _read_plane(mask_path)
            except FileNotFoundError:
                print(f"Warning: Could not load {img_path} or {mask_path}")
                continue
            to_dev = lambda p: p.to(dev) if torch.is_tensor(p) else torch.from_numpy(p).to(dev)
            image, mask = to_dev(image), to_dev(mask)
            mask = (mask > 0).to(torch.uint8) * 255
            pair = torch.stack([image, mask])
            stem = os.path.splitext(os.path.basename(img_path))[0]
            for i in range(self.augmentations_per_image):
                out = self.pipeline.apply_planes(pair)
                if self._to_disk:
                    a_img = os.path.join(self.temp_dir, f"{stem}_aug{i}.tif")
                    a_msk = os.path.join(self.temp_dir, f"{stem}_mask{i}.tif")
                    host = out.cpu().numpy()
                    if not (cv2.imwrite(a_img, host[0]) and cv2.imwrite(a_msk, host[1])):
                        print(f"Warning: Failed to save augmented files for {stem}")
                        continue
                else:
                    a_img = f"{_MEM_PREFIX}{tag}/{self.temp_dir}/{stem}_aug{i}.tif"
                    a_msk = f"{_MEM_PREFIX}{tag}/{self.temp_dir}/{stem}_mask{i}.tif"
                    _MEM_PLANES[a_img], _MEM_PLANES[a_msk] = out[0], out[1]
                    self._mem += [a_img, a_msk]
                all_images.append(a_img)
                all_masks.append(a_msk)
        print(f"Training data: {len(train_images)} original + {len(all_images) - len(train_images)} "
              f"augmented = {len(all_images)} total")
        return all_images, all_masks

    def cleanup(self):
        for name in self._mem:
            _MEM_PLANES.pop(name, None)
        self._mem = []
        if os.path.exists(self.temp_dir):
            shutil.rmtree(self.temp_dir)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.cleanup()
        return False
