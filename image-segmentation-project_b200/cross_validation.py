"""K-fold cross-validation and model comparison -- API mirror of the reference
`cross_validation.py` (CrossValidator :20-298, ModelComparator :301-415, quick_* :419-447).

Same classes, constructor arguments, method names and returned dict structures, so
`loss_function_cross_validation.py`, `model_configuration_experiment.py` and
`augmentation_amount_experiment.py` import it unchanged.  The difference is the execution model:
the reference walks `for config: for fold:` sequentially on one device; here every (config, fold)
pair is an independent *unit* handed to `b2seg.shard.run_units`, which runs one unit per GPU when
the script is launched with one process per GPU (torchrun) and gathers the metric dicts on the
host.  Folds come from the same seeded `KFold` (reference :63), so the sharded run trains exactly
the units the sequential run would, and the merged result has the reference's layout.
"""
import time
from collections import defaultdict
from typing import Callable, Dict, List, Tuple

import numpy as np
import torch
from sklearn.model_selection import KFold, train_test_split

from b2seg import shard
from b2seg.optim import FusedAdam
from dataset import load_original_data
from losses import get_loss_function
from train import train_model
from utils import get_device


def _summarise(fold_results: List[Dict]) -> Dict:
    """{metric}_mean / {metric}_std over folds, `training_time` left out (reference :181-184)."""
    per_metric = defaultdict(list)
    for res in fold_results:
        for k, v in res.items():
            if k != 'training_time':
                per_metric[k].append(v)
    out = {}
    for k, vals in per_metric.items():
        out[f'{k}_mean'] = np.mean(vals)
        out[f'{k}_std'] = np.std(vals)
    return out


def _unit_cost(config: Dict, n_train: int) -> float:
    """Relative run time of a unit for load balancing: epochs x images x backbone weight."""
    heavy = 4.5 if str(config.get('backbone', '')).endswith('50') else 1.0
    return float(config.get('num_epochs', 1)) * max(n_train, 1) * heavy


class CrossValidator:
    """K-fold CV over all images of one type (no held-out test set)."""

    def __init__(self, data_dir: str = "manual_labels", image_type: str = 'W', n_splits: int = 5,
                 random_state: int = 42, augmentations_per_image: int = 3, verbose: bool = True):
        self.n_splits = n_splits
        self.random_state = random_state
        self.augmentations_per_image = augmentations_per_image
        # under torchrun (RANK / WORLD_SIZE / LOCAL_RANK set) this joins the rendezvous and binds the
        # process to cuda:LOCAL_RANK; otherwise (0, 1) and the loops below run sequentially
        self.rank, self.world_size = shard.init_from_env()
        self._announce = verbose              # every rank reports the units it ran
        self.verbose = verbose and self.rank == 0
        self.data = load_original_data(data_dir, image_type)
        self.image_paths = self.data['image_paths']
        self.mask_paths = self.data['mask_paths']
        if self.verbose:
            print(f"Loaded {len(self.image_paths)} {image_type} images for CV")

    # ---- folds -------------------------------------------------------------------------
    def create_cv_folds(self, indices: List[int] = None) -> List[Tuple[List[int], List[int]]]:
        """[(train_indices, val_indices)] x n_splits from a shuffled, seeded KFold over `indices`."""
        if indices is None:
            indices = list(range(len(self.image_paths)))
        splitter = KFold(n_splits=self.n_splits, shuffle=True, random_state=self.random_state)
        folds = [([indices[i] for i in tr], [indices[i] for i in va]) for tr, va in splitter.split(indices)]
        if self.verbose:
            print(f"Created {self.n_splits} CV folds from {len(indices)} samples")
        return folds

    def _paths(self, idx: List[int]):
        return [self.image_paths[i] for i in idx], [self.mask_paths[i] for i in idx]

    # ---- one unit ----------------------------------------------------------------------
    def _fit(self, model_class: Callable, config: Dict, train_images, train_masks,
             val_images, val_masks) -> Dict:
        """Build model / loss / Adam / plateau scheduler as the reference does (:80-101) and run
        `train_model` quietly; returns its full result dict."""
        device = get_device()
        model = model_class(n_classes=1, backbone=config['backbone'], pretrained=config['pretrained'],
                            use_attention=config['use_attention']).to(device)
        criterion = get_loss_function(config)
        optimizer = FusedAdam(model.parameters(), lr=config['learning_rate'],  # a torch.optim.Adam
                              weight_decay=config.get('weight_decay', 1e-5))
        scheduler = torch.optim.lr_scheduler.ReduceLROnPlateau(optimizer, mode='max', factor=0.5,
                                                               patience=3, min_lr=1e-6)
        quiet = dict(config, verbose=False, save_plots=False)
        return train_model(model=model, train_images=train_images, train_masks=train_masks,
                           val_images=val_images, val_masks=val_masks, criterion=criterion,
                           optimizer=optimizer, scheduler=scheduler, num_epochs=config['num_epochs'],
                           device=device, config=quiet,
                           augmentations_per_image=self.augmentations_per_image, save_plots=False)

    def train_single_model(self, model_class: Callable, config: Dict, train_images: List[str],
                           train_masks: List[str], val_images: List[str], val_masks: List[str]) -> Dict:
        """Final validation metrics of one training run (reference :76-122)."""
        return self._fit(model_class, config, train_images, train_masks, val_images, val_masks)['final_val_metrics']

    def _run_unit(self, unit) -> Dict:
        model_class, config, fold_idx, (tr, va) = unit
        t0 = time.time()
        tr_i, tr_m = self._paths(tr)
        va_i, va_m = self._paths(va)
        metrics = dict(self.train_single_model(model_class, config, tr_i, tr_m, va_i, va_m))
        metrics['training_time'] = time.time() - t0
        if self._announce:
            print(f"  [gpu {self.rank}] {config.get('name', 'Model')} fold {fold_idx + 1}/{self.n_splits}: "
                  f"IoU: {metrics['iou']:.4f} ({metrics['training_time']:.1f}s)", flush=True)
        return metrics

    def _run_grid(self, model_configs: List[Tuple], folds) -> List[List[Dict]]:
        """All (config, fold) units, sharded; returns fold_results per config in config order."""
        units = [(mc, cfg, f, folds[f]) for mc, cfg in model_configs for f in range(len(folds))]
        costs = [_unit_cost(u[1], len(u[3][0])) for u in units]
        flat = shard.run_units(units, self._run_unit, costs if self.world_size > 1 else None)
        k = len(folds)
        return [flat[c * k:(c + 1) * k] for c in range(len(model_configs))]

    # ---- public loops ------------------------------------------------------------------
    def cross_validate_single_model(self, model_class: Callable, config: Dict,
                                    indices: List[int] = None) -> Dict:
        """{'config', 'fold_results', 'cv_summary'} for one configuration (reference :124-199)."""
        if self.verbose:
            print(f"\nCross-validating {config.get('name', 'Model')}...")
        folds = self.create_cv_folds(indices)
        fold_results = self._run_grid([(model_class, config)], folds)[0]
        summary = _summarise(fold_results)
        if self.verbose:
            print(f"  Overall: {summary['iou_mean']:.4f} ± {summary['iou_std']:.4f}")
        return {'config': config, 'fold_results': fold_results, 'cv_summary': summary}

    def compare_multiple_models(self, model_configs: List[Tuple], indices: List[int] = None) -> Dict:
        """Every configuration on the SAME folds (reference :201-298)."""
        if self.verbose:
            print(f"\nComparing {len(model_configs)} models with {self.n_splits}-fold CV...")
        folds = self.create_cv_folds(indices)
        per_config = self._run_grid(model_configs, folds)
        individual, comparison = {}, {}
        for (_, config), fold_results in zip(model_configs, per_config):
            name = config.get('name', 'Unknown')
            summary = _summarise(fold_results)
            individual[name] = {'config': config, 'fold_results': fold_results, 'cv_summary': summary}
            comparison[name] = summary
            if self.verbose:
                print(f"  {name}: {summary['iou_mean']:.4f} ± {summary['iou_std']:.4f}")
        if self.verbose:
            print(f"\n{'=' * 60}\nMODEL COMPARISON SUMMARY\n{'=' * 60}")
            ranked = sorted(comparison.items(), key=lambda kv: kv[1]['iou_mean'], reverse=True)
            for i, (name, s) in enumerate(ranked):
                print(f"{i + 1:2d}. {name:25}: {s['iou_mean']:.4f} ± {s['iou_std']:.4f}")
        return {'individual_results': individual, 'comparison_summary': comparison,
                'cv_folds_used': len(folds)}


class ModelComparator:
    """CV on a training split plus a generalisation run on the held-out split."""

    def __init__(self, data_dir: str = "manual_labels", image_type: str = 'W', test_size: float = 0.2,
                 n_splits: int = 5, random_state: int = 42, augmentations_per_image: int = 3,
                 verbose: bool = True):
        self.cv = CrossValidator(data_dir, image_type, n_splits, random_state,
                                 augmentations_per_image, verbose)
        self.test_size = test_size
        self.random_state = random_state
        self.verbose = self.cv.verbose
        everything = list(range(len(self.cv.image_paths)))
        self.train_indices, self.test_indices = train_test_split(everything, test_size=test_size,
                                                                 random_state=random_state)
        if self.verbose:
            print(f"Created train/test split: {len(self.train_indices)}/{len(self.test_indices)}")

    def run_cv_comparison(self, model_configs: List[Tuple]) -> Dict:
        return self.cv.compare_multiple_models(model_configs, self.train_indices)

    def _generalisation_unit(self, unit) -> Dict:
        model_class, config = unit
        tr_i, tr_m = self.cv._paths(self.train_indices)
        te_i, te_m = self.cv._paths(self.test_indices)
        t0 = time.time()
        res = self.cv._fit(model_class, config, tr_i, tr_m, te_i, te_m)
        return {'final_train_iou': res['best_iou'], 'test_metrics': dict(res['final_val_metrics']),
                'training_time': time.time() - t0}

    def evaluate_generalization(self, model_configs: List[Tuple]) -> Dict:
        """Train each configuration on the whole training split, score on the test split
        (reference :328-415); one configuration per GPU."""
        if self.verbose:
            print(f"\n{'=' * 60}\nGENERALIZATION EVALUATION\n{'=' * 60}")
        units = list(model_configs)
        costs = [_unit_cost(cfg, len(self.train_indices)) for _, cfg in units]
        outs = shard.run_units(units, self._generalisation_unit, costs if self.cv.world_size > 1 else None)
        results = {cfg.get('name', 'Unknown'): out for (_, cfg), out in zip(units, outs)}
        if self.verbose:
            for name, out in results.items():
                print(f"  {name}: Test IoU: {out['test_metrics']['iou']:.4f}")
            print(f"\n{'=' * 60}\nGENERALIZATION SUMMARY\n{'=' * 60}")
            ranked = sorted(results.items(), key=lambda kv: kv[1]['test_metrics']['iou'], reverse=True)
            for i, (name, out) in enumerate(ranked):
                gap = out['final_train_iou'] - out['test_metrics']['iou']
                print(f"{i + 1:2d}. {name:25}: Test IoU = {out['test_metrics']['iou']:.4f}, "
                      f"Overfitting = {gap:.4f}")
        return results


def quick_cv(model_class: Callable, config: Dict, data_dir: str = "manual_labels", image_type: str = 'W',
             n_splits: int = 5, augmentations_per_image: int = 3) -> Dict:
    cv = CrossValidator(data_dir=data_dir, image_type=image_type, n_splits=n_splits,
                        augmentations_per_image=augmentations_per_image)
    return cv.cross_validate_single_model(model_class, config)


def quick_model_comparison(model_configs: List[Tuple], data_dir: str = "manual_labels", image_type: str = 'W',
                           n_splits: int = 5, augmentations_per_image: int = 3,
                           include_generalization: bool = False) -> Dict:
    if include_generalization:
        comparator = ModelComparator(data_dir=data_dir, image_type=image_type, n_splits=n_splits,
                                     augmentations_per_image=augmentations_per_image)
        return {'cv_results': comparator.run_cv_comparison(model_configs),
                'generalization_results': comparator.evaluate_generalization(model_configs)}
    cv = CrossValidator(data_dir=data_dir, image_type=image_type, n_splits=n_splits,
                        augmentations_per_image=augmentations_per_image)
    return cv.compare_multiple_models(model_configs)


if __name__ == "__main__":
    print("cross_validation.py: sharded K-fold cross-validation over the B200 training step")
