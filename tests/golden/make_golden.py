"""Generate tests/golden/* by running the REFERENCE's own modules (build container only).

    python tests/golden/make_golden.py            # needs /root/reference

The reference ships no tests or golden vectors, so its own code is run
here on seeded synthetic inputs and the results are committed as small fixtures.  The GPU
box has no /root/reference; tests read only the fixtures.
albumentations / matplotlib / seaborn are stubbed at import time only (they are not on the
model / loss / metric arithmetic path).
"""
import json
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("B2SEG_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)


def _stub(name):
    m = types.ModuleType(name)

    class _Any:
        def __init__(self, *a, **k):
            self.args, self.kwargs = a, k

        def __call__(self, *a, **k):
            raise RuntimeError(f"{name} stub called")

    def _getattr(attr):
        if attr.startswith("__"):
            raise AttributeError(attr)
        return _Any

    m.__getattr__ = _getattr
    sys.modules[name] = m
    return m


for n in ("albumentations", "matplotlib", "matplotlib.pyplot", "seaborn"):
    _stub(n)
sys.path.insert(0, REF)
import advanced_models as ref_models  # noqa: E402
import losses as ref_losses  # noqa: E402
import utils as ref_utils  # noqa: E402

from oracle import seg_losses, unet_functional  # noqa: E402

torch.set_num_threads(8)
LOSS_NAMES = list(seg_losses.CLOSED_FORM) + ["exponential_logarithmic", "distance_map_bce", "boundary"]


def loss_fixtures():
    out = {"kat": {}, "small": {}}
    arrays = {}
    for tag, shape in (("kat", (4, 1, 256, 256)), ("small", (2, 1, 32, 48))):
        logits, target = seg_losses.kat_inputs(shape)
        out[tag]["logits_sum"] = float(logits.double().sum())
        out[tag]["target_sum"] = float(target.sum())
        for name in LOSS_NAMES:
            crit = ref_losses.get_loss_function({"loss_fn": name})
            x = logits.clone().requires_grad_(True)
            val = crit(x, target)
            val.backward()
            out[tag][name] = {"loss": float(val), "grad_l1": float(x.grad.abs().sum())}
            if tag == "small":
                arrays[f"grad_{name}"] = x.grad.numpy().copy()
        # non-default hyper-parameters exercise every config key of losses.py:345-396
        cfg = {"smooth": 0.5, "beta": 3.0, "focal_alpha": 0.4, "focal_gamma": 1.5, "loss_alpha": 0.3,
               "alpha_dice": 0.2, "alpha_bce": 0.5, "alpha_focal": 0.3, "tversky_alpha": 0.3,
               "tversky_beta": 0.7, "focal_tversky_gamma": 1.33, "sensitivity_weight": 0.7}
        out[tag]["custom_cfg"] = cfg
        for name in seg_losses.CLOSED_FORM:
            crit = ref_losses.get_loss_function(dict(cfg, loss_fn=name))
            x = logits.clone().requires_grad_(True)
            val = crit(x, target)
            val.backward()
            out[tag][name + "@custom"] = {"loss": float(val), "grad_l1": float(x.grad.abs().sum())}
        with torch.no_grad():
            out[tag]["metrics"] = ref_utils.calculate_metrics(torch.sigmoid(logits), target)
            out[tag]["iou_score"] = float(ref_utils.iou_score(logits, target))
            out[tag]["utils_dice_loss"] = float(ref_utils.dice_loss(logits, target))
            out[tag]["utils_combo_loss"] = float(ref_utils.combo_loss(logits, target))
    # unknown name falls back to ComboLoss (losses.py:401-403)
    crit = ref_losses.get_loss_function({"loss_fn": "no_such_loss"})
    out["fallback_class"] = type(crit).__name__
    with open(os.path.join(HERE, "loss_kat.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    np.savez_compressed(os.path.join(HERE, "loss_small_grads.npz"), **arrays)


def model_fixtures():
    meta = {}
    arrays = {}
    for backbone, attn, size, batch in (("resnet34", True, 64, 2), ("resnet34", False, 64, 2),
                                        ("resnet50", False, 64, 2), ("resnet34", True, 96, 1)):
        tag = f"{backbone}_{'attn' if attn else 'noattn'}_{size}"
        torch.manual_seed(7)
        model = ref_models.UNetWithBackbone(n_classes=1, backbone=backbone, pretrained=False,
                                            use_attention=attn)
        sd = unet_functional.synth_state(model.state_dict(), seed=3)
        model.load_state_dict(sd)
        g = torch.Generator().manual_seed(11)
        x = torch.rand(batch, 1, size, size, generator=g)
        t = (torch.rand(batch, 1, size, size, generator=g) > 0.8).float()
        crit = ref_losses.get_loss_function({"loss_fn": "combo"})
        model.train()
        logits = model(x)
        loss = crit(logits, t)
        loss.backward()
        gnorm = {k: float(p.grad.double().norm()) for k, p in model.named_parameters()}
        new_sd = model.state_dict()
        model.eval()
        with torch.no_grad():
            logits_eval = model(x)
        arrays[tag + "_logits_train"] = logits.detach().numpy()
        arrays[tag + "_logits_eval"] = logits_eval.numpy()
        arrays[tag + "_bn1_running_mean"] = new_sd["bn1.running_mean"].numpy()
        arrays[tag + "_bn1_running_var"] = new_sd["bn1.running_var"].numpy()
        arrays[tag + "_grad_input_conv"] = model.input_conv.weight.grad.numpy()
        arrays[tag + "_grad_conv_final"] = model.conv_final.weight.grad.numpy()
        meta[tag] = {"loss": float(loss), "grad_norms": gnorm, "n_state": len(new_sd),
                     "n_params": sum(p.numel() for p in model.parameters()),
                     "state_keys_sha": __import__("hashlib").sha1(
                         "\n".join(f"{k}:{tuple(v.shape)}" for k, v in new_sd.items()).encode()).hexdigest()}
        print(tag, meta[tag]["loss"], meta[tag]["n_params"], meta[tag]["n_state"])
    with open(os.path.join(HERE, "model_kat.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)
    np.savez_compressed(os.path.join(HERE, "model_kat.npz"), **arrays)


if __name__ == "__main__":
    loss_fixtures()
    model_fixtures()
    print("golden fixtures written to", HERE)
