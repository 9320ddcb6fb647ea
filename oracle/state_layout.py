"""Names and shapes of the reference model's state_dict (oracle; tests only), rebuilt from
torchvision's resnet plus the decoder spec of advanced_models.py:89-100,119-130,158-195."""
from collections import OrderedDict

import torch

WIDTHS = {"resnet34": (64, 128, 256, 512), "resnet50": (256, 512, 1024, 2048)}


def _bn(out, prefix, c):
    out[prefix + ".weight"] = torch.empty(c)
    out[prefix + ".bias"] = torch.empty(c)
    out[prefix + ".running_mean"] = torch.empty(c)
    out[prefix + ".running_var"] = torch.empty(c)
    out[prefix + ".num_batches_tracked"] = torch.empty((), dtype=torch.long)


def reference_state_shapes(backbone="resnet34", use_attention=True):
    from torchvision import models
    net = getattr(models, backbone)(weights=None)
    out = OrderedDict()
    out["input_conv.weight"] = torch.empty(64, 1, 7, 7)
    _bn(out, "bn1", 64)
    for li in range(1, 5):
        for k, v in getattr(net, f"layer{li}").state_dict().items():
            out[f"enc{li}.{k}"] = torch.empty(v.shape, dtype=v.dtype)
    w = WIDTHS[backbone]
    c1_skip = 64
    up_out = {4: w[2], 3: w[1], 2: w[0], 1: 32 if backbone == "resnet34" else 64}
    up_in = {4: w[3], 3: w[2], 2: w[1], 1: w[0]}
    skip = {4: w[2], 3: w[1], 2: w[0], 1: c1_skip}
    for k in (4, 3, 2, 1):
        out[f"upconv{k}.weight"] = torch.empty(up_in[k], up_out[k], 2, 2)
        out[f"upconv{k}.bias"] = torch.empty(up_out[k])
        cin, cout = skip[k] + up_out[k], up_out[k]
        out[f"decoder{k}.0.weight"] = torch.empty(cout, cin, 3, 3)
        out[f"decoder{k}.0.bias"] = torch.empty(cout)
        _bn(out, f"decoder{k}.1", cout)
        out[f"decoder{k}.3.weight"] = torch.empty(cout, cout, 3, 3)
        out[f"decoder{k}.3.bias"] = torch.empty(cout)
        _bn(out, f"decoder{k}.4", cout)
    out["upconv0.weight"] = torch.empty(up_out[1], 16, 2, 2)
    out["upconv0.bias"] = torch.empty(16)
    out["conv_final.weight"] = torch.empty(1, 16, 1, 1)
    out["conv_final.bias"] = torch.empty(1)
    if use_attention:
        fint = {4: w[2] // 2, 3: w[1] // 2, 2: w[0] // 2, 1: 32}
        for k in (4, 3, 2, 1):
            a = f"attention{k}"
            for br, cin in (("W_g", up_out[k]), ("W_x", skip[k])):
                out[f"{a}.{br}.0.weight"] = torch.empty(fint[k], cin, 1, 1)
                out[f"{a}.{br}.0.bias"] = torch.empty(fint[k])
                _bn(out, f"{a}.{br}.1", fint[k])
            out[f"{a}.psi.0.weight"] = torch.empty(1, fint[k], 1, 1)
            out[f"{a}.psi.0.bias"] = torch.empty(1)
            _bn(out, f"{a}.psi.1", 1)
        for k in (4, 3, 2, 1):
            c = up_out[k]
            out[f"ch_attention{k}.fc.0.weight"] = torch.empty(c // 16, c, 1, 1)
            out[f"ch_attention{k}.fc.2.weight"] = torch.empty(c, c // 16, 1, 1)
    return out
