"""Functional fp32 CPU restatement of UNetWithBackbone (oracle; tests only).

The network is evaluated straight from a state_dict with torch.nn.functional
calls, so it shares no module code with either the reference or the product.
State-dict keys are the reference's (advanced_models.py:64-195 + torchvision
resnet naming).  Citations give the reference lines each stage restates.
"""
import torch
import torch.nn.functional as F

RESNET_LAYERS = {"resnet34": ("basic", (3, 4, 6, 3)), "resnet50": ("bottleneck", (3, 4, 6, 3))}


def synth_state(reference_state, seed=0):
    """Deterministic, name-keyed synthetic weights with the shapes of `reference_state`.

    Used both by tests/golden/make_golden.py (loaded into the *reference* module) and by the
    tests (loaded into oracle and product), so all three see identical parameters without
    shipping a 100 MB checkpoint."""
    import zlib
    out = {}
    for k, v in reference_state.items():
        g = torch.Generator().manual_seed(seed * 100003 + (zlib.crc32(k.encode()) & 0xFFFFFF))
        if k.endswith("num_batches_tracked"):
            out[k] = torch.zeros_like(v)
        elif k.endswith("running_var"):
            out[k] = 0.5 + torch.rand(v.shape, generator=g)
        elif k.endswith("running_mean"):
            out[k] = 0.1 * torch.randn(v.shape, generator=g)
        elif v.dim() == 1 and ("bn" in k or ".1.weight" in k or ".4.weight" in k or "downsample.1.weight" in k) \
                and k.endswith("weight"):
            out[k] = 0.75 + 0.5 * torch.rand(v.shape, generator=g)  # BN gamma
        elif v.dim() == 1:
            out[k] = 0.1 * torch.randn(v.shape, generator=g)  # biases / BN beta
        else:
            fan_in = v[0].numel()
            out[k] = torch.randn(v.shape, generator=g) * (2.0 / fan_in) ** 0.5
    return out


class _BN:
    """BatchNorm2d semantics: batch statistics (biased var) when training, running otherwise;
    running stats updated with momentum 0.1 and the unbiased variance."""

    def __init__(self, sd, prefix, training, new_stats):
        self.sd, self.p, self.training, self.new_stats = sd, prefix, training, new_stats

    def __call__(self, x):
        sd, p = self.sd, self.p
        if self.training and self.new_stats is not None:
            with torch.no_grad():
                n = x.numel() // x.shape[1]
                mean = x.mean(dim=(0, 2, 3))
                var_u = x.var(dim=(0, 2, 3), unbiased=True) if n > 1 else torch.zeros_like(mean)
                self.new_stats[p + ".running_mean"] = 0.9 * sd[p + ".running_mean"] + 0.1 * mean
                self.new_stats[p + ".running_var"] = 0.9 * sd[p + ".running_var"] + 0.1 * var_u
                self.new_stats[p + ".num_batches_tracked"] = sd[p + ".num_batches_tracked"] + 1
        return F.batch_norm(x, None if self.training else sd[p + ".running_mean"],
                            None if self.training else sd[p + ".running_var"],
                            sd[p + ".weight"], sd[p + ".bias"], self.training, 0.1, 1e-5)


_STORE = [None]


def _store(t):
    return _STORE[0](t) if _STORE[0] is not None else t


def bf16_round(t):
    """Round to the nearest bf16 value, keep fp32 storage (straight-through for autograd)."""
    return t + (t.detach().to(torch.bfloat16).float() - t.detach())


def unet_forward(sd, x, backbone="resnet34", use_attention=True, training=True, new_stats=None,
                 taps=None, storage_round=None):
    """logits = UNetWithBackbone.forward(x) (advanced_models.py:264-357) from state dict `sd`.

    `taps` (optional dict) receives named intermediate activations for layer-level parity.
    `storage_round` (e.g. bf16_round) is applied to every conv weight, every conv output and every
    stored activation: the "same arithmetic, bf16 storage" comparator for the B200 path, which
    keeps activations in bf16 between kernels (all sums still accumulate in fp32)."""
    kind, depths = RESNET_LAYERS[backbone]
    if storage_round is not None:
        return _unet_forward_rounded(sd, x, backbone, use_attention, training, new_stats, taps, storage_round)

    def bn(prefix):
        return _BN(sd, prefix, training, new_stats)

    def tap(name, v):
        if taps is not None:
            taps[name] = v
        return v

    # stem: 7x7 s2 conv (1 input channel), BN, ReLU (:268-270); 3x3 s2 max-pool (:272)
    x1 = F.relu(bn("bn1")(F.conv2d(x, sd["input_conv.weight"], None, 2, 3)))
    tap("x1", x1)
    h = F.max_pool2d(x1, 3, 2, 1)
    skips = [x1]
    for li, nblocks in enumerate(depths, start=1):  # enc1..enc4 (:273-276), torchvision blocks
        for bi in range(nblocks):
            pre = f"enc{li}.{bi}"
            stride = 2 if (li > 1 and bi == 0) else 1
            idt = h
            if kind == "basic":
                o = F.relu(bn(pre + ".bn1")(F.conv2d(h, sd[pre + ".conv1.weight"], None, stride, 1)))
                o = bn(pre + ".bn2")(F.conv2d(o, sd[pre + ".conv2.weight"], None, 1, 1))
            else:  # Bottleneck v1.5: stride on the 3x3
                o = F.relu(bn(pre + ".bn1")(F.conv2d(h, sd[pre + ".conv1.weight"])))
                o = F.relu(bn(pre + ".bn2")(F.conv2d(o, sd[pre + ".conv2.weight"], None, stride, 1)))
                o = bn(pre + ".bn3")(F.conv2d(o, sd[pre + ".conv3.weight"]))
            if pre + ".downsample.0.weight" in sd:
                idt = bn(pre + ".downsample.1")(F.conv2d(h, sd[pre + ".downsample.0.weight"], None, stride))
            h = F.relu(o + idt)
        tap(f"x{li + 1}", h)
        skips.append(h)
    x1, x2, x3, x4, x5 = skips

    def gate(k, g, s):  # AttentionGate.forward (:28-40)
        a = f"attention{k}"
        g1 = bn(a + ".W_g.1")(F.conv2d(g, sd[a + ".W_g.0.weight"], sd[a + ".W_g.0.bias"]))
        s1 = bn(a + ".W_x.1")(F.conv2d(s, sd[a + ".W_x.0.weight"], sd[a + ".W_x.0.bias"]))
        psi = F.relu(g1 + s1)
        psi = torch.sigmoid(bn(a + ".psi.1")(F.conv2d(psi, sd[a + ".psi.0.weight"], sd[a + ".psi.0.bias"])))
        return _store(s * psi)

    def squeeze_excite(k, d):  # ChannelAttention.forward (:57-61); fc shared by both pools
        c = f"ch_attention{k}"

        def fc(v):
            return F.conv2d(F.relu(F.conv2d(v, sd[c + ".fc.0.weight"])), sd[c + ".fc.2.weight"])

        w = fc(F.adaptive_avg_pool2d(d, 1)) + fc(F.adaptive_max_pool2d(d, 1))
        return _store(d * torch.sigmoid(w))

    def dec_block(k, d):  # _decoder_block (:197-205)
        p = f"decoder{k}"
        d = F.relu(bn(p + ".1")(F.conv2d(d, sd[p + ".0.weight"], sd[p + ".0.bias"], 1, 1)))
        return F.relu(bn(p + ".4")(F.conv2d(d, sd[p + ".3.weight"], sd[p + ".3.bias"], 1, 1)))

    d = x5
    for k, skip in ((4, x4), (3, x3), (2, x2), (1, x1)):  # (:284-334)
        d = F.conv_transpose2d(d, sd[f"upconv{k}.weight"], sd[f"upconv{k}.bias"], 2)
        if k == 1 and d.shape != skip.shape:  # centre crop only when both diffs > 0 (:318-325)
            dh, dw = skip.shape[2] - d.shape[2], skip.shape[3] - d.shape[3]
            if dh > 0 and dw > 0:
                skip = skip[:, :, dh // 2:dh // 2 + d.shape[2], dw // 2:dw // 2 + d.shape[3]]
        s = gate(k, d, skip) if use_attention else skip
        d = dec_block(k, torch.cat((s, d), dim=1))  # skip FIRST (:288,292,...)
        if use_attention:
            d = squeeze_excite(k, d)
        tap(f"d{k}", d)
    d0 = F.conv_transpose2d(d, sd["upconv0.weight"], sd["upconv0.bias"], 2)  # (:337)
    if d0.shape[2] != x.shape[2] or d0.shape[3] != x.shape[3]:  # (:340-347)
        dh, dw = d0.shape[2] - x.shape[2], d0.shape[3] - x.shape[3]
        if dh > 0 or dw > 0:
            d0 = d0[:, :, dh // 2:dh // 2 + x.shape[2], dw // 2:dw // 2 + x.shape[3]]
    return F.conv2d(d0, sd["conv_final.weight"], sd["conv_final.bias"])  # (:350)


def _unet_forward_rounded(sd, x, backbone, use_attention, training, new_stats, taps, q):
    """Same network with `q` applied at every point where the B200 path stores bf16."""
    import contextlib
    orig = dict(conv2d=F.conv2d, conv_transpose2d=F.conv_transpose2d, relu=F.relu, batch_norm=F.batch_norm)

    def conv2d(inp, w, b=None, *a, **k):
        if w.shape[1] == 1 and w.shape[2] == 7:      # stem: fp32 weights and input, bf16 output
            return q(orig["conv2d"](inp, w, b, *a, **k))
        if w.shape[0] == 1 or inp.shape[2] * inp.shape[3] == 1:   # 1-channel convs / SE MLP: fp32
            return orig["conv2d"](inp, w, b, *a, **k)
        return q(orig["conv2d"](inp, q(w), b, *a, **k))

    def conv_transpose2d(inp, w, b=None, *a, **k):
        return q(orig["conv_transpose2d"](inp, q(w), b, *a, **k))

    def relu(t, *a, **k):
        return q(orig["relu"](t, *a, **k))

    F.conv2d, F.conv_transpose2d, F.relu = conv2d, conv_transpose2d, relu
    _STORE[0] = q
    try:
        return unet_forward(sd, x, backbone, use_attention, training, new_stats, taps, None)
    finally:
        _STORE[0] = None
        F.conv2d, F.conv_transpose2d, F.relu = orig["conv2d"], orig["conv_transpose2d"], orig["relu"]


def forward_backward(sd, x, target, loss_fn, **kw):
    """One oracle training step: returns (logits, loss, {param_name: grad})."""
    names = [k for k, v in sd.items() if v.is_floating_point() and "running_" not in k]
    leaf = {k: (sd[k].detach().clone().requires_grad_(True) if k in names else sd[k]) for k in sd}
    logits = unet_forward(leaf, x, **kw)
    loss = loss_fn(logits, target)
    grads = torch.autograd.grad(loss, [leaf[k] for k in names], allow_unused=True)
    return logits.detach(), float(loss), {k: g for k, g in zip(names, grads)}
