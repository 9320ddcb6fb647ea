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


def bf16_round(t):
    """Round to the nearest bf16 value, keep fp32 storage (straight-through for autograd)."""
    return t + (t.detach().to(torch.bfloat16).float() - t.detach())


def teacher_store(values, computed=None, grad_round=None, grads=None):
    """`store` hook that replaces every named storage point by the tensor in `values` (NCHW fp32,
    e.g. what the B200 plan holds in HBM) with a straight-through gradient, and records what the
    oracle itself computed from the substituted upstream values in `computed`.  Downstream layers
    and the whole backward pass then see exactly the product's forward activations, so layer-level
    parity is free of the chaotic bf16 drift a free-running comparison accumulates.
    `grad_round` (e.g. bf16_round) is applied to the gradient arriving at every storage point:
    the statistical twin of the product's bf16 gradient buffers.  `grads` (dict) receives the
    gradient w.r.t. every storage point."""
    def store(name, t, pre=None):
        if computed is not None:
            computed[name] = t.detach()
        v = values.get(name)
        if v is not None:
            v = v.to(t.dtype).reshape(t.shape)
            if pre is not None:
                # ReLU output: gradient flows where the SUBSTITUTED activation is positive, so a
                # pre-activation within rounding distance of 0 cannot flip the mask between the
                # two implementations (a handful of such flips dominate heavy-tailed gradients)
                lin = pre * (v > 0).to(pre.dtype)
                t = v + (lin - lin.detach())
            else:
                t = t + (v - t.detach())
        if grads is not None and t.requires_grad:  # gradient w.r.t. this storage point
            t.register_hook(lambda g, name=name: grads.__setitem__(name, g.detach().clone()))
        if grad_round is not None and t.requires_grad:
            # the product keeps activation gradients in bf16 at these same points
            t.register_hook(lambda g: grad_round(g.detach()))
        return t
    return store


def unet_forward(sd, x, backbone="resnet34", use_attention=True, training=True, new_stats=None,
                 taps=None, storage_round=None, store=None):
    """logits = UNetWithBackbone.forward(x) (advanced_models.py:264-357) from state dict `sd`.

    `taps` (optional dict) receives named intermediate activations for layer-level parity.
    `storage_round` (e.g. bf16_round) is applied to every tensor-core conv weight and to every
    activation the B200 path stores in HBM between kernels: the "same arithmetic, bf16 storage"
    comparator (all sums still accumulate in fp32).  `store(name, tensor) -> tensor` is called at
    each of those storage points after rounding (see teacher_store)."""
    kind, depths = RESNET_LAYERS[backbone]
    q = storage_round if storage_round is not None else (lambda t: t)

    def st(name, t):
        t = q(t)
        if store is not None:
            t = store(name, t)
        if taps is not None:
            taps[name] = t
        return t

    def relu_st(name, pre):  # ReLU whose output is a storage point
        t = q(F.relu(pre))
        if store is not None:
            t = store(name, t, pre)
        if taps is not None:
            taps[name] = t
        return t

    def bn(prefix):
        return _BN(sd, prefix, training, new_stats)

    def conv(name, inp, wkey, bkey=None, stride=1, pad=0):  # tensor-core conv: bf16 weights + output
        return st(name, F.conv2d(inp, q(sd[wkey]), sd[bkey] if bkey else None, stride, pad))

    # stem: 7x7 s2 conv (1 input channel, fp32 weights), BN, ReLU (:268-270); 3x3 s2 max-pool (:272)
    y0 = st("stem.y", F.conv2d(x, sd["input_conv.weight"], None, 2, 3))
    x1 = relu_st("x1", (bn("bn1")(y0)))
    h = F.max_pool2d(x1, 3, 2, 1)
    skips = [x1]
    for li, nblocks in enumerate(depths, start=1):  # enc1..enc4 (:273-276), torchvision blocks
        for bi in range(nblocks):
            pre = f"enc{li}.{bi}"
            stride = 2 if (li > 1 and bi == 0) else 1
            idt = h
            if kind == "basic":
                o = relu_st(pre + ".a1", (bn(pre + ".bn1")(conv(pre + ".y1", h, pre + ".conv1.weight", None, stride, 1))))
                o = bn(pre + ".bn2")(conv(pre + ".y2", o, pre + ".conv2.weight", None, 1, 1))
            else:  # Bottleneck v1.5: stride on the 3x3
                o = relu_st(pre + ".a1", (bn(pre + ".bn1")(conv(pre + ".y1", h, pre + ".conv1.weight"))))
                o = relu_st(pre + ".a2", (bn(pre + ".bn2")(conv(pre + ".y2", o, pre + ".conv2.weight", None, stride, 1))))
                o = bn(pre + ".bn3")(conv(pre + ".y3", o, pre + ".conv3.weight"))
            if pre + ".downsample.0.weight" in sd:
                idt = bn(pre + ".downsample.1")(conv(pre + ".yd", h, pre + ".downsample.0.weight", None, stride))
            h = relu_st(pre + ".out", (o + idt))
        if taps is not None:
            taps[f"x{li + 1}"] = h
        skips.append(h)
    x1, x2, x3, x4, x5 = skips

    def gate(k, g, s):  # AttentionGate.forward (:28-40)
        a = f"attention{k}"
        g1 = bn(a + ".W_g.1")(conv(f"att{k}.yg", g, a + ".W_g.0.weight", a + ".W_g.0.bias"))
        s1 = bn(a + ".W_x.1")(conv(f"att{k}.ys", s, a + ".W_x.0.weight", a + ".W_x.0.bias"))
        psi = relu_st(f"att{k}.pin", (g1 + s1))
        pre = F.conv2d(psi, sd[a + ".psi.0.weight"], sd[a + ".psi.0.bias"])  # 1 channel: fp32 path
        if store is not None:
            pre = store(f"att{k}.pre", pre)
        psi = torch.sigmoid(bn(a + ".psi.1")(pre))
        return st(f"att{k}.out", s * psi)

    def squeeze_excite(k, d):  # ChannelAttention.forward (:57-61); fc shared by both pools (fp32)
        c = f"ch_attention{k}"

        def fc(v):
            return F.conv2d(F.relu(F.conv2d(v, sd[c + ".fc.0.weight"])), sd[c + ".fc.2.weight"])

        w = fc(F.adaptive_avg_pool2d(d, 1)) + fc(F.adaptive_max_pool2d(d, 1))
        return st(f"se{k}.out", d * torch.sigmoid(w))

    def dec_block(k, d):  # _decoder_block (:197-205)
        p = f"decoder{k}"
        d = relu_st(f"dec{k}.a", (bn(p + ".1")(conv(f"dec{k}.ya", d, p + ".0.weight", p + ".0.bias", 1, 1))))
        return relu_st(f"dec{k}.b", (bn(p + ".4")(conv(f"dec{k}.yb", d, p + ".3.weight", p + ".3.bias", 1, 1))))

    d = x5
    for k, skip in ((4, x4), (3, x3), (2, x2), (1, x1)):  # (:284-334)
        d = st(f"up{k}", F.conv_transpose2d(d, q(sd[f"upconv{k}.weight"]), sd[f"upconv{k}.bias"], 2))
        if k == 1 and d.shape != skip.shape:  # centre crop only when both diffs > 0 (:318-325)
            dh, dw = skip.shape[2] - d.shape[2], skip.shape[3] - d.shape[3]
            if dh > 0 and dw > 0:
                skip = skip[:, :, dh // 2:dh // 2 + d.shape[2], dw // 2:dw // 2 + d.shape[3]]
        s = gate(k, d, skip) if use_attention else skip
        d = dec_block(k, torch.cat((s, d), dim=1))  # skip FIRST (:288,292,...)
        if use_attention:
            d = squeeze_excite(k, d)
        if taps is not None:
            taps[f"d{k}"] = d
    d0 = st("d0", F.conv_transpose2d(d, q(sd["upconv0.weight"]), sd["upconv0.bias"], 2))  # (:337)
    if d0.shape[2] != x.shape[2] or d0.shape[3] != x.shape[3]:  # (:340-347)
        dh, dw = d0.shape[2] - x.shape[2], d0.shape[3] - x.shape[3]
        if dh > 0 or dw > 0:
            d0 = d0[:, :, dh // 2:dh // 2 + x.shape[2], dw // 2:dw // 2 + x.shape[3]]
    logits = F.conv2d(d0, sd["conv_final.weight"], sd["conv_final.bias"])  # (:350), fp32 path
    if store is not None:
        logits = store("logits", logits)
    return logits


def forward_backward(sd, x, target, loss_fn, **kw):
    """One oracle training step: returns (logits, loss, {param_name: grad})."""
    names = [k for k, v in sd.items() if v.is_floating_point() and "running_" not in k]
    leaf = {k: (sd[k].detach().clone().requires_grad_(True) if k in names else sd[k]) for k in sd}
    logits = unet_forward(leaf, x, **kw)
    loss = loss_fn(logits, target)
    grads = torch.autograd.grad(loss, [leaf[k] for k in names], allow_unused=True)
    return logits.detach(), float(loss), {k: g for k, g in zip(names, grads)}
