"""U-Net with a ResNet encoder, attention gates and channel attention -- API mirror of the
reference `advanced_models.py` (UNetWithBackbone :64-357, AttentionGate :7-40,
ChannelAttention :43-61) running on the B200 kernels of libb2seg.

What stays identical to the reference: constructor signature and defaults, module / parameter
names (hence `state_dict()` keys, shapes and order), parameter initialisation (same torch RNG
consumption: torchvision builds the encoder containers, nn.Conv2d / nn.ConvTranspose2d /
nn.BatchNorm2d the rest), `pretrained=True` weight surgery for the grayscale stem, fp32
`nn.Parameter`s for `torch.optim.Adam`, `.train()/.eval()` BatchNorm semantics, logits
`[B,1,H,W]` fp32 that take part in autograd.

What differs: `forward` never calls those containers.  It replays a static plan of hand-written
sm_100a kernels (b2seg/unet_plan.py): tcgen05 implicit-GEMM convolutions in bf16 with fp32
accumulation, fused BatchNorm/ReLU/residual passes, skip-concat by construction.  There is no
CPU or eager-PyTorch fallback: on a CPU tensor / CPU parameters `forward` raises.
"""
import os

import torch
import torch.nn as nn
from torchvision import models

from b2seg import _lib as L
from b2seg.unet_plan import UNetPlan


class AttentionGate(nn.Module):
    """Parameter container of the additive attention gate (reference :7-40).  The arithmetic
    runs inside UNetWithBackbone's plan (1x1 tcgen05 convs + fused gate kernels)."""

    def __init__(self, F_g, F_l, F_int):
        super().__init__()
        self.W_g = nn.Sequential(nn.Conv2d(F_g, F_int, kernel_size=1, stride=1, padding=0, bias=True),
                                 nn.BatchNorm2d(F_int))
        self.W_x = nn.Sequential(nn.Conv2d(F_l, F_int, kernel_size=1, stride=1, padding=0, bias=True),
                                 nn.BatchNorm2d(F_int))
        self.psi = nn.Sequential(nn.Conv2d(F_int, 1, kernel_size=1, stride=1, padding=0, bias=True),
                                 nn.BatchNorm2d(1), nn.Sigmoid())
        self.relu = nn.ReLU(inplace=True)

    def forward(self, g, x):
        raise L.B2SegError("AttentionGate is fused into UNetWithBackbone's kernel plan; call the U-Net")


class ChannelAttention(nn.Module):
    """Parameter container of the avg+max squeeze-excite block (reference :43-61)."""

    def __init__(self, in_channels, reduction_ratio=16):
        super().__init__()
        self.avg_pool = nn.AdaptiveAvgPool2d(1)
        self.max_pool = nn.AdaptiveMaxPool2d(1)
        self.fc = nn.Sequential(nn.Conv2d(in_channels, in_channels // reduction_ratio, 1, bias=False),
                                nn.ReLU(inplace=True),
                                nn.Conv2d(in_channels // reduction_ratio, in_channels, 1, bias=False))
        self.sigmoid = nn.Sigmoid()

    def forward(self, x):
        raise L.B2SegError("ChannelAttention is fused into UNetWithBackbone's kernel plan; call the U-Net")


class _UNetStep(torch.autograd.Function):
    """logits = plan.forward(x); backward replays the planned backward sequence and hands the
    flat-buffer views to autograd as parameter gradients."""

    @staticmethod
    def forward(ctx, model, need_grad, x, *params):
        plan = model._plan_for(x, need_grad)
        logits = plan.run_forward(x)
        ctx.plan, ctx.generation, ctx.model = plan, plan.generation, model
        return logits.clone()

    @staticmethod
    def backward(ctx, dlogits):
        plan, model = ctx.plan, ctx.model
        if not plan.need_grad:
            raise L.B2SegError("backward through a forward that was planned without gradients")
        if plan.generation != ctx.generation:
            raise L.B2SegError("UNetWithBackbone.backward: the saved activations were overwritten by a "
                               "later forward of the same shape and mode; call backward before it")
        params = list(plan.params.values())
        # Gradients live in plan.gflat.  If a parameter's .grad still aliases it (zero_grad(
        # set_to_none=False) or gradient accumulation), keep the old values across the overwrite.
        lo, hi = plan.gflat.data_ptr(), plan.gflat.data_ptr() + plan.gflat.numel() * 4
        aliased = [p.grad is not None and lo <= p.grad.data_ptr() < hi for p in params]
        carry = plan.gflat.clone() if any(aliased) else None
        hook = model._grad_hook
        if hook is not None and getattr(hook, "bucketed", False) and len(plan.buckets) > 1:
            # data parallel: one graph per gradient bucket, its all-reduce right behind it
            plan.run_backward(dlogits.contiguous(), on_bucket=hook.bucket)
            hook.finish(plan.gflat)
            model._dp_buckets = len(plan.buckets)
        else:
            plan.run_backward(dlogits.contiguous())
            if hook is not None:
                hook(plan.gflat)
                model._dp_buckets = 1
        if carry is not None:
            plan.gflat.add_(carry)
        grads = []
        for (name, p), al in zip(plan.params.items(), aliased):
            if not p.requires_grad or al:
                grads.append(None)  # aliased: p.grad already shows old + new through the flat buffer
            elif carry is not None:
                grads.append(plan.gview(name) - carry[plan.goff[name]:plan.goff[name] + p.numel()].view(p.shape))
            else:
                grads.append(plan.gview(name))
        return (None, None, None, *grads)


class UNetWithBackbone(nn.Module):
    """UNetWithBackbone(n_classes=1, backbone='resnet34', pretrained=True, use_attention=True)."""

    def __init__(self, n_classes=1, backbone='resnet34', pretrained=True, use_attention=True):
        super().__init__()
        self.use_attention = use_attention
        self.backbone_name = backbone
        if backbone == 'densenet121':
            raise NotImplementedError(
                "backbone='densenet121' cannot run in the reference either (decoder4 expects 512 "
                "channels, the encoder yields 1024: advanced_models.py:132-155,284-293) and no "
                "experiment uses it; the B200 path implements resnet34 and resnet50")
        if backbone not in ('resnet34', 'resnet50'):
            raise ValueError(f"unknown backbone {backbone!r}")
        r50 = backbone == 'resnet50'
        net = getattr(models, backbone)(weights='IMAGENET1K_V1' if pretrained else None)
        self.input_conv = nn.Conv2d(1, 64, kernel_size=7, stride=2, padding=3, bias=False)
        if pretrained:  # grayscale stem = sum of the RGB filters (reference :78,108)
            self.input_conv.weight.data = torch.sum(net.conv1.weight.data, dim=1, keepdim=True)
        self.bn1, self.relu, self.maxpool = net.bn1, net.relu, net.maxpool
        self.enc1, self.enc2, self.enc3, self.enc4 = net.layer1, net.layer2, net.layer3, net.layer4
        w = (256, 512, 1024, 2048) if r50 else (64, 128, 256, 512)
        low = 64 if r50 else 32
        self.upconv4 = nn.ConvTranspose2d(w[3], w[2], kernel_size=2, stride=2)
        self.decoder4 = self._decoder_block(2 * w[2], w[2])
        self.upconv3 = nn.ConvTranspose2d(w[2], w[1], kernel_size=2, stride=2)
        self.decoder3 = self._decoder_block(2 * w[1], w[1])
        self.upconv2 = nn.ConvTranspose2d(w[1], w[0], kernel_size=2, stride=2)
        self.decoder2 = self._decoder_block(2 * w[0], w[0])
        self.upconv1 = nn.ConvTranspose2d(w[0], low, kernel_size=2, stride=2)
        self.decoder1 = self._decoder_block(64 + low, low)
        self.upconv0 = nn.ConvTranspose2d(low, 16, kernel_size=2, stride=2)
        self.conv_final = nn.Conv2d(16, n_classes, kernel_size=1)
        if use_attention:
            self.attention4 = AttentionGate(F_g=w[2], F_l=w[2], F_int=w[2] // 2)
            self.attention3 = AttentionGate(F_g=w[1], F_l=w[1], F_int=w[1] // 2)
            self.attention2 = AttentionGate(F_g=w[0], F_l=w[0], F_int=w[0] // 2)
            self.attention1 = AttentionGate(F_g=low, F_l=64, F_int=32)
            self.ch_attention4 = ChannelAttention(w[2])
            self.ch_attention3 = ChannelAttention(w[1])
            self.ch_attention2 = ChannelAttention(w[0])
            self.ch_attention1 = ChannelAttention(low)
        self._plans = {}
        self._grad_hook = None  # data-parallel runs install the flat-gradient all-reduce here

    @staticmethod
    def _decoder_block(in_channels, out_channels):
        return nn.Sequential(
            nn.Conv2d(in_channels, out_channels, kernel_size=3, padding=1), nn.BatchNorm2d(out_channels),
            nn.ReLU(inplace=True),
            nn.Conv2d(out_channels, out_channels, kernel_size=3, padding=1), nn.BatchNorm2d(out_channels),
            nn.ReLU(inplace=True))

    # plans are cached per (batch, H, W, training, needs-grad); rebuilt if parameters moved
    def _plan_for(self, x, need_grad):
        key = (x.shape[0], x.shape[2], x.shape[3], self.training, bool(need_grad))
        plan = self._plans.pop(key, None)
        if plan is not None and plan.params_moved():
            self._plans.clear()
            plan = None
        if plan is None:
            max_plans = int(os.environ.get("B2SEG_MAX_PLANS", "6"))
            while len(self._plans) >= max_plans:  # least recently used first (dict order = use order)
                self._plans.pop(next(iter(self._plans)))
            plan = UNetPlan(self, x.shape[0], x.shape[2], x.shape[3], self.training, need_grad)
        self._plans[key] = plan  # (re-)inserted last: most recently used
        return plan

    def __getstate__(self):  # copy.deepcopy / pickle: plans hold device handles, rebuild lazily
        d = self.__dict__.copy()
        d["_plans"], d["_grad_hook"] = {}, None
        return d

    def _apply(self, fn, *a, **k):  # .to()/.cuda()/.float(): parameter storage changes
        self._plans = {}
        return super()._apply(fn, *a, **k)

    def forward(self, x, return_features=False):
        if return_features:
            raise NotImplementedError("return_features=True (reference :352-356) is not used by any "
                                      "training or experiment script and is not on the B200 path")
        if x.dim() != 4 or x.shape[1] != 1:
            raise ValueError(f"expected a [B,1,H,W] grayscale batch, got {tuple(x.shape)}")
        p0 = self.input_conv.weight
        if not x.is_cuda or not p0.is_cuda:
            raise L.B2SegError("UNetWithBackbone runs on a Blackwell GPU only (no CPU fallback): move the "
                               "model and the batch to CUDA")
        if x.device != p0.device:
            raise L.B2SegError(f"input on {x.device}, parameters on {p0.device}")
        params = list(self.parameters())
        need_grad = torch.is_grad_enabled() and any(p.requires_grad for p in params)
        return _UNetStep.apply(self, need_grad, x, *params)
