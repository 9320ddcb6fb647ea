#!/usr/bin/env python
"""Launches of the warp kernel at the hbm_bench shape (128 planes x 1024^2) for an ncu capture:
    ncu --set full --clock-control none --import-source on -k regex:warp_u8 -s 4 -c 3 -o gpurun_out/warp python tools/warp_profile.py
Order of launches after 4 warm-ups: affine bilinear / reflect, affine nearest / constant, remap bilinear / reflect."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "image-segmentation-project_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)
import torch  # noqa: E402
from b2seg import augment  # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator(device="cpu").manual_seed(1234)
n, size = 64, 1024
planes = torch.randint(0, 256, (2 * n, size, size), generator=g, dtype=torch.uint8).to(dev)
m = augment.affine_matrix(size, size, 1.03, (0.02, -0.03), 11.0, (3.0, -2.0))
yy, xx = torch.meshgrid(torch.arange(size, dtype=torch.float32), torch.arange(size, dtype=torch.float32), indexing="ij")
mx = (xx + 3.0 * torch.sin(yy / 17.0)).to(dev)
my = (yy + 3.0 * torch.cos(xx / 13.0)).to(dev)
for _ in range(4):
    augment.warp_affine(planes, m, augment.LINEAR, augment.REFLECT)
torch.cuda.synchronize()
augment.warp_affine(planes, m, augment.LINEAR, augment.REFLECT)
augment.warp_affine(planes, m, augment.NEAREST, augment.CONSTANT)
augment.remap(planes, mx, my)
torch.cuda.synchronize()
