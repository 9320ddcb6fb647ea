"""Runs the affine warp of tools/hbm_bench.py a few times (target of an ncu capture)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "image-segmentation-project_b200")); sys.path.insert(0, ROOT)
import torch
from b2seg import augment
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(1)
planes = torch.randint(0, 256, (128, 1024, 1024), generator=g, dtype=torch.uint8).to(dev)
m = augment.affine_matrix(1024, 1024, 1.03, (0.02, -0.03), 11.0, (3.0, -2.0))
for _ in range(3):
    augment.warp_affine(planes, m, augment.LINEAR, augment.REFLECT)
torch.cuda.synchronize()
print("ok")
