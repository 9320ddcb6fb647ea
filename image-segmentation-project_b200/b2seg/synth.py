"""Synthetic microscopy tiles (SURVEY section 8d recipe): a noisy background with filled, blurred
ellipses ("cells") and the ellipse union as the binary mask.  Pure torch-CPU with an explicit
generator, so the oracle, the reference arm and the B200 path all see identical bytes."""
import math

import torch
import torch.nn.functional as F


def cell_tiles(batch, size, seed=1234, cells_per_256=12):
    """-> (images [B,1,S,S] fp32 in [0,1], masks [B,1,S,S] fp32 {0,1})."""
    g = torch.Generator().manual_seed(seed)
    ncell = max(1, int(round(cells_per_256 * (size / 256.0) ** 2)))
    yy, xx = torch.meshgrid(torch.arange(size, dtype=torch.float32), torch.arange(size, dtype=torch.float32),
                            indexing="ij")
    masks = torch.zeros(batch, 1, size, size)
    for b in range(batch):
        cx = torch.rand(ncell, generator=g) * size
        cy = torch.rand(ncell, generator=g) * size
        ax = 6 + 12 * torch.rand(ncell, generator=g)
        ay = 6 + 12 * torch.rand(ncell, generator=g)
        th = math.pi * torch.rand(ncell, generator=g)
        m = torch.zeros(size, size, dtype=torch.bool)
        for i in range(ncell):
            dx, dy = xx - cx[i], yy - cy[i]
            u = dx * torch.cos(th[i]) + dy * torch.sin(th[i])
            v = -dx * torch.sin(th[i]) + dy * torch.cos(th[i])
            m |= (u / ax[i]) ** 2 + (v / ay[i]) ** 2 <= 1.0
        masks[b, 0] = m.float()
    img = 0.35 + 0.3 * masks
    k = torch.arange(-4, 5, dtype=torch.float32)
    k = torch.exp(-k ** 2 / (2 * 1.5 ** 2))
    k = (k / k.sum())
    img = F.conv2d(F.pad(img, (4, 4, 0, 0), mode="replicate"), k.view(1, 1, 1, 9))
    img = F.conv2d(F.pad(img, (0, 0, 4, 4), mode="replicate"), k.view(1, 1, 9, 1))
    img = (img + 0.05 * torch.randn(img.shape, generator=g)).clamp_(0.0, 1.0)
    return img.contiguous(), masks.contiguous()
