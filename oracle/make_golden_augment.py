"""Generates tests/golden/augment_golden.npz from the REAL reference classes (perseus/detector/augmentations.py
imported from /root/reference with `kornia` stubbed — the custom classes only need torch).  Run in the build container:

    python oracle/make_golden_augment.py

The fixture holds the inputs, the torch draws the reference made (replayed under the same seed, in the reference's own
call order) and the reference outputs, so the oracle restatement can be pinned where /root/reference does not exist.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "tests", "golden", "augment_golden.npz")


def import_reference_augmentations():
    if "kornia" not in sys.modules:
        k = types.ModuleType("kornia")
        k.augmentation = types.ModuleType("kornia.augmentation")
        k.geometry = types.ModuleType("kornia.geometry")
        k.geometry.conversions = types.ModuleType("kornia.geometry.conversions")
        sys.modules.update({"kornia": k, "kornia.augmentation": k.augmentation, "kornia.geometry": k.geometry,
                            "kornia.geometry.conversions": k.geometry.conversions})
    if "/root/reference" not in sys.path:
        sys.path.insert(0, "/root/reference")
    import perseus.detector.augmentations as ra

    return ra


def synthetic_batch(B: int, H: int, W: int, seed: int) -> torch.Tensor:
    """(B, 5, H, W): RGB in [0,1], depth in data units (metres / 0.035), a rectangular 'cube' segmentation."""
    g = torch.Generator().manual_seed(seed)
    rgb = torch.rand(B, 3, H, W, generator=g)
    depth = (0.15 + 0.4 * torch.rand(B, H, W, generator=g)) / 0.035
    seg = torch.zeros(B, H, W)
    for b in range(B):
        y0, x0 = int(torch.randint(0, H // 2, (1,), generator=g)), int(torch.randint(0, W // 2, (1,), generator=g))
        h, w = int(torch.randint(H // 8, H // 2, (1,), generator=g)), int(torch.randint(W // 8, W // 2, (1,), generator=g))
        seg[b, y0:y0 + h, x0:x0 + w] = 1.0
        depth[b, y0:y0 + h, x0:x0 + w] *= 0.6  # the cube is closer than its background
    return torch.cat([rgb, depth[:, None], seg[:, None]], dim=1)


def main():
    ra = import_reference_augmentations()
    out = {}
    B, H, W = 6, 24, 20
    images = synthetic_batch(B, H, W, seed=3)
    images[5, 4] = 0.0  # an image without a cube: its transplant is rejected by the ratio bounds or accepted as donor-only
    # --- RandomTransplantationWithDepth: the one draw is torch.randint(1, B, (B,)) at :207
    torch.manual_seed(11)
    donors = (torch.arange(B) + torch.randint(1, B, (B,))) % B
    torch.manual_seed(11)
    ref = ra.RandomTransplantationWithDepth()(images.clone())
    out.update(tp_images=images.numpy(), tp_donors=donors.numpy(), tp_out=ref.numpy())
    # --- depth classes: replay dropout + rand (+ randn) in the reference's order
    d = images[:, 3].clone()
    torch.manual_seed(12)
    m = F.dropout(torch.ones_like(d), p=0.5, training=True)
    u = torch.rand_like(d)
    torch.manual_seed(12)
    ref = ra.DepthBiasAugmentation(dev=0.02, p_bias=0.5, cube_scale=0.035)(d.clone())
    out.update(bias_in=d.numpy(), bias_mask=m.numpy(), bias_u=u.numpy(), bias_out=ref.numpy())
    torch.manual_seed(13)
    z = torch.randn_like(d)
    torch.manual_seed(13)
    ref = ra.DepthGaussianNoiseAugmentation(std=0.005)(d.clone())
    out.update(noise_z=z.numpy(), noise_out=ref.numpy())
    torch.manual_seed(14)
    mn = F.dropout(torch.ones_like(d), p=0.5, training=True)
    un = torch.rand_like(d)
    mf = F.dropout(torch.ones_like(d), p=0.5, training=True)
    uf = torch.rand_like(d)
    torch.manual_seed(14)
    ref = ra.DepthPlaneAugmentation()(d.clone())
    out.update(pl_mn=mn.numpy(), pl_un=un.numpy(), pl_mf=mf.numpy(), pl_uf=uf.numpy(), pl_out=ref.numpy())
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
