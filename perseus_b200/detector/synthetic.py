"""Seeded random-init weights for benchmarks and smoke runs (there is no network for checkpoints).

Same recipe the parity tests use: the module tree built under torch.manual_seed(seed) (torchvision init:
kaiming-normal convs, BN gamma = 1 / beta = 0, default Linear init; reference models.py:20-32 minus the
ImageNet download), then BatchNorm affine parameters and running statistics re-drawn so that BN folding
is exercised (gamma ~ U(0.5, 1.5), beta ~ N(0, 0.1), mean ~ N(0, 0.1), var ~ U(0.5, 2)).
"""
from __future__ import annotations

import torch


def random_init_state_dict(num_channels: int = 3, n_keypoints: int = 8, seed: int = 0, randomize_bn: bool = True) -> dict:
    from .models import KeypointCNN

    torch.manual_seed(seed)
    sd = KeypointCNN(n_keypoints=n_keypoints, num_channels=num_channels).state_dict()
    if randomize_bn:
        g = torch.Generator().manual_seed(seed + 1)
        for k in sorted(sd.keys()):
            v = sd[k]
            is_bn = "bn" in k or "downsample.1" in k
            if k.endswith("running_var"):
                v.copy_(torch.rand(v.shape, generator=g) * 1.5 + 0.5)
            elif k.endswith("running_mean"):
                v.copy_(torch.randn(v.shape, generator=g) * 0.1)
            elif is_bn and k.endswith(".weight"):
                v.copy_(torch.rand(v.shape, generator=g) + 0.5)
            elif is_bn and k.endswith(".bias"):
                v.copy_(torch.randn(v.shape, generator=g) * 0.1)
    return sd
