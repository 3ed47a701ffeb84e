"""Shared helpers for the -m gpu parity tests (all compute goes through the C ABI)."""
import os

import numpy as np
import torch

from oracle import detector_oracle as do


def dev():
    return torch.device("cuda", 0)


_MODELS = {}


def native_model(C=3, precision="bf16", seed=0, randomize_bn=True, max_batch=8):
    """perseus_b200 KeypointCNN on cuda:0 loaded with the oracle's seeded weights."""
    from perseus_b200.detector import KeypointCNN

    key = (C, precision, seed, randomize_bn)
    if key not in _MODELS:
        sd = do.seeded_state_dict(C, 8, seed=seed, randomize_bn=randomize_bn)
        m = KeypointCNN(num_channels=C, precision=precision, max_batch=max_batch)
        m.load_state_dict(sd)
        m.eval().to(dev())
        _MODELS[key] = (m, sd)
    return _MODELS[key]


def px_err(a: torch.Tensor, b: torch.Tensor, H=256, W=256):
    """max / mean pixel distance per coordinate between two normalised outputs (B, 16)."""
    d = (a.double().cpu() - b.double().cpu()).abs() * (W - 1) / 2.0
    return float(d.max()), float(d.mean())


def synth_u8_frames(B, h=256, w=256, ch=3, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, 256, (B, h, w, ch), dtype=torch.uint8, generator=g)


def synth_depth_m(B, h=256, w=256, seed=0, bad=True):
    """metres in [0.1, 0.5] with 1% NaN and 0.5% +-Inf injected."""
    g = torch.Generator().manual_seed(seed + 100)
    d = torch.rand(B, h, w, generator=g) * 0.4 + 0.1
    if bad:
        u = torch.rand(B, h, w, generator=g)
        d[u < 0.01] = float("nan")
        d[(u >= 0.01) & (u < 0.0125)] = float("inf")
        d[(u >= 0.0125) & (u < 0.015)] = float("-inf")
    return d


def frames_view(model, buf: torch.Tensor, B: int):
    """The ingested frames buffer as a (B, H+6, W+8, 4) tensor of the plan's element type."""
    dt = {"bf16": torch.bfloat16, "fp16x2": torch.float16}.get(model.precision, torch.float32)
    n = B * (model.H + 6) * (model.W + 8) * 4
    return buf.view(dt)[:n].view(B, model.H + 6, model.W + 8, 4)


def nhwc_to_p8(x: torch.Tensor) -> torch.Tensor:
    """(B, H, W, C) bf16 -> P8 (B, C/8, H+2, W+2, 8) bf16 with a zero border (csrc/layout.h)."""
    B, H, W, C = x.shape
    out = torch.zeros(B, C // 8, H + 2, W + 2, 8, dtype=x.dtype, device=x.device)
    out[:, :, 1:-1, 1:-1] = x.view(B, H, W, C // 8, 8).permute(0, 3, 1, 2, 4)
    return out


def p8_to_nhwc(p: torch.Tensor):
    """P8 -> ((B, H, W, C) interior, max |border|)."""
    B, C8, Hp, Wp, _ = p.shape
    border = p.clone().float()
    border[:, :, 1:-1, 1:-1] = 0
    x = p[:, :, 1:-1, 1:-1].permute(0, 2, 3, 1, 4).reshape(B, Hp - 2, Wp - 2, C8 * 8)
    return x, float(border.abs().max())
