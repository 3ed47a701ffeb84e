"""Pixel-coordinate (de)normalisation used around the detector.

The reference calls kornia.geometry.denormalize_pixel_coordinates (validate.py:144-153,
validate_real.py:76-78, scripts/streaming.py:129-131) and normalize_pixel_coordinates
(augmentations.py:481).  kornia's definition: factor = 2 / (size - 1).clamp(eps);
denormalize = (x + 1) / factor; normalize = factor * x - 1; last axis is (x, y) vs (width, height).
"""
from __future__ import annotations

import torch

from .. import _lib


def denormalize_pixel_coordinates(pixel_coordinates: torch.Tensor, height: int, width: int) -> torch.Tensor:
    """[-1, 1] -> pixels, on the GPU through pb200_decode_pixels.  (..., 2) float cuda tensor."""
    if pixel_coordinates.shape[-1] != 2:
        raise ValueError(f"Input pixel_coordinates must be of shape (*, 2). Got {tuple(pixel_coordinates.shape)}")
    if not pixel_coordinates.is_cuda:
        raise RuntimeError("denormalize_pixel_coordinates: cuda tensor required (no CPU fallback)")
    x = pixel_coordinates.to(torch.float32).contiguous()
    out = torch.empty_like(x)
    with torch.cuda.device(x.device):
        _lib.check(
            _lib.lib().pb200_decode_pixels(x.data_ptr(), out.data_ptr(), x.numel() // 2, height, width,
                                          _lib.current_stream_ptr())
        )
    return out.to(pixel_coordinates.dtype)


def normalize_pixel_coordinates(pixel_coordinates: torch.Tensor, height: int, width: int, eps: float = 1e-8):
    """pixels -> [-1, 1] (label preparation, augmentations.py:481): plain tensor arithmetic, host-side
    bookkeeping on the label path, not on the detector hot path."""
    if pixel_coordinates.shape[-1] != 2:
        raise ValueError(f"Input pixel_coordinates must be of shape (*, 2). Got {tuple(pixel_coordinates.shape)}")
    hw = torch.tensor([width, height], device=pixel_coordinates.device, dtype=pixel_coordinates.dtype)
    factor = torch.tensor(2.0, device=pixel_coordinates.device, dtype=pixel_coordinates.dtype) / (hw - 1).clamp(eps)
    return factor * pixel_coordinates - 1
