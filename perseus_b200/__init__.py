"""perseus_b200 — B200-native (sm_100a) implementation of perseus's detector hot path.

Mirrors the reference package layout for the pieces on the path:
  perseus_b200.detector.models.KeypointCNN   <- perseus/detector/models.py
  perseus_b200.detector.decode               <- kornia (de)normalize_pixel_coordinates call sites
  perseus_b200.smoother.factors              <- perseus/smoother/factors.py
Everything computes through libperseus_b200.so (include/perseus_b200.h); there is no fallback.
"""
from pathlib import Path

ROOT = str(Path(__file__).parent.parent)  # same constant as perseus/__init__.py:4

__all__ = ["ROOT"]
