"""Drop-in for perseus/detector/augmentations.py (AugmentationConfig :259-324, KeypointAugmentation :327-488) on the
fused sm_100a augmentation kernels (csrc/augment.cu behind pb200_augment / pb200_plasma).

Same constructor, same `forward(images, pixel_coordinates) -> (images, normalised coords)`; the batch must live on a
CUDA device (there is no CPU path: a CPU tensor raises).  What the reference does with ~60 kornia / torch launches
and a dozen full-size temporaries per batch is two launches here (+ one for the plasma map): the per-SAMPLE random
draws (donor image, affine matrix, erase rectangles, colour gains, operator order, blur sigma, shadow parameters) are
sampled on the host with a torch.Generator and handed to the kernel as two small tables; the per-PIXEL depth noise is
drawn inside the kernel (Philox4x32-10).  The distributions follow the kornia generators the reference configures
(augmentations.py:351-392); kornia itself is not needed.
"""
from __future__ import annotations

import ctypes
import functools
import math
from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np
import torch
from torch import nn

from .. import _lib

NUM_RGB_CHANNELS = 3
DEPTH_CHANNEL_INDEX = 3
IP, FP = 12, 32
AFFINE, PLANCKIAN, JIGGLE, BLUR, SHADOW = 1, 2, 4, 8, 16


@dataclass(frozen=True)
class AugmentationConfig:
    """Field for field the reference's AugmentationConfig (augmentations.py:259-324)."""

    cube_scale: float = 0.035
    random_transplantation_with_depth: bool = True
    random_affine: bool = True
    degrees: float = 90
    translate: Tuple[float, float] = (0.1, 0.1)
    scale: Tuple[float, float] = (0.9, 1.5)
    shear: float = 0.1
    random_erasing: bool = True
    planckian_jitter: bool = True
    color_jiggle: bool = True
    brightness: float = 0.2
    contrast: float = 0.4
    saturation: float = 0.4
    hue: float = 0.025
    blur: bool = True
    random_plasma_shadow: bool = True
    random_bias: bool = True
    dev_bias: float = 0.02
    p_bias: float = 0.5
    depth_gaussian_noise: bool = True
    std_gaussian_noise: float = 0.005
    random_near_plane: bool = True
    scaled_near_plane_mean: float = 0.1
    dev_near_plane: float = 0.05
    p_near_plane: float = 0.5
    near_value: float = 0.0
    random_far_plane: bool = True
    scaled_far_plane_mean: float = 0.5
    dev_far_plane: float = 0.05
    p_far_plane: float = 0.5
    far_value: float = 0.0


class AugConfigStruct(ctypes.Structure):
    """pb200_aug_config (include/perseus_b200.h)."""

    _fields_ = [
        ("cube_scale", ctypes.c_float), ("noise_cube_scale", ctypes.c_float),
        ("lb_seg_ratio", ctypes.c_float), ("ub_seg_ratio", ctypes.c_float),
        ("training", ctypes.c_int),
        ("depth_bias", ctypes.c_int), ("dev_bias", ctypes.c_float), ("p_bias", ctypes.c_float),
        ("depth_noise", ctypes.c_int), ("std_noise", ctypes.c_float),
        ("near_plane", ctypes.c_int), ("near_mean", ctypes.c_float), ("near_dev", ctypes.c_float),
        ("p_near", ctypes.c_float), ("near_value", ctypes.c_float),
        ("far_plane", ctypes.c_int), ("far_mean", ctypes.c_float), ("far_dev", ctypes.c_float),
        ("p_far", ctypes.c_float), ("far_value", ctypes.c_float),
    ]


def kernel_config(cfg: AugmentationConfig, train: bool, training: bool = True) -> dict:
    """The depth-stage / transplantation constants as KeypointAugmentation.__init__ wires them (:395-431)."""
    planes = cfg.random_near_plane or cfg.random_far_plane
    return dict(
        cube_scale=cfg.cube_scale, noise_cube_scale=0.035,  # DepthGaussianNoiseAugmentation(std) keeps its default scale
        lb_seg_ratio=0.02, ub_seg_ratio=0.7, training=int(training),
        depth_bias=int(cfg.random_bias and train), dev_bias=cfg.dev_bias, p_bias=cfg.p_bias,
        depth_noise=int(cfg.depth_gaussian_noise and train), std_noise=cfg.std_gaussian_noise,
        near_plane=int(planes and cfg.random_near_plane), near_mean=cfg.scaled_near_plane_mean,
        near_dev=cfg.dev_near_plane, p_near=cfg.p_near_plane if train else 0.0, near_value=cfg.near_value,
        far_plane=int(planes and cfg.random_far_plane), far_mean=cfg.scaled_far_plane_mean,
        far_dev=cfg.dev_far_plane, p_far=cfg.p_far_plane if train else 0.0, far_value=cfg.far_value,
    )


def config_struct(d: dict) -> AugConfigStruct:
    s = AugConfigStruct()
    for name, _ in AugConfigStruct._fields_:
        setattr(s, name, d[name])
    return s


# ---- geometry (kornia conventions) ----------------------------------------------------------------------------------
def affine_matrix(angle_deg, tx, ty, scale, shear_x_deg, shear_y_deg, H, W) -> np.ndarray:
    """RandomAffine's 3x3 pixel-space matrix: kornia get_affine_matrix2d(translations, centre ((W-1)/2, (H-1)/2), scale,
    angle, sx, sy) = [rotation by -angle with scale about the centre, + translation] @ shear-about-the-centre."""
    cx, cy = (W - 1) / 2.0, (H - 1) / 2.0
    a = math.radians(-angle_deg)
    al, be = scale * math.cos(a), scale * math.sin(a)
    T = np.array([[al, be, (1 - al) * cx - be * cy + tx], [-be, al, be * cx + (1 - al) * cy + ty], [0, 0, 1]])
    sx, sy = math.tan(math.radians(shear_x_deg)), math.tan(math.radians(shear_y_deg))
    Sh = np.array([[1, -sx, sx * cy], [-sy, 1 + sx * sy, sy * (cx - sx * cy)], [0, 0, 1]])
    return T @ Sh


def sampling_map(M: np.ndarray, H: int, W: int) -> np.ndarray:
    """Source position (pixels) of output pixel (x, y, 1) under kornia.warp_affine(align_corners=False): the matrix is
    normalised as if [-1, 1] spanned [0, size-1] while affine_grid / grid_sample use the half-pixel convention, i.e.
    src = D2 @ M^-1 @ D1 with D1: x -> (x + 0.5)(W-1)/W and D2: p -> p W/(W-1) - 0.5 (rows likewise with H)."""
    D1 = np.array([[(W - 1) / W, 0, 0.5 * (W - 1) / W], [0, (H - 1) / H, 0.5 * (H - 1) / H], [0, 0, 1]])
    D2 = np.array([[W / (W - 1), 0, -0.5], [0, H / (H - 1), -0.5], [0, 0, 1]])
    return D2 @ np.linalg.inv(M) @ D1


def affine_matrices(ang, tx, ty, sc, shx, shy, H, W) -> np.ndarray:
    """affine_matrix over arrays -> (B, 3, 3)."""
    B = len(ang)
    cx, cy = (W - 1) / 2.0, (H - 1) / 2.0
    a = np.radians(-np.asarray(ang, dtype=np.float64))
    al, be = sc * np.cos(a), sc * np.sin(a)
    T = np.zeros((B, 3, 3))
    T[:, 0, 0], T[:, 0, 1], T[:, 0, 2] = al, be, (1 - al) * cx - be * cy + tx
    T[:, 1, 0], T[:, 1, 1], T[:, 1, 2] = -be, al, be * cx + (1 - al) * cy + ty
    T[:, 2, 2] = 1
    sx, sy = np.tan(np.radians(shx)), np.tan(np.radians(shy))
    Sh = np.zeros((B, 3, 3))
    Sh[:, 0, 0], Sh[:, 0, 1], Sh[:, 0, 2] = 1, -sx, sx * cy
    Sh[:, 1, 0], Sh[:, 1, 1], Sh[:, 1, 2] = -sy, 1 + sx * sy, sy * (cx - sx * cy)
    Sh[:, 2, 2] = 1
    return T @ Sh


def sampling_maps(M: np.ndarray, H: int, W: int) -> np.ndarray:
    """sampling_map over a (B, 3, 3) stack."""
    D1 = np.array([[(W - 1) / W, 0, 0.5 * (W - 1) / W], [0, (H - 1) / H, 0.5 * (H - 1) / H], [0, 0, 1]])
    D2 = np.array([[W / (W - 1), 0, -0.5], [0, H / (H - 1), -0.5], [0, 0, 1]])
    return D2 @ np.linalg.inv(M) @ D1


def gaussian_taps(sigma: float, ksize: int = 5) -> np.ndarray:
    x = np.arange(ksize, dtype=np.float64) - ksize // 2
    g = np.exp(-(x ** 2) / (2.0 * sigma ** 2))
    return (g / g.sum()).astype(np.float32)


@functools.lru_cache(maxsize=1)
def planckian_table() -> np.ndarray:
    """(25, 2) red / blue gains relative to green for black bodies at 3000..15000 K (RandomPlanckianJitter
    mode='blackbody' picks one row per sample).  kornia's literal coefficient table is not available offline; these
    come from the usual piecewise fit of black-body chromaticity (T. Helland) — pass `planckian_table=` to
    KeypointAugmentation to substitute another table."""
    rows = []
    for T in range(3000, 15001, 500):
        t = T / 100.0
        r = 255.0 if t <= 66 else 329.698727446 * (t - 60) ** -0.1332047592
        g = 99.4708025861 * math.log(t) - 161.1195681661 if t <= 66 else 288.1221695283 * (t - 60) ** -0.0755148492
        b = 255.0 if t >= 66 else (0.0 if t <= 19 else 138.5177312231 * math.log(t - 10) - 305.0447927307)
        r, g, b = (min(max(v, 0.0), 255.0) for v in (r, g, b))
        rows.append((r / g, b / g))
    return np.asarray(rows, dtype=np.float32)


# ---- per-sample draws (the kornia parameter generators the reference configures) -----------------------------------
class _Pool:
    """All of a batch's per-sample uniforms in ONE generator call (rows are consumed in a fixed order)."""

    ROWS = 40

    def __init__(self, g: torch.Generator, n: int):
        self.u = torch.rand(self.ROWS, n, generator=g, dtype=torch.float64).numpy()
        self.i = 0

    def next(self) -> np.ndarray:
        self.i += 1
        return self.u[self.i - 1]


def _uniform(g: "_Pool", n: int, lo: float, hi: float) -> np.ndarray:
    return lo + (hi - lo) * g.next()


def _erase_rects(g, B, H, W, p, scale, ratio):
    """kornia RectangleEraseGenerator: area ~ U(scale) * H * W; aspect ~ U(ratio) (two-sided when ratio spans 1);
    h = round(sqrt(area * aspect)), w = round(sqrt(area / aspect)) clamped to [1, size]; corner ~ U * (size - len + 1)."""
    on = g.next() < p
    area = _uniform(g, B, scale[0], scale[1]) * H * W
    if ratio[0] < 1.0 < ratio[1]:
        a1, a2 = _uniform(g, B, ratio[0], 1.0), _uniform(g, B, 1.0, ratio[1])
        aspect = np.where(np.round(_uniform(g, B, 0.0, 1.0)).astype(bool), a1, a2)
    else:
        aspect = _uniform(g, B, ratio[0], ratio[1])
    h = np.clip(np.round(np.sqrt(area * aspect)), 1, H).astype(np.int64)
    w = np.clip(np.round(np.sqrt(area / aspect)), 1, W).astype(np.int64)
    x = (_uniform(g, B, 0.0, 1.0) * (W - w + 1)).astype(np.int64)
    y = (_uniform(g, B, 0.0, 1.0) * (H - h + 1)).astype(np.int64)
    w = np.where(on, w, 0)
    return np.stack([x, y, w, h], axis=1).astype(np.int32)


def sample_params(cfg: AugmentationConfig, train: bool, B: int, C: int, H: int, W: int, g: torch.Generator,
                  planck: Optional[np.ndarray] = None):
    """Returns (iparams (B, 12) int32, fparams (B, 32) float32, roughness (B,) float32) — layout in perseus_b200.h."""
    ip = np.zeros((B, IP), dtype=np.int32)
    fp = np.zeros((B, FP), dtype=np.float32)
    ip[:, 0] = -1
    ip[:, 9] = 0x3210
    fp[:, 0], fp[:, 4], fp[:, 6], fp[:, 10] = 1, 1, 1, 1  # identity matrices
    fp[:, 12:17] = 1
    rough = np.full(B, 0.5, dtype=np.float32)
    if not train:
        return ip, fp, rough
    g = _Pool(g, B)
    if cfg.random_transplantation_with_depth and C == 5 and B > 1:  # augmentations.py:201-208
        shift = 1 + np.minimum((g.next() * (B - 1)).astype(np.int64), B - 2)  # randint(1, B)
        ip[:, 0] = (np.arange(B) + shift) % B
    if cfg.random_affine:  # RandomAffine(degrees, translate, scale, shear), p = 0.5
        on = g.next() < 0.5
        ang = _uniform(g, B, -cfg.degrees, cfg.degrees)
        tx = _uniform(g, B, -cfg.translate[0] * W, cfg.translate[0] * W)
        ty = _uniform(g, B, -cfg.translate[1] * H, cfg.translate[1] * H)
        sc = _uniform(g, B, cfg.scale[0], cfg.scale[1])
        shx = _uniform(g, B, -cfg.shear, cfg.shear)  # a scalar shear gives x-shear only
        M = affine_matrices(ang, tx, ty, sc, shx, np.zeros(B), H, W)  # (B, 3, 3)
        S = sampling_maps(M, H, W)
        fp[on, 0:6] = M[on, :2].reshape(-1, 6)
        fp[on, 6:12] = S[on, :2].reshape(-1, 6)
        ip[:, 10] |= np.where(on, AFFINE, 0).astype(np.int32)
    if cfg.random_erasing:  # :361-372
        ip[:, 1:5] = _erase_rects(g, B, H, W, 0.5, (0.02, 0.1), (2.0, 3.0))
        ip[:, 5:9] = _erase_rects(g, B, H, W, 0.5, (0.02, 0.05), (0.8, 1.2))
    if cfg.planckian_jitter:  # RandomPlanckianJitter(mode="blackbody"), p = 0.5
        table = planckian_table() if planck is None else np.asarray(planck, dtype=np.float32)
        on = g.next() < 0.5
        idx = np.minimum((g.next() * len(table)).astype(np.int64), len(table) - 1)
        fp[:, 12:14] = table[idx]
        ip[:, 10] |= np.where(on, PLANCKIAN, 0).astype(np.int32)
    if cfg.color_jiggle:  # ColorJiggle(brightness, contrast, saturation, hue), p = 1, operator order = randperm(4)
        fp[:, 14] = _uniform(g, B, 1 - cfg.brightness, 1 + cfg.brightness)
        fp[:, 15] = _uniform(g, B, 1 - cfg.contrast, 1 + cfg.contrast)
        fp[:, 16] = _uniform(g, B, 1 - cfg.saturation, 1 + cfg.saturation)
        fp[:, 17] = _uniform(g, B, -cfg.hue, cfg.hue) * 2.0 * math.pi
        perm = np.argsort(np.stack([g.next() for _ in range(4)], axis=1), axis=1).astype(np.int32)  # uniform random permutation
        ip[:, 9] = perm[:, 0] | (perm[:, 1] << 4) | (perm[:, 2] << 8) | (perm[:, 3] << 12)
        ip[:, 10] |= JIGGLE
    if cfg.blur:  # RandomGaussianBlur((5, 5), (3.0, 8.0), p=0.5)
        on = g.next() < 0.5
        sig = _uniform(g, B, 3.0, 8.0)
        taps = np.exp(-((np.arange(5.0) - 2.0)[None, :] ** 2) / (2.0 * sig[:, None] ** 2))
        fp[:, 18:23] = taps / taps.sum(axis=1, keepdims=True)
        ip[:, 10] |= np.where(on, BLUR, 0).astype(np.int32)
    if cfg.random_plasma_shadow:  # RandomPlasmaShadow(): roughness (0.1, 0.7), intensity (-1, 0), quantity (0, 1), p = 0.5
        on = g.next() < 0.5
        rough = _uniform(g, B, 0.1, 0.7).astype(np.float32)
        fp[:, 23] = _uniform(g, B, -1.0, 0.0)
        fp[:, 24] = _uniform(g, B, 0.0, 1.0)
        ip[:, 10] |= np.where(on, SHADOW, 0).astype(np.int32)
    return ip, fp, rough


# ---- raw calls (also what the parity tests drive) -------------------------------------------------------------------
def plasma_map(roughness: torch.Tensor, seed: int, H: int, W: int) -> torch.Tensor:
    """(B, H, W) float32 diamond-square map in [0, 1] on roughness's device."""
    lib = _lib.lib()
    B = roughness.shape[0]
    S = lib.pb200_plasma_side(H, W)
    scratch = torch.empty(B, S, S, dtype=torch.float32, device=roughness.device)
    out = torch.empty(B, H, W, dtype=torch.float32, device=roughness.device)
    _lib.check(lib.pb200_plasma(roughness.data_ptr(), ctypes.c_uint64(seed), scratch.data_ptr(), out.data_ptr(), B, H, W,
                                _lib.current_stream_ptr()))
    return out


def augment_batch(images: torch.Tensor, iparams: torch.Tensor, fparams: torch.Tensor, plasma: Optional[torch.Tensor],
                  cfg: dict, seed: int, keypoints: Optional[torch.Tensor] = None, c_out: Optional[int] = None):
    """One pb200_augment call.  images (B, C, H, W) float32 cuda contiguous; tables on the same device."""
    if not images.is_cuda:
        raise RuntimeError("perseus_b200 augmentation runs on a CUDA device only (no CPU path)")
    lib = _lib.lib()
    images = images.contiguous().float()
    B, C, H, W = images.shape
    c_out = C if c_out is None else c_out
    out = torch.empty(B, c_out, H, W, dtype=torch.float32, device=images.device)
    counts = torch.empty(max(B, 1), dtype=torch.int32, device=images.device)
    kp_in = kp_out = None
    K = 0
    if keypoints is not None:
        kp_in = keypoints.contiguous().float().reshape(B, -1, 2)
        K = kp_in.shape[1]
        kp_out = torch.empty_like(kp_in)
    st = config_struct(cfg)
    _lib.check(lib.pb200_augment(images.data_ptr(), iparams.data_ptr(), fparams.data_ptr(), _lib.ptr(plasma),
                                 ctypes.byref(st), ctypes.c_uint64(seed), counts.data_ptr(), _lib.ptr(kp_in),
                                 out.data_ptr(), _lib.ptr(kp_out), B, C, c_out, H, W, K, _lib.current_stream_ptr()))
    return out, kp_out


class KeypointAugmentation(nn.Module):
    """Drop-in for the reference class (augmentations.py:327-488)."""

    def __init__(self, cfg: AugmentationConfig, train: bool = True, seed: int = 0,
                 planckian_table: Optional[np.ndarray] = None, out_channels: Optional[int] = None) -> None:
        """cfg, train: as the reference.  Additive: seed (host generator + Philox key), planckian_table, and
        out_channels = write only the first k channels (what train.py:284 slices afterwards anyway)."""
        super().__init__()
        self.out_channels = out_channels
        self.cfg = cfg
        self.train_mode = train  # the reference stores this as .train, shadowing nn.Module.train()
        self.generator = torch.Generator().manual_seed(seed)
        self.planck = planckian_table
        self._calls = 0
        self._seed = seed

    def forward(self, images: torch.Tensor, pixel_coordinates: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        leading = pixel_coordinates.shape[:-1]  # :447
        unbatched = images.dim() == 3
        x = images.unsqueeze(0) if unbatched else images
        B, C, H, W = x.shape
        kp = pixel_coordinates.reshape(B, -1, 2)
        ip, fp, rough = sample_params(self.cfg, self.train_mode, B, C, H, W, self.generator, self.planck)
        dev = x.device
        ip_d = torch.from_numpy(ip).to(dev, non_blocking=True)
        fp_d = torch.from_numpy(fp).to(dev, non_blocking=True)
        self._calls += 1
        seed = (self._seed * 0x9E3779B97F4A7C15 + self._calls) & 0xFFFFFFFFFFFFFFFF
        plasma = None
        if self.train_mode and self.cfg.random_plasma_shadow and (ip[:, 10] & SHADOW).any():
            plasma = plasma_map(torch.from_numpy(rough).to(dev), seed, H, W)
        out, kp_out = augment_batch(x, ip_d, fp_d, plasma, kernel_config(self.cfg, self.train_mode, self.training),
                                    seed, kp, self.out_channels)
        return out, kp_out.reshape(*leading, -1)  # :488
