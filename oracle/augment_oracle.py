"""CPU oracle for the train-time augmentation path.  TEST INFRASTRUCTURE ONLY: imported by
tests/ (and nothing on the product path); the product is perseus_b200/csrc/augment.cu.

What it restates (reference = /root/reference, perseus/detector/augmentations.py):
  * transplant                 RandomTransplantationWithDepth.forward :189-251, with the donor indices (drawn at :207
                               by torch.randint) made an explicit argument.  PINNED: tests/test_oracle_augment.py runs
                               the real reference class (kornia stubbed) under a seed, replays its randint draw, and
                               compares bit for bit; golden fixture tests/golden/augment_golden.npz.
  * depth_bias / depth_noise / depth_planes
                               DepthBiasAugmentation :42-50, DepthGaussianNoiseAugmentation :76-78,
                               DepthPlaneAugmentation :137-168, with the per-pixel draws (F.dropout mask, rand, randn)
                               made explicit arguments.  PINNED against the real classes by replaying their torch draws.
  * kornia operators           RandomAffine / RandomErasing / RandomPlanckianJitter / ColorJiggle / RandomGaussianBlur /
                               RandomPlasmaShadow as configured at :351-392 and normalize_pixel_coordinates :481.
                               kornia (>=0.7.1, pyproject.toml:20) is ABSENT from this image and from /root/reference, and
                               the reference has no test of its augmentations -> these follow kornia's published
                               definitions (warp_affine = normalize_homography + F.affine_grid + F.grid_sample with
                               align_corners=False; adjust_brightness / adjust_contrast / adjust_saturation / adjust_hue;
                               rgb_to_hsv / hsv_to_rgb; gaussian_blur2d with 'reflect' border; diamond_square) and are
                               PARITY UNPINNED by reference vectors.
  * philox4x32_10              Salmon et al. 2011 (Random123 known-answer vectors checked in the test); the kernel's
                               per-pixel noise source, so kernel-vs-oracle comparisons are draw-for-draw.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F

IP, FP = 12, 32  # PB200_AUG_IPARAMS / PB200_AUG_FPARAMS (include/perseus_b200.h)
AFFINE, PLANCKIAN, JIGGLE, BLUR, SHADOW = 1, 2, 4, 8, 16


# ---- Philox4x32-10 -------------------------------------------------------------------------------------------------
def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised over numpy uint32 arrays (counters); keys are python ints.  Returns 4 uint32 arrays."""
    M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) for c in np.broadcast_arrays(c0, c1, c2, c3))
    mask = np.uint64(0xFFFFFFFF)
    k0, k1 = int(k0) & 0xFFFFFFFF, int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & mask, p1 >> np.uint64(32), p1 & mask
        c0, c1, c2, c3 = hi1 ^ c1 ^ np.uint64(k0), lo1, hi0 ^ c3 ^ np.uint64(k1), lo0
        k0 = (k0 + 0x9E3779B9) & 0xFFFFFFFF
        k1 = (k1 + 0xBB67AE85) & 0xFFFFFFFF
    return tuple(c.astype(np.uint32) for c in (c0, c1, c2, c3))


def u01(w: np.ndarray) -> np.ndarray:
    """24-bit uniform in [0, 1) as float32 (torch.rand's float32 construction)."""
    return (w >> np.uint32(8)).astype(np.float32) * np.float32(2.0 ** -24)


def pixel_words(seed: int, B: int, H: int, W: int, stream: int):
    pix = np.arange(H * W, dtype=np.uint32)[None, :]
    b = np.arange(B, dtype=np.uint32)[:, None]
    w = philox4x32_10(pix, b, np.uint32(stream), np.uint32(0), seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    return [x.reshape(B, H, W) for x in w]


# ---- RandomTransplantationWithDepth (augmentations.py:189-251) -----------------------------------------------------
def transplant(images: torch.Tensor, donor_indices: torch.Tensor, lb: float = 0.02, ub: float = 0.7):
    """Returns (new_images, accepted (B,) bool, counts (B,) int).  donor_indices < 0: sample left alone."""
    if images.shape[-3] != 5 or images.dim() <= 3:  # :201-202
        B = images.shape[0] if images.dim() > 3 else 1
        return images, torch.zeros(B, dtype=torch.bool), torch.zeros(B, dtype=torch.int64)
    use = donor_indices >= 0
    di = torch.where(use, donor_indices, torch.arange(images.shape[0]))
    rgb, depth, seg = images[:, :3], images[:, 3], images[:, 4]  # :204-206
    donor = images[di]  # :208
    acc_cube = seg == 1.0  # :217
    masks = ~acc_cube  # :218
    d_acc = depth * acc_cube  # :222
    d_don = donor[:, 3] * acc_cube  # :223
    masks = masks | (d_don < d_acc)  # :224-225
    don_cube = donor[:, 4] == 1.0  # :228
    masks = torch.where(don_cube, torch.zeros_like(masks), masks)  # :229
    new_rgb = torch.where(masks.unsqueeze(1), donor[:, :3], rgb)  # :232-234
    new_depth = torch.where(masks, donor[:, 3], depth)[:, None]  # :235-237
    new_seg = (1.0 - masks.float()).unsqueeze(1)  # :238
    new_seg = torch.where(don_cube & ~acc_cube, torch.zeros_like(new_seg[:, 0]), new_seg[:, 0])[:, None]  # :239-243
    new = torch.cat([new_rgb, new_depth, new_seg], dim=1)  # :244
    ratios = torch.mean(new_seg, dim=(-2, -1)).squeeze(-1)  # :247
    ok = (ratios >= lb) & (ratios <= ub) & use  # :248
    counts = torch.where(use, new_seg.sum(dim=(-3, -2, -1)).round().long(), torch.zeros(images.shape[0], dtype=torch.int64))
    return torch.where(ok.view(-1, 1, 1, 1), new, images), ok, counts  # :249


# ---- depth operators with explicit draws (augmentations.py:17-169) --------------------------------------------------
def dropout_mask(u: torch.Tensor, p: float, training: bool) -> torch.Tensor:
    """F.dropout(ones, p, training) given the uniform draw: 0 with probability p else 1/(1-p)."""
    if not training or p <= 0.0:
        return torch.ones_like(u)
    if p >= 1.0:
        return torch.zeros_like(u)
    keep = torch.tensor(1.0, dtype=torch.float32) / (torch.tensor(1.0, dtype=torch.float32) - torch.tensor(p, dtype=torch.float32))
    return torch.where(u >= p, keep, torch.zeros_like(u))


def depth_bias(d, mask, u, dev=0.02, cube_scale=0.035):
    s = cube_scale * d  # :42
    bias = dev * mask * 2 * (u - 0.5)  # :46
    return (s + bias) / cube_scale  # :49-50


def depth_noise(d, z, std=0.005, cube_scale=0.035):
    s = cube_scale * d  # :76
    return (s + std * z) / cube_scale  # :77-78


def depth_planes(d, mask_n, u_n, mask_f, u_f, near=True, near_mean=0.1, near_dev=0.05, near_value=0.0, far=True,
                 far_mean=0.5, far_dev=0.05, far_value=0.0, cube_scale=0.035):
    s = cube_scale * d  # :137
    if near:
        plane = near_mean + near_dev * mask_n * 2 * (u_n - 0.5)  # :145-148
        s = torch.where(s < plane, near_value * torch.ones_like(s), s)  # :149-152
    if far:
        plane = far_mean + far_dev * mask_f * 2 * (u_f - 0.5)  # :158-161
        s = torch.where(s > plane, far_value * torch.ones_like(s), s)  # :162-165
    return s / cube_scale  # :167


def box_muller(u1: np.ndarray, u2: np.ndarray) -> np.ndarray:
    """z = sqrt(-2 ln(1 - u1)) cos(2 pi u2), float32 (the kernel's normal draw)."""
    one, two_pi = np.float32(1.0), np.float32(6.283185307179586)
    return (np.sqrt(np.float32(-2.0) * np.log(one - u1)) * np.cos(two_pi * u2)).astype(np.float32)


def depth_ops(d: torch.Tensor, cfg: dict, seed: int) -> torch.Tensor:
    """The kernel's depth stage: Philox streams 0 (bias mask, bias u, normal u1, u2) and 1 (near mask, near u, far
    mask, far u) per (pixel, image)."""
    B, H, W = d.shape
    w0 = [torch.from_numpy(u01(w)) for w in pixel_words(seed, B, H, W, 0)]
    w1 = [torch.from_numpy(u01(w)) for w in pixel_words(seed, B, H, W, 1)]
    tr = bool(cfg["training"])
    if cfg["depth_bias"]:
        d = depth_bias(d, dropout_mask(w0[0], cfg["p_bias"], tr), w0[1], cfg["dev_bias"], cfg["cube_scale"])
    if cfg["depth_noise"]:
        z = torch.from_numpy(box_muller(w0[2].numpy(), w0[3].numpy()))
        d = depth_noise(d, z, cfg["std_noise"], cfg["noise_cube_scale"])
    if cfg["near_plane"] or cfg["far_plane"]:
        d = depth_planes(d, dropout_mask(w1[0], cfg["p_near"], tr), w1[1], dropout_mask(w1[2], cfg["p_far"], tr), w1[3],
                         bool(cfg["near_plane"]), cfg["near_mean"], cfg["near_dev"], cfg["near_value"],
                         bool(cfg["far_plane"]), cfg["far_mean"], cfg["far_dev"], cfg["far_value"], cfg["cube_scale"])
    return d


# ---- kornia geometry: RandomAffine -> warp_affine -------------------------------------------------------------------
def affine_matrix(angle_deg, tx, ty, scale, shear_x_deg, shear_y_deg, H, W) -> np.ndarray:
    """kornia.geometry.get_affine_matrix2d as RandomAffine calls it (float64, 3x3, pixel coordinates):
    rotation by -angle about the centre ((W-1)/2, (H-1)/2) with isotropic scale, + translation, then @ shear."""
    cx, cy = (W - 1) / 2.0, (H - 1) / 2.0
    a = math.radians(-angle_deg)  # get_affine_matrix2d passes -angle to get_rotation_matrix2d
    # get_rotation_matrix2d (OpenCV convention): [[s cos, s sin], [-s sin, s cos]] about the centre
    al, be = scale * math.cos(a), scale * math.sin(a)
    T = np.array([[al, be, (1 - al) * cx - be * cy], [-be, al, be * cx + (1 - al) * cy], [0, 0, 1]], dtype=np.float64)
    T[0, 2] += tx
    T[1, 2] += ty
    sx, sy = math.tan(math.radians(shear_x_deg)), math.tan(math.radians(shear_y_deg))
    # get_shear_matrix2d(center, sx, sy)
    Sh = np.array([[1, -sx, sx * cy], [-sy, 1 + sx * sy, sy * (cx - sx * cy)], [0, 0, 1]], dtype=np.float64)
    return T @ Sh


def warp_affine_kornia(img: torch.Tensor, M: torch.Tensor) -> torch.Tensor:
    """kornia.geometry.warp_affine(img, M[:, :2], (H, W), 'bilinear', 'zeros', align_corners=False): the homography is
    normalised with normal_transform_pixel ([0, size-1] <-> [-1, 1]), inverted, fed to affine_grid + grid_sample."""
    B, _, H, W = img.shape
    N = torch.tensor([[2.0 / (W - 1), 0, -1], [0, 2.0 / (H - 1), -1], [0, 0, 1]], dtype=torch.float64)
    dst_norm_trans_src_norm = N @ M.double() @ torch.linalg.inv(N)
    src_norm_trans_dst_norm = torch.linalg.inv(dst_norm_trans_src_norm)
    grid = F.affine_grid(src_norm_trans_dst_norm[:, :2].to(img.dtype), [B, img.shape[1], H, W], align_corners=False)
    return F.grid_sample(img, grid, mode="bilinear", padding_mode="zeros", align_corners=False)


def warp_direct(img: torch.Tensor, S: torch.Tensor) -> torch.Tensor:
    """Bilinear gather at source position S @ (x, y, 1) (pixels), zeros outside — the kernel's formulation."""
    B, C, H, W = img.shape
    ys, xs = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    S = S.to(torch.float32).view(B, 6, 1, 1)
    sx = S[:, 0] * xs + (S[:, 1] * ys + S[:, 2])
    sy = S[:, 3] * xs + (S[:, 4] * ys + S[:, 5])
    fx, fy = torch.floor(sx), torch.floor(sy)
    wx1, wy1 = sx - fx, sy - fy
    wx0, wy0 = 1 - wx1, 1 - wy1
    out = torch.zeros_like(img)
    bidx = torch.arange(B).view(B, 1, 1).expand(B, H, W)
    for dy, dx, w in ((0, 0, wx0 * wy0), (0, 1, wx1 * wy0), (1, 0, wx0 * wy1), (1, 1, wx1 * wy1)):
        yy, xx = fy + dy, fx + dx
        ok = (yy >= 0) & (yy < H) & (xx >= 0) & (xx < W)
        yi, xi = yy.clamp(0, H - 1).long(), xx.clamp(0, W - 1).long()
        v = img[bidx, :, yi, xi].permute(0, 3, 1, 2)  # (B, C, H, W)
        out = out + v * (w * ok).unsqueeze(1)
    return out


# ---- kornia colour operators -----------------------------------------------------------------------------------------
TWO_PI = 2.0 * math.pi


def rgb_to_hsv(img: torch.Tensor, eps: float = 1e-8) -> torch.Tensor:
    mx, arg = img.max(-3)
    mn = img.min(-3)[0]
    delta = mx - mn
    v = mx
    s = delta / (mx + eps)
    delta = torch.where(delta == 0, torch.ones_like(delta), delta)
    rc, gc, bc = torch.unbind(mx.unsqueeze(-3) - img, dim=-3)
    h = torch.stack([bc - gc, (rc - bc) + 2.0 * delta, (gc - rc) + 4.0 * delta], dim=-3) / delta.unsqueeze(-3)
    # first maximum wins ties (r, then g, then b)
    r, g = img[..., 0, :, :], img[..., 1, :, :]
    arg = torch.where(r == mx, torch.zeros_like(arg), torch.where(g == mx, torch.ones_like(arg), 2 * torch.ones_like(arg)))
    h = torch.gather(h, -3, arg.unsqueeze(-3)).squeeze(-3)
    h = (h / 6.0) % 1.0
    return torch.stack([TWO_PI * h, s, v], dim=-3)


def hsv_to_rgb(img: torch.Tensor) -> torch.Tensor:
    h = img[..., 0, :, :] / TWO_PI
    s, v = img[..., 1, :, :], img[..., 2, :, :]
    hi = torch.floor(h * 6) % 6
    f = ((h * 6) % 6) - hi
    p, q, t = v * (1 - s), v * (1 - f * s), v * (1 - (1 - f) * s)
    hi = hi.long()
    table = torch.stack([v, q, p, p, t, v, t, v, v, q, p, p, p, p, t, v, v, q], dim=-3)
    idx = torch.stack([hi, hi + 6, hi + 12], dim=-3)
    return torch.gather(table, -3, idx)


def colour_ops(rgb: torch.Tensor, flags: int, order: int, fp: np.ndarray) -> torch.Tensor:
    """One sample (3, H, W): Planckian gains, then ColorJiggle's operators in the sampled order."""
    if flags & PLANCKIAN:
        gains = torch.tensor([fp[12], 1.0, fp[13]], dtype=torch.float32).view(3, 1, 1)
        rgb = (rgb * gains).clamp(max=1.0)
    if flags & JIGGLE:
        for k in range(4):
            op = (order >> (4 * k)) & 15
            if op == 0:
                rgb = (rgb + (np.float32(fp[14]) - np.float32(1.0))).clamp(0, 1)
            elif op == 1:
                rgb = (rgb * np.float32(fp[15])).clamp(0, 1)
            elif op == 2:
                hsv = rgb_to_hsv(rgb)
                hsv = torch.stack([hsv[0], (hsv[1] * np.float32(fp[16])).clamp(0, 1), hsv[2]])
                rgb = hsv_to_rgb(hsv)
            else:
                hsv = rgb_to_hsv(rgb)
                hsv = torch.stack([torch.fmod(hsv[0] + np.float32(fp[17]), TWO_PI), hsv[1], hsv[2]])
                rgb = hsv_to_rgb(hsv)
    return rgb


def gaussian_taps(sigma: float, ksize: int = 5) -> np.ndarray:
    """kornia.filters.get_gaussian_kernel1d."""
    x = np.arange(ksize, dtype=np.float64) - ksize // 2
    g = np.exp(-(x ** 2) / (2.0 * sigma ** 2))
    return (g / g.sum()).astype(np.float32)


def blur5(rgb: torch.Tensor, taps: np.ndarray) -> torch.Tensor:
    """gaussian_blur2d((5, 5), separable, border_type='reflect') on one sample (3, H, W)."""
    k = torch.from_numpy(np.asarray(taps, dtype=np.float32))
    x = F.pad(rgb.unsqueeze(0), (2, 2, 2, 2), mode="reflect")
    x = F.conv2d(x, k.view(1, 1, 1, 5).expand(3, 1, 1, 5), groups=3)
    x = F.conv2d(x, k.view(1, 1, 5, 1).expand(3, 1, 5, 1), groups=3)
    return x[0]


# ---- the whole pipeline the kernel fuses (KeypointAugmentation.forward :434-488) -----------------------------------
def augment(images: torch.Tensor, iparams: np.ndarray, fparams: np.ndarray, plasma, cfg: dict, seed: int,
            keypoints=None, c_out: int | None = None, kornia_warp: bool = False):
    """images (B, C, H, W) float32; iparams (B, 12) int32; fparams (B, 32) float32 (layout: include/perseus_b200.h).
    kornia_warp=True evaluates the warp through affine_grid/grid_sample from the forward matrix M (the semantic
    definition) instead of the sampling map S (the kernel's formulation)."""
    B, C, H, W = images.shape
    c_out = C if c_out is None else c_out
    ip, fp = np.asarray(iparams), np.asarray(fparams, dtype=np.float32)
    x = images.clone()
    if C == 5:  # :465-466
        x, _, _ = transplant(x, torch.from_numpy(ip[:, 0].astype(np.int64)), cfg["lb_seg_ratio"], cfg["ub_seg_ratio"])
    aff = torch.from_numpy((ip[:, 10] & AFFINE) != 0)
    if aff.any():  # :468-469 RandomAffine (only the selected samples)
        if kornia_warp:
            M = torch.eye(3, dtype=torch.float64).repeat(B, 1, 1)
            M[:, :2] = torch.from_numpy(fp[:, 0:6].astype(np.float64)).view(B, 2, 3)
            warped = warp_affine_kornia(x, M)
        else:
            warped = warp_direct(x, torch.from_numpy(fp[:, 6:12]))
        x = torch.where(aff.view(B, 1, 1, 1), warped, x)
    for e in range(2):  # RandomErasing x2, value 0, every channel
        for b in range(B):
            ex, ey, ew, eh = (int(v) for v in ip[b, 1 + 4 * e:5 + 4 * e])
            if ew > 0:
                x[b, :, ey:ey + eh, ex:ex + ew] = 0.0
    out = x.clone()
    for b in range(B):  # :472-473 RGB operators
        flags, order = int(ip[b, 10]), int(ip[b, 9])
        rgb = colour_ops(x[b, :3], flags, order, fp[b])
        if flags & BLUR:
            rgb = blur5(rgb, fp[b, 18:23])
        if (flags & SHADOW) and plasma is not None:
            shade = (plasma[b] < np.float32(fp[b, 24])).to(torch.float32) * np.float32(fp[b, 23])
            rgb = (rgb + shade.unsqueeze(0)).clamp(0, 1)
        out[b, :3] = rgb
    if C > 3:  # :476-477 depth operators
        out[:, 3] = depth_ops(out[:, 3].contiguous(), cfg, seed)
    kp_out = None
    if keypoints is not None:  # kornia maps key points by the forward matrix; :481 normalises
        kp = keypoints.clone().to(torch.float32).reshape(B, -1, 2)
        M = torch.from_numpy(fp[:, 0:6]).view(B, 2, 3)
        moved = torch.stack([M[:, 0, 0:1] * kp[..., 0] + (M[:, 0, 1:2] * kp[..., 1] + M[:, 0, 2:3]),
                             M[:, 1, 0:1] * kp[..., 0] + (M[:, 1, 1:2] * kp[..., 1] + M[:, 1, 2:3])], dim=-1)
        kp = torch.where(aff.view(B, 1, 1), moved, kp)
        hw = torch.tensor([W, H], dtype=torch.float32)
        factor = torch.tensor(2.0) / (hw - 1).clamp(1e-8)
        kp_out = factor * kp - 1
    return out[:, :c_out], kp_out


# ---- diamond-square plasma (kornia.contrib.diamond_square as RandomPlasmaShadow uses it) ---------------------------
def plasma_side(H: int, W: int) -> int:
    s = 2
    while s + 1 < max(H, W):
        s <<= 1
    return s + 1


def plasma(roughness: np.ndarray, seed: int, H: int, W: int) -> np.ndarray:
    """(B, H, W) float32 in [0, 1]: 3x3 uniform [-1, 1) seed grid, then diamond / square refinements with the noise
    amplitude multiplied by the roughness at each level; min-max normalised over the H x W crop.  Noise of grid point
    (y, x) of image b: 2 * u01(Philox(seed; y*S+x, b, 2).word0) - 1."""
    B = len(roughness)
    S = plasma_side(H, W)
    pt = np.arange(S * S, dtype=np.uint32)[None, :]
    bb = np.arange(B, dtype=np.uint32)[:, None]
    w = philox4x32_10(pt, bb, np.uint32(2), np.uint32(0), seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)[0]
    noise = (np.float32(2.0) * u01(w) - np.float32(1.0)).reshape(B, S, S)
    g = np.zeros((B, S, S), dtype=np.float32)
    st0 = (S - 1) // 2
    g[:, ::st0, ::st0] = noise[:, ::st0, ::st0]
    amp = np.ones(B, dtype=np.float32)
    rough = np.asarray(roughness, dtype=np.float32)
    st = st0
    while st >= 2:
        half = st // 2
        amp = amp * rough
        a = amp[:, None, None]
        c = g[:, ::st, ::st]
        mean = np.float32(0.25) * ((c[:, :-1, :-1] + c[:, :-1, 1:]) + (c[:, 1:, :-1] + c[:, 1:, 1:]))
        g[:, half::st, half::st] = mean + a * noise[:, half::st, half::st]
        for rows, cols in ((slice(0, S, st), slice(half, S, st)), (slice(half, S, st), slice(0, S, st))):
            ys, xs = np.arange(S)[rows], np.arange(S)[cols]
            Y, X = np.meshgrid(ys, xs, indexing="ij")
            total = np.zeros((B,) + Y.shape, dtype=np.float32)
            cnt = np.zeros(Y.shape, dtype=np.float32)
            for dy, dx in ((-half, 0), (0, -half), (0, half), (half, 0)):  # the kernel's summation order
                yy, xx = Y + dy, X + dx
                ok = (yy >= 0) & (yy < S) & (xx >= 0) & (xx < S)
                v = g[:, np.clip(yy, 0, S - 1), np.clip(xx, 0, S - 1)]
                total = total + np.where(ok, v, np.float32(0.0)).astype(np.float32)
                cnt = cnt + ok.astype(np.float32)
            g[:, Y, X] = total / cnt + a * noise[:, Y, X]
        st = half
    crop = g[:, :H, :W]
    mn, mx = crop.min(axis=(1, 2), keepdims=True), crop.max(axis=(1, 2), keepdims=True)
    rng = mx - mn
    return np.where(rng > 0, (crop - mn) / np.where(rng > 0, rng, 1), 0).astype(np.float32)
