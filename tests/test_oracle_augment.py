"""Pins oracle/augment_oracle.py (CPU): the reference's own augmentation classes where they run without kornia
(transplantation + the three depth classes: golden fixture made from the REAL classes, and the classes themselves
when /root/reference exists), Random123's Philox known answers, and the internal consistency of the kornia-defined
geometry (sampling map == affine_grid/grid_sample definition; key points follow the image)."""
import os
import sys

import numpy as np
import pytest
import torch

from oracle import augment_oracle as ao
from perseus_b200.detector import augmentations as host

GOLD = os.path.join(os.path.dirname(__file__), "golden", "augment_golden.npz")
HAVE_REF = os.path.isdir("/root/reference/perseus")


def test_philox_known_answers():
    """Random123 kat_vectors: philox4x32 10 rounds."""
    kats = [
        ((0, 0, 0, 0), (0, 0), (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
        ((0xFFFFFFFF,) * 4, (0xFFFFFFFF,) * 2, (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
        ((0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344), (0xA4093822, 0x299F31D0),
         (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1)),
    ]
    for ctr, key, want in kats:
        got = ao.philox4x32_10(*(np.uint32([c]) for c in ctr), *key)
        assert tuple(int(g[0]) for g in got) == want


def test_transplant_and_depth_match_reference_golden():
    g = np.load(GOLD)
    images = torch.from_numpy(g["tp_images"])
    out, ok, counts = ao.transplant(images, torch.from_numpy(g["tp_donors"]))
    assert torch.equal(out, torch.from_numpy(g["tp_out"]))
    assert ok.any() and (counts >= 0).all()
    d = torch.from_numpy(g["bias_in"])
    assert torch.equal(ao.depth_bias(d, torch.from_numpy(g["bias_mask"]), torch.from_numpy(g["bias_u"])),
                       torch.from_numpy(g["bias_out"]))
    assert torch.equal(ao.depth_noise(d, torch.from_numpy(g["noise_z"])), torch.from_numpy(g["noise_out"]))
    got = ao.depth_planes(d, *(torch.from_numpy(g[k]) for k in ("pl_mn", "pl_un", "pl_mf", "pl_uf")))
    assert torch.equal(got, torch.from_numpy(g["pl_out"]))
    # the reference's dropout masks are {0, 1/(1-p)}: what dropout_mask builds from a uniform draw
    assert set(np.unique(g["bias_mask"]).tolist()) == {0.0, 2.0}
    u = torch.rand(1000, generator=torch.Generator().manual_seed(0))
    assert set(ao.dropout_mask(u, 0.5, True).unique().tolist()) == {0.0, 2.0}
    assert torch.equal(ao.dropout_mask(u, 0.5, False), torch.ones_like(u))
    assert torch.equal(ao.dropout_mask(u, 0.0, True), torch.ones_like(u))  # val mode: p = 0 (augmentations.py:419-431)


@pytest.mark.skipif(not HAVE_REF, reason="/root/reference only exists in the build container")
def test_transplant_matches_reference_class_live():
    from oracle.make_golden_augment import import_reference_augmentations, synthetic_batch

    ra = import_reference_augmentations()
    for seed, (B, H, W) in enumerate(((2, 16, 16), (5, 32, 24), (8, 20, 36))):
        images = synthetic_batch(B, H, W, seed=100 + seed)
        torch.manual_seed(seed)
        donors = (torch.arange(B) + torch.randint(1, B, (B,))) % B
        torch.manual_seed(seed)
        want = ra.RandomTransplantationWithDepth()(images.clone())
        got, _, _ = ao.transplant(images, donors)
        assert torch.equal(got, want)
    # 4-channel batches are returned untouched (augmentations.py:201-202)
    x4 = synthetic_batch(3, 8, 8, 0)[:, :4]
    assert torch.equal(ra.RandomTransplantationWithDepth()(x4), ao.transplant(x4, torch.tensor([1, 2, 0]))[0])


def _random_affine(rng, H, W):
    return dict(angle_deg=rng.uniform(-90, 90), tx=rng.uniform(-0.1 * W, 0.1 * W), ty=rng.uniform(-0.1 * H, 0.1 * H),
                scale=rng.uniform(0.9, 1.5), shear_x_deg=rng.uniform(-0.1, 0.1), shear_y_deg=0.0, H=H, W=W)


def test_sampling_map_is_the_affine_grid_definition():
    """The kernel's S-formulation == kornia.warp_affine's normalise -> invert -> affine_grid -> grid_sample chain."""
    rng = np.random.default_rng(0)
    for H, W in ((32, 32), (24, 40), (256, 256)):
        B = 3
        img = torch.rand(B, 5, H, W, generator=torch.Generator().manual_seed(1))
        Ms, Ss = [], []
        for _ in range(B):
            p = _random_affine(rng, H, W)
            M = host.affine_matrix(**p)
            np.testing.assert_allclose(M, ao.affine_matrix(**p), rtol=0, atol=1e-12)
            Ms.append(M)
            Ss.append(host.sampling_map(M, H, W)[:2].reshape(-1))
        a = ao.warp_affine_kornia(img, torch.from_numpy(np.stack(Ms)))
        b = ao.warp_direct(img, torch.from_numpy(np.stack(Ss)))
        assert float((a - b).abs().max()) < 2e-4


def test_keypoints_follow_the_image():
    """A bright dot at p must land at M @ p after the warp (direction convention of S vs M)."""
    H = W = 64
    rng = np.random.default_rng(2)
    for _ in range(5):
        p = _random_affine(rng, H, W)
        p["scale"] = 1.0
        M = host.affine_matrix(**p)
        src = np.array([rng.uniform(24, 40), rng.uniform(24, 40), 1.0])
        img = torch.zeros(1, 1, H, W)
        ys, xs = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
        img[0, 0] = torch.exp(-((xs - src[0]) ** 2 + (ys - src[1]) ** 2) / (2 * 1.5 ** 2))
        S = torch.from_numpy(host.sampling_map(M, H, W)[:2].reshape(1, 6))
        out = ao.warp_direct(img, S)[0, 0]
        w = out / out.sum()
        got = np.array([float((w * xs).sum()), float((w * ys).sum())])
        want = (M @ src)[:2]
        assert np.abs(got - want).max() < 0.6, (got, want)  # half-pixel convention quirk of kornia included


def test_hsv_round_trip_and_colour_ops():
    g = torch.Generator().manual_seed(0)
    rgb = torch.rand(3, 16, 16, generator=g)
    back = ao.hsv_to_rgb(ao.rgb_to_hsv(rgb))
    assert float((back - rgb).abs().max()) < 1e-5
    fp = np.zeros(32, dtype=np.float32)
    fp[12:17] = 1
    same = ao.colour_ops(rgb, ao.JIGGLE | ao.PLANCKIAN, 0x3210, fp)  # unit gains / zero hue shift: identity
    assert float((same - rgb).abs().max()) < 1e-5
    taps = ao.gaussian_taps(3.0)
    assert abs(float(taps.sum()) - 1) < 1e-6 and np.allclose(taps, host.gaussian_taps(3.0))
    flat = torch.full((3, 12, 12), 0.25)
    assert float((ao.blur5(flat, taps) - flat).abs().max()) < 1e-6


def test_plasma_oracle_properties():
    p = ao.plasma(np.float32([0.1, 0.5, 0.7]), seed=5, H=40, W=56)
    assert p.shape == (3, 40, 56) and p.dtype == np.float32
    assert np.allclose(p.min(axis=(1, 2)), 0) and np.allclose(p.max(axis=(1, 2)), 1)
    # rougher maps have more fine-scale energy
    tv = np.abs(np.diff(p, axis=2)).mean(axis=(1, 2))
    assert tv[0] < tv[1] < tv[2]
    assert np.array_equal(p, ao.plasma(np.float32([0.1, 0.5, 0.7]), seed=5, H=40, W=56))
    assert ao.plasma_side(256, 256) == 257 and ao.plasma_side(257, 100) == 257 and ao.plasma_side(258, 3) == 513


def test_host_sampler_ranges_and_val_mode():
    cfg = host.AugmentationConfig()
    g = torch.Generator().manual_seed(0)
    B, H, W = 64, 256, 256
    ip, fp, rough = host.sample_params(cfg, True, B, 5, H, W, g)
    assert ip.shape == (B, host.IP) and fp.shape == (B, host.FP) and ip.dtype == np.int32 and fp.dtype == np.float32
    donors = ip[:, 0]
    assert ((donors >= 0) & (donors < B) & (donors != np.arange(B))).all()  # a donor differs from its acceptor (:207)
    for e in (0, 1):
        x, y, w, h = (ip[:, 1 + 4 * e + k] for k in range(4))
        on = w > 0
        assert (x[on] >= 0).all() and (x[on] + w[on] <= W).all() and (y[on] + h[on] <= H).all()
    assert 0.2 < (ip[:, 10] & host.AFFINE).astype(bool).mean() < 0.8
    for b in range(B):
        assert sorted((ip[b, 9] >> (4 * k)) & 15 for k in range(4)) == [0, 1, 2, 3]
    assert (fp[:, 14] >= 0.8).all() and (fp[:, 14] <= 1.2).all() and (np.abs(fp[:, 17]) <= 0.025 * 2 * np.pi + 1e-6).all()
    assert np.allclose(fp[:, 18:23].sum(axis=1), 1, atol=1e-5)
    assert (rough >= 0.1).all() and (rough <= 0.7).all() and (fp[:, 23] <= 0).all() and (fp[:, 23] >= -1).all()
    # val mode: nothing but the deterministic-probability depth planes + normalisation (augmentations.py:419-431)
    ipv, fpv, _ = host.sample_params(cfg, False, 4, 4, H, W, g)
    assert (ipv[:, 10] == 0).all() and (ipv[:, 0] == -1).all() and (ipv[:, [3, 7]] == 0).all()
    kc = host.kernel_config(cfg, train=False)
    assert kc["depth_bias"] == 0 and kc["depth_noise"] == 0 and kc["near_plane"] == 1 and kc["p_near"] == 0.0
    assert host.planckian_table().shape == (25, 2)


def test_oracle_pipeline_identity_and_kornia_warp_agree():
    """Full oracle pipeline: identity parameters leave a 3-channel batch unchanged; S- and M-driven warps agree."""
    B, H, W = 3, 32, 32
    from oracle.make_golden_augment import synthetic_batch

    x = synthetic_batch(B, H, W, 5)
    cfg = host.kernel_config(host.AugmentationConfig(random_near_plane=False, random_far_plane=False), train=False)
    ip, fp, _ = host.sample_params(host.AugmentationConfig(), False, B, 5, H, W, torch.Generator().manual_seed(0))
    kp = torch.rand(B, 8, 2) * (W - 1)
    out, kpn = ao.augment(x, ip, fp, None, cfg, seed=1, keypoints=kp)
    assert torch.equal(out, x)
    assert torch.allclose(kpn, 2 * kp / (W - 1) - 1, atol=1e-6)
    ip2, fp2, _ = host.sample_params(host.AugmentationConfig(), True, B, 5, H, W, torch.Generator().manual_seed(3))
    ip2[:, 10] |= host.AFFINE
    for b in range(B):
        M = host.affine_matrix(30.0 * (b + 1), 2.0, -1.0, 1.1, 0.05, 0.0, H, W)
        fp2[b, 0:6] = M[:2].reshape(-1)
        fp2[b, 6:12] = host.sampling_map(M, H, W)[:2].reshape(-1)
    ip2[:, 10] &= ~(host.BLUR | host.SHADOW)
    a, _ = ao.augment(x, ip2, fp2, None, cfg, 1, kp, kornia_warp=False)
    b_, _ = ao.augment(x, ip2, fp2, None, cfg, 1, kp, kornia_warp=True)
    # colour operators are discontinuous in places (hue wrap, first-max ties): compare where both are smooth
    assert float((a - b_).abs().median()) < 1e-5 and float(((a - b_).abs() > 1e-2).float().mean()) < 5e-3


def test_batched_host_geometry_equals_scalar():
    rng = np.random.default_rng(4)
    B, H, W = 7, 48, 80
    ps = [_random_affine(rng, H, W) for _ in range(B)]
    M = host.affine_matrices(*(np.array([p[k] for p in ps]) for k in ("angle_deg", "tx", "ty", "scale", "shear_x_deg",
                                                                      "shear_y_deg")), H, W)
    S = host.sampling_maps(M, H, W)
    for b, p in enumerate(ps):
        np.testing.assert_allclose(M[b], host.affine_matrix(**p), atol=1e-12)
        np.testing.assert_allclose(S[b], host.sampling_map(M[b], H, W), atol=1e-12)
