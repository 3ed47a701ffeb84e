"""Pins the detector oracle (oracle/detector_oracle.py) — CPU only.

The reference has no detector tests, so the pins are (i) the REAL reference class
imported from /root/reference where it exists (this container), bit for bit, and (ii) the committed
fixture tests/golden/detector_golden.npz that oracle/make_golden.py produced from that class."""
import os

import numpy as np
import pytest
import torch

from oracle import detector_oracle as do
from oracle.make_golden import golden_input

GOLD = os.path.join(os.path.dirname(__file__), "golden", "detector_golden.npz")
HAVE_REF = os.path.isdir("/root/reference/perseus")


@pytest.mark.parametrize("C", [3, 4])
def test_restatement_matches_golden(C):
    g = np.load(GOLD)
    for bn in (0, 1):
        tag = f"c{C}_bn{bn}"
        sd = do.seeded_state_dict(C, 8, seed=0, randomize_bn=bool(bn))
        wsum = sum(float(v.double().abs().sum()) for v in sd.values())
        assert wsum == pytest.approx(float(g[f"{tag}_wsum"]), rel=1e-12), "seeded weights do not reproduce"
        x = golden_input(C)
        assert float(x.double().sum()) == pytest.approx(float(g[f"{tag}_xsum"]), rel=1e-12)
        y = do.reference_forward(sd, x, torch.float32, num_threads=8).numpy()
        # same torchvision code path; thread count may change oneDNN's reduction order
        assert np.abs(y - g[f"{tag}_out32"]).max() <= 2e-5 * max(1.0, np.abs(g[f"{tag}_out32"]).max())
        y64 = do.reference_forward(sd, x, torch.float64).numpy()
        assert np.abs(y64 - g[f"{tag}_out64"]).max() <= 1e-10


@pytest.mark.skipif(not HAVE_REF, reason="/root/reference only exists in the build container")
def test_restatement_is_the_reference_class():
    Ref = do.import_reference_keypoint_cnn()
    for C in (3, 4):
        sd = do.seeded_state_dict(C, 8, seed=0)
        ref = Ref(num_channels=C)
        assert list(ref.state_dict().keys()) == list(sd.keys())
        assert all(ref.state_dict()[k].shape == sd[k].shape for k in sd)
        ref.load_state_dict(sd)
        ref.eval()
        x = golden_input(C, B=1)
        torch.set_num_threads(8)
        with torch.no_grad():
            want = ref(x)
        got = do.reference_forward(sd, x, torch.float32, num_threads=8)
        assert torch.equal(want, got)
        assert (ref.n_keypoints, ref.num_channels, ref.H, ref.W) == (8, C, 256, 256)


def test_u8_normalisation_variants_agree():
    """streaming.py:69 (float64 divide, then .float()) == data.py:84 (float32 divide) for all bytes —
    the ingest kernel's fp32 k/255.0f division equals both."""
    k = np.arange(256, dtype=np.uint8)
    a = torch.from_numpy(k / 255.0).float().numpy()
    b = np.asarray(k, dtype=np.float32) / 255.0
    assert b.dtype == np.float32 and np.array_equal(a, b)


def test_streaming_ingest_shapes_and_values():
    rng = np.random.default_rng(0)
    bgra = rng.integers(0, 256, (376, 672, 4), dtype=np.uint8)
    depth = rng.uniform(0.1, 0.5, (376, 672)).astype(np.float32)
    depth[5, 7] = np.nan
    depth[200, 300] = np.inf
    depth[190, 330] = -np.inf
    depth[188, 336] = np.nan
    x = do.streaming_ingest(bgra, depth, 4)
    assert x.shape == (4, 256, 256) and x.dtype == torch.float32
    y0, x0 = 376 // 2 - 128, 672 // 2 - 128
    assert x[0, 0, 0] == np.float32(bgra[y0, x0, 2] / 255.0)  # R comes from BGRA[...,2]
    assert x[2, 10, 20] == np.float32(bgra[y0 + 10, x0 + 20, 0] / 255.0)
    assert x[3, 200 - y0, 300 - x0] == 0 and x[3, 190 - y0, 330 - x0] == 0 and x[3, 188 - y0, 336 - x0] == 0
    assert x[3, 1, 1] == np.float32(depth[y0 + 1, x0 + 1]) / np.float32(0.035)
    assert do.streaming_ingest(bgra, depth, 3).shape == (3, 256, 256)
    assert do.streaming_ingest(bgra, None, 3).shape == (3, 256, 256)


@pytest.mark.skipif(not HAVE_REF, reason="/root/reference only exists in the build container")
def test_depth_clip_matches_reference_class():
    """DepthPlaneAugmentation imported from the reference (kornia stubbed: the class only needs
    torch) with dev = 0 equals depth_plane_clip without jitter."""
    import sys
    import types

    if "kornia" not in sys.modules:
        k = types.ModuleType("kornia")
        k.augmentation = types.ModuleType("kornia.augmentation")
        k.geometry = types.ModuleType("kornia.geometry")
        k.geometry.conversions = types.ModuleType("kornia.geometry.conversions")
        k.augmentation.AugmentationSequential = object
        sys.modules.update({"kornia": k, "kornia.augmentation": k.augmentation, "kornia.geometry": k.geometry,
                            "kornia.geometry.conversions": k.geometry.conversions})
    do.import_reference_keypoint_cnn()  # puts /root/reference on sys.path
    try:
        from perseus.detector.augmentations import DepthPlaneAugmentation  # type: ignore
    except Exception as e:  # pragma: no cover
        pytest.skip(f"reference augmentations not importable here: {e}")
    aug = DepthPlaneAugmentation(near=True, near_mean=0.1, near_dev=0.0, p_near=0.0, far=True, far_mean=0.5,
                                 far_dev=0.0, p_far=0.0, cube_scale=0.035)
    g = torch.Generator().manual_seed(0)
    depth = torch.rand(2, 64, 64, generator=g) * 20.0
    assert torch.equal(aug(depth), do.depth_plane_clip(depth))


def test_denormalize_definition():
    xy = torch.tensor([[-1.0, -1.0], [1.0, 1.0], [0.0, 0.5]])
    px = do.denormalize_pixel_coordinates(xy, 256, 256)
    assert torch.allclose(px, torch.tensor([[0.0, 0.0], [255.0, 255.0], [127.5, 191.25]]))
    back = do.normalize_pixel_coordinates(px, 256, 256)
    assert torch.allclose(back, xy, atol=1e-6)


def test_bf16_emulation_tracks_fp32_reference():
    sd = do.seeded_state_dict(3, 8, seed=0)
    x = golden_input(3, B=1)
    ref = do.reference_forward(sd, x, torch.float32)
    emu = do.bf16_emulated_forward(sd, x)
    # bf16 noise on random-init weights: ~1e-2 relative, far from an indexing bug
    assert (emu - ref).abs().max() < 0.05 * ref.abs().max()


@pytest.mark.parametrize("C", [3, 4])
def test_bf16x3_emulation_meets_the_bf16_tolerance_on_the_golden_vectors(C):
    """The split-bf16 operand model (a_hi*w_hi + a_lo*w_hi + a_hi*w_lo) stays within north_star's 0.05 px of the
    reference's golden outputs — the bound the GPU bf16x3 mode is tested against."""
    g = np.load(GOLD)
    sd = do.seeded_state_dict(C, 8, seed=0, randomize_bn=True)
    x = golden_input(C)
    y = do.bf16x3_emulated_forward(sd, x)
    err = float(((y.double() - torch.from_numpy(g[f"c{C}_bn1_out32"]).double()).abs() * 127.5).max())
    assert err <= 0.05, err


def test_fp16x2_emulation_meets_the_bf16_mode_tolerance_on_rgb_golden():
    """The arithmetic of the "fp16x2" tcgen05 mode (fp16 activations x fp16 hi + lo weights, exact accumulation),
    emulated on the CPU, is within north_star's 0.05 px of the REAL reference's fp32 outputs (golden fixture) on RGB;
    on random-init RGBD (outputs up to 3100 px) it is not (0.25-0.30 px), which the GPU tests report in plain pixels."""
    import numpy as np

    from oracle.make_golden import golden_input

    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "detector_golden.npz"))
    for bn in (0, 1):
        sd = do.seeded_state_dict(3, 8, seed=0, randomize_bn=bool(bn))
        y = do.fp16x2_emulated_forward(sd, golden_input(3))
        err = float((y.double() - torch.from_numpy(g[f"c3_bn{bn}_out32"]).double()).abs().max() * 127.5)
        assert err <= 0.04, err
    sd = do.seeded_state_dict(4, 8, seed=0, randomize_bn=True)
    y = do.fp16x2_emulated_forward(sd, golden_input(4))
    err = float((y.double() - torch.from_numpy(g["c4_bn1_out32"]).double()).abs().max() * 127.5)
    assert 0.05 < err <= 0.4, err
