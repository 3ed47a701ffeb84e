"""Parity of the fused augmentation kernels (csrc/augment.cu, through the C ABI) against oracle/augment_oracle.py on the
same explicit per-sample parameters and the same Philox streams.  Reference: perseus/detector/augmentations.py."""
import numpy as np
import pytest
import torch

from oracle import augment_oracle as ao
from oracle.make_golden_augment import synthetic_batch
from perseus_b200.detector import augmentations as host

pytestmark = pytest.mark.gpu

ALL_OFF = dict(random_transplantation_with_depth=False, random_affine=False, random_erasing=False, planckian_jitter=False,
               color_jiggle=False, blur=False, random_plasma_shadow=False, random_bias=False, depth_gaussian_noise=False,
               random_near_plane=False, random_far_plane=False)


def _run(x, ip, fp, plasma, kcfg, seed, kp=None, c_out=None):
    dev = torch.device("cuda", 0)
    out, kpo = host.augment_batch(x.to(dev), torch.from_numpy(ip).to(dev), torch.from_numpy(fp).to(dev),
                                  None if plasma is None else plasma.to(dev), kcfg, seed,
                                  None if kp is None else kp.to(dev), c_out)
    torch.cuda.synchronize()
    return out.cpu(), None if kpo is None else kpo.cpu()


def _close(got, want, atol=2e-5, outliers=0.0, rtol=2e-5):
    d = (got - want).abs()
    bad = float((d > atol + rtol * want.abs()).float().mean())
    assert bad <= outliers, (bad, float(d.max()))


def _params(cfg, B, C, H, W, seed):
    return host.sample_params(cfg, True, B, C, H, W, torch.Generator().manual_seed(seed))


def test_plasma_matches_oracle():
    dev = torch.device("cuda", 0)
    for (H, W), seed in (((256, 256), 1), ((40, 56), 2), ((100, 130), 3)):
        rough = np.float32([0.1, 0.35, 0.7, 0.5])
        got = host.plasma_map(torch.from_numpy(rough).to(dev), seed, H, W).cpu().numpy()
        want = ao.plasma(rough, seed, H, W)
        assert np.abs(got - want).max() <= 1e-6


def test_transplantation_only_is_exact():
    cfg = host.AugmentationConfig(**{**ALL_OFF, "random_transplantation_with_depth": True})
    for (B, H, W), seed in (((6, 24, 20), 0), ((16, 64, 64), 1), ((3, 33, 70), 2)):
        x = synthetic_batch(B, H, W, seed)
        x[0, 4] = 0.0  # no cube: ratio 0 < 0.02 unless the donor brings one -> exercises the rejection branch
        ip, fp, _ = _params(cfg, B, 5, H, W, seed)
        kc = host.kernel_config(cfg, True)
        got, _ = _run(x, ip, fp, None, kc, 0)
        want, _ = ao.augment(x, ip, fp, None, kc, 0)
        assert torch.equal(got, want)
        _, ok, _ = ao.transplant(x, torch.from_numpy(ip[:, 0].astype(np.int64)))
        assert ok.any()
    # 4-channel input: no transplantation, channels sliced (train.py:284)
    x4 = synthetic_batch(4, 32, 32, 9)[:, :4].contiguous()
    ip, fp, _ = _params(cfg, 4, 4, 32, 32, 0)
    got, _ = _run(x4, ip, fp, None, host.kernel_config(cfg, True), 0, c_out=3)
    assert torch.equal(got, x4[:, :3])


@pytest.mark.parametrize("which", ["random_affine", "random_erasing", "planckian_jitter", "color_jiggle", "blur",
                                   "random_plasma_shadow"])
def test_single_operator(which):
    cfg = host.AugmentationConfig(**{**ALL_OFF, which: True})
    B, H, W = 12, 48, 80
    x = synthetic_batch(B, H, W, 4)
    ip, fp, rough = _params(cfg, B, 5, H, W, 7)
    flag = {"random_affine": host.AFFINE, "planckian_jitter": host.PLANCKIAN, "blur": host.BLUR,
            "random_plasma_shadow": host.SHADOW}.get(which)
    if flag:
        assert (ip[:, 10] & flag).any() and not (ip[:, 10] & flag).all() or B < 4
    plasma = torch.from_numpy(ao.plasma(rough, 3, H, W)) if which == "random_plasma_shadow" else None
    kc = host.kernel_config(cfg, True)
    kp = torch.rand(B, 8, 2) * torch.tensor([W - 1.0, H - 1.0])
    got, gkp = _run(x, ip, fp, plasma, kc, 0, kp)
    want, wkp = ao.augment(x, ip, fp, plasma, kc, 0, kp)
    if which in ("random_erasing", "planckian_jitter", "random_plasma_shadow"):
        assert torch.equal(got, want)
    else:  # hue wrap / first-max ties are discontinuous: a handful of pixels may land on the other side
        _close(got, want, atol=5e-5, outliers=2e-4 if which == "color_jiggle" else 0.0)
    assert float((gkp - wkp).abs().max()) < 1e-5
    if which == "random_affine":
        assert float((got - x).abs().max()) > 0.1  # something moved


@pytest.mark.parametrize("order", [0x3210, 0x0123, 0x1302, 0x2031])
def test_colour_jiggle_orders(order):
    cfg = host.AugmentationConfig(**{**ALL_OFF, "color_jiggle": True})
    B, H, W = 4, 32, 32
    x = synthetic_batch(B, H, W, 6)[:, :3].contiguous()
    ip, fp, _ = _params(cfg, B, 3, H, W, order)
    ip[:, 9] = order
    kc = host.kernel_config(cfg, True)
    got, _ = _run(x, ip, fp, None, kc, 0)
    want, _ = ao.augment(x, ip, fp, None, kc, 0)
    _close(got, want, atol=5e-5, outliers=5e-4)


def test_depth_operators_share_the_philox_streams():
    """bias + Gaussian noise + near/far planes: the kernel's draws are the oracle's draws (bit-equal uniforms; the
    normal goes through logf/cosf -> 1e-5 in data units; plane flips only where the value sits on a threshold)."""
    for train, training in ((True, True), (True, False), (False, True)):
        cfg = host.AugmentationConfig(**{**ALL_OFF, "random_bias": True, "depth_gaussian_noise": True,
                                         "random_near_plane": True, "random_far_plane": True})
        B, H, W = 5, 40, 64
        x = synthetic_batch(B, H, W, 8)[:, :4].contiguous()
        x[:, 3] *= 1.6  # push part of the depth beyond the far plane
        ip, fp, _ = host.sample_params(cfg, False, B, 4, H, W, torch.Generator())
        kc = host.kernel_config(cfg, train, training)
        got, _ = _run(x, ip, fp, None, kc, 1234567890123)
        want, _ = ao.augment(x, ip, fp, None, kc, 1234567890123)
        assert torch.equal(got[:, :3], want[:, :3])
        _close(got[:, 3], want[:, 3], atol=2e-5, outliers=1e-4)
        zeros = float((want[:, 3] == 0).float().mean())
        assert 0.02 < zeros < 0.98  # both planes cut something
        if not train:  # val mode: no bias / noise; the planes still jitter
            kept = want[:, 3] != 0
            assert torch.equal(want[:, 3][kept], (0.035 * x[:, 3] / 0.035)[kept])


@pytest.mark.parametrize("shape", [(16, 64, 64), (5, 40, 72), (8, 256, 256)])
def test_full_pipeline_default_config(shape):
    B, H, W = shape
    cfg = host.AugmentationConfig()
    x = synthetic_batch(B, H, W, 21)
    ip, fp, rough = _params(cfg, B, 5, H, W, 5)
    plasma = torch.from_numpy(ao.plasma(rough, 77, H, W))
    kc = host.kernel_config(cfg, True)
    kp = torch.rand(B, 16) * (W - 1)
    got, gkp = _run(x, ip, fp, plasma, kc, 99, kp, c_out=4)
    want, wkp = ao.augment(x, ip, fp, plasma, kc, 99, kp, c_out=4)
    assert got.shape == (B, 4, H, W)
    _close(got[:, :3], want[:, :3], atol=1e-4, outliers=1e-3)
    _close(got[:, 3], want[:, 3], atol=1e-3, outliers=2e-3)
    assert float((gkp - wkp).abs().max()) < 1e-5


def test_module_drop_in_and_failure_modes():
    dev = torch.device("cuda", 0)
    B, H, W = 8, 256, 256
    x = synthetic_batch(B, H, W, 2).to(dev)
    kp = (torch.rand(B, 16) * 255).to(dev)
    aug = host.KeypointAugmentation(host.AugmentationConfig(), train=True, seed=1)
    out, coords = aug(x, kp)
    assert out.shape == (B, 5, H, W) and coords.shape == (B, 16) and torch.isfinite(out).all()
    assert float(out[:, :3].min()) >= 0 and float(out[:, :3].max()) <= 1
    out2, _ = aug(x, kp)
    assert not torch.equal(out, out2)  # fresh draws every call
    # val mode (validate.py:126-127): images untouched except the depth planes; coords normalised
    val = host.KeypointAugmentation(host.AugmentationConfig(), train=False)
    x4 = x[:, :4].contiguous()
    vo, vc = val(x4, kp.reshape(B, 8, 2))
    assert vc.shape == (B, 8, 2) and torch.allclose(vc, 2 * kp.reshape(B, 8, 2) / 255 - 1, atol=1e-6)
    assert torch.equal(vo[:, :3], x4[:, :3])
    kept = (vo[:, 3] != 0).cpu()
    xc = x4[:, 3].cpu()  # IEEE division on the host (torch's CUDA div-by-scalar multiplies by the reciprocal)
    assert torch.equal(vo[:, 3].cpu()[kept], (0.035 * xc / 0.035)[kept])
    with pytest.raises(RuntimeError):
        val(x4.cpu(), kp.cpu())


def test_no_write_outside_the_output_tensors():
    """Outputs carved out of a sentinel-filled arena (compute-sanitizer is not available on the pool): the kernels must not
    touch a byte outside [B][C_out][H][W] / [B][K][2] / [B][H][W] even on ragged tiles."""
    dev = torch.device("cuda", 0)
    B, H, W, K = 3, 37, 70, 8
    cfg = host.AugmentationConfig()
    x = synthetic_batch(B, H, W, 13).to(dev).contiguous().float()
    ip, fp, rough = _params(cfg, B, 5, H, W, 3)
    lib = host._lib.lib()
    guard = 4096
    n_out, n_kp, n_pl = B * 4 * H * W, B * K * 2, B * H * W
    arena = torch.full((guard + n_out + guard + n_kp + guard + n_pl + guard,), 7.5, dtype=torch.float32, device=dev)
    o0, k0, p0 = guard, guard + n_out + guard, guard + n_out + guard + n_kp + guard
    out, kpo, pl = arena[o0:o0 + n_out], arena[k0:k0 + n_kp], arena[p0:p0 + n_pl]
    S = lib.pb200_plasma_side(H, W)
    scratch = torch.empty(B, S, S, dtype=torch.float32, device=dev)
    rough_d = torch.from_numpy(rough).to(dev)
    import ctypes

    host._lib.check(lib.pb200_plasma(rough_d.data_ptr(), ctypes.c_uint64(5), scratch.data_ptr(), pl.data_ptr(), B, H, W,
                                     host._lib.current_stream_ptr()))
    counts = torch.empty(B, dtype=torch.int32, device=dev)
    kp = (torch.rand(B, K, 2) * 30).to(dev)
    st = host.config_struct(host.kernel_config(cfg, True))
    ip_d, fp_d = torch.from_numpy(ip).to(dev), torch.from_numpy(fp).to(dev)
    host._lib.check(lib.pb200_augment(x.data_ptr(), ip_d.data_ptr(), fp_d.data_ptr(),
                                      pl.data_ptr(), ctypes.byref(st), ctypes.c_uint64(9), counts.data_ptr(), kp.data_ptr(),
                                      out.data_ptr(), kpo.data_ptr(), B, 5, 4, H, W, K, host._lib.current_stream_ptr()))
    torch.cuda.synchronize()
    mask = torch.ones_like(arena, dtype=torch.bool)
    for a, n in ((o0, n_out), (k0, n_kp), (p0, n_pl)):
        mask[a:a + n] = False
    assert bool((arena[mask] == 7.5).all())
    assert bool((out != 7.5).any()) and torch.isfinite(out).all() and torch.isfinite(kpo).all() and torch.isfinite(pl).all()
