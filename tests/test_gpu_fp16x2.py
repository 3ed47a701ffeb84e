"""-m gpu: the "fp16x2" tcgen05 mode (PB200_PREC_FP16X2) through the C ABI vs the oracle.

fp16 activations x (fp16 hi + fp16 lo) weights, two MMAs per K step on ONE activation tile.  It is the tensor-core
mode that meets north_star's bf16-mode bound (<= 0.05 px vs the reference PyTorch detector on identical inputs and
random-init weights) on RGB at speed; on random-init RGBD (outputs reach +-3100 px) no 16-bit-operand mode does, and the
tests below say so with plain-pixel numbers instead of rescaling the bound.

Tolerances, in pixels and unscaled:
  * single layers vs F.conv2d in fp64 on the SAME fp16 activations and (hi + lo) weights: 2 fp16 ulps of the value
  * whole forward vs the reference fp32 forward (golden fixture made by the REAL reference class): <= 0.05 px RGB
    (measured on B200: 0.024 px golden inputs, 0.023 px on 8 sampled frames of a bs-256 u8 batch)
  * whole forward vs the operand-matched emulation `fp16x2_emulated_forward` (exact accumulation): <= 0.04 px RGB
    (measured 0.019-0.031).  This is NOT a tight kernel check -- the single-layer tests above are: the tensor core's
    fp32 accumulation differs from exact accumulation by ~2e-5 relative (the split-bf16 mode, whose emulation is exact
    to 0.0015 px, measures 0.006 px), which flips ~2 % of the fp16 roundings by one ulp, i.e. it re-draws part of the
    rounding noise that IS the mode's error.  Kernel and emulation are two samples of the same error distribution.
"""
import os

import numpy as np
import pytest
import torch

from oracle import detector_oracle as do
from oracle.make_golden import golden_input
from tests import gpu_util as gu

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "detector_golden.npz")


def _w_hi_lo(w: torch.Tensor) -> torch.Tensor:
    """the value the kernels multiply with: fp16(w) + fp16(w - fp16(w)), in float64"""
    hi = w.half().float()
    lo = (w - hi).half().float()
    return hi.double() + lo.double()


X2_CASES = [
    # (B, H, W, Cin, Cout, R, stride, pad, residual, relu)
    (2, 64, 64, 64, 64, 3, 1, 1, True, True),      # layer1 shape on the generic kernel (N = 64)
    (3, 64, 64, 64, 128, 3, 2, 1, False, True),    # layer2.0.conv1 (stride 2, parity maps)
    (2, 32, 32, 128, 128, 3, 1, 1, True, True),    # layer2
    (2, 16, 16, 256, 256, 3, 1, 1, True, True),    # layer3: N tile 256 (two 32 KB weight halves per stage)
    (3, 16, 16, 256, 512, 3, 2, 1, False, True),   # layer4.0.conv1: 8x8 images, 2 per tile, ragged tail
    (1, 8, 8, 512, 512, 3, 1, 1, True, True),      # layer4 at batch 1: half-empty tile
]


@pytest.mark.parametrize("block_n", ["0", "128"])
@pytest.mark.parametrize("cluster", ["1", "p", "2"])
@pytest.mark.parametrize("case", X2_CASES, ids=[str(c) for c in X2_CASES])
def test_x2_conv_layer_vs_exact_operand_reference(case, cluster, block_n, monkeypatch):
    from perseus_b200 import _lib

    monkeypatch.setenv("PB200_FORCE_CLUSTER", cluster)
    monkeypatch.setenv("PB200_FORCE_BLOCK_N", block_n)
    B, H, W, Cin, Cout, R, stride, pad, use_res, relu = case
    g = torch.Generator().manual_seed(sum(case) + 7)
    x = torch.randn(B, H, W, Cin, generator=g).half()
    w = torch.randn(Cout, Cin, R, R, generator=g) / (R * R * Cin) ** 0.5
    bias = torch.randn(Cout, generator=g)
    Ho, Wo = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - R) // stride + 1
    res = torch.randn(B, Ho, Wo, Cout, generator=g).half() if use_res else None
    want = torch.nn.functional.conv2d(x.double().permute(0, 3, 1, 2), _w_hi_lo(w), bias.double(), stride, pad)
    want = want.permute(0, 2, 3, 1)
    if use_res:
        want = want + res.double()
    if relu:
        want = want.relu()
    xd = x.cuda()
    outd = torch.full((B, Ho, Wo, Cout), float("nan"), dtype=torch.float16, device="cuda")
    resd = res.cuda() if use_res else None
    wc, bc = w.contiguous(), bias.contiguous()
    _lib.check(_lib.lib().pb200_debug_conv_f16x2(
        xd.data_ptr(), B, H, W, Cin, wc.data_ptr(), bc.data_ptr(), Cout, R, stride, pad,
        None if resd is None else resd.data_ptr(), int(relu), outd.data_ptr(), None, 0, None, _lib.current_stream_ptr()))
    torch.cuda.synchronize()
    got = outd.double().cpu()
    assert torch.isfinite(got).all(), "unwritten (NaN) outputs"
    err = (got - want).abs()
    tol = 2.0 ** -10 * want.abs() + 2e-3  # 2 fp16 ulps; near zero the weight lo term's own rounding (2^-22 w) dominates
    bad = err > tol
    assert not bad.any(), f"{int(bad.sum())} / {bad.numel()} outputs off; max err {float(err.max()):.5f}; first bad idx {bad.nonzero()[:4].tolist()}"


@pytest.mark.parametrize("cluster", ["1", "p"])
@pytest.mark.parametrize("case", [(3, 32, 32, 128, 64), (2, 16, 16, 256, 128), (3, 8, 8, 512, 256)], ids=str)
def test_x2_conv_with_fused_projection(case, cluster, monkeypatch):
    """conv3x3(y1) + conv1x1/2(x) accumulated in ONE TMEM tile, hi and lo weight halves of BOTH (K-extension)."""
    from perseus_b200 import _lib

    monkeypatch.setenv("PB200_FORCE_CLUSTER", cluster)
    B, H, W, C, Cds = case
    g = torch.Generator().manual_seed(sum(case))
    y1 = torch.randn(B, H, W, C, generator=g).half()
    x = torch.randn(B, 2 * H, 2 * W, Cds, generator=g).half()
    w = torch.randn(C, C, 3, 3, generator=g) / (9 * C) ** 0.5
    wds = torch.randn(C, Cds, generator=g) / Cds ** 0.5
    bias = torch.randn(C, generator=g)
    want = torch.nn.functional.conv2d(y1.double().permute(0, 3, 1, 2), _w_hi_lo(w), bias.double(), 1, 1)
    want = want + torch.nn.functional.conv2d(x.double().permute(0, 3, 1, 2), _w_hi_lo(wds)[:, :, None, None], None, 2, 0)
    want = want.permute(0, 2, 3, 1).relu()
    out = torch.full((B, H, W, C), float("nan"), dtype=torch.float16, device="cuda")
    y1d, xd = y1.cuda(), x.cuda()
    _lib.check(_lib.lib().pb200_debug_conv_f16x2(
        y1d.data_ptr(), B, H, W, C, w.contiguous().data_ptr(), bias.data_ptr(), C, 3, 1, 1, None, 1, out.data_ptr(),
        xd.data_ptr(), Cds, wds.contiguous().data_ptr(), _lib.current_stream_ptr()))
    torch.cuda.synchronize()
    got = out.double().cpu()
    assert torch.isfinite(got).all(), "unwritten (NaN) outputs"
    err = (got - want).abs()
    bad = err > 2.0 ** -10 * want.abs() + 2e-3
    assert not bad.any(), f"{int(bad.sum())} / {bad.numel()} outputs off; max err {float(err.max()):.5f}"


@pytest.mark.parametrize("reverse", [False, True], ids=["fwd", "rev"])
@pytest.mark.parametrize("single", [False, True], ids=["pair", "single"])
@pytest.mark.parametrize("B,use_res,relu,nhwc", [(1, False, True, False), (3, True, True, False), (2, True, True, True),
                                                 (5, False, False, False), (11, True, True, False)])
def test_x2_conv_p8_vs_exact_operand_reference(B, use_res, relu, nhwc, single, reverse, monkeypatch):
    """layer1 conv on the P8 layout, fp16, ONE N = 128 MMA per K step over [W_hi ; W_lo] (the epilogue adds the halves)."""
    from perseus_b200 import _lib

    # default: the CTA-pair kernel (tiles 2j / 2j+1 on two SMs, odd tile counts leave a dead peer tile); single: one CTA per tile
    if single:
        monkeypatch.setenv("PB200_P8_NOPAIR", "1")
    else:
        monkeypatch.delenv("PB200_P8_NOPAIR", raising=False)
    # rev: the tile list walked back to front (what every other layer1 launch of the plan does: launch_conv_p8)
    if reverse:
        monkeypatch.setenv("PB200_DEBUG_P8_REVERSE", "1")
    else:
        monkeypatch.delenv("PB200_DEBUG_P8_REVERSE", raising=False)

    H = W = 64
    g = torch.Generator().manual_seed(200 + B)
    x = torch.randn(B, H, W, 64, generator=g).half()
    w = torch.randn(64, 64, 3, 3, generator=g) / 24.0
    bias = torch.randn(64, generator=g)
    res = torch.randn(B, H, W, 64, generator=g).half() if use_res else None
    want = torch.nn.functional.conv2d(x.double().permute(0, 3, 1, 2), _w_hi_lo(w), bias.double(), 1, 1).permute(0, 2, 3, 1)
    if use_res:
        want = want + res.double()
    if relu:
        want = want.relu()
    slack = torch.zeros(4096, dtype=torch.uint8, device="cuda")  # noqa: F841
    xin = torch.cat([gu.nhwc_to_p8(x).cuda().flatten(), torch.zeros(2048, dtype=torch.float16, device="cuda")])
    resd = gu.nhwc_to_p8(res).cuda() if use_res else None
    if nhwc:
        out = torch.full((B, H, W, 64), float("nan"), dtype=torch.float16, device="cuda")
    else:
        out = torch.full((B, 8, H + 2, W + 2, 8), float("nan"), dtype=torch.float16, device="cuda")
    wc, bc = w.contiguous(), bias.contiguous()
    _lib.check(_lib.lib().pb200_debug_conv_p8_f16x2(xin.data_ptr(), B, H, W, wc.data_ptr(), bc.data_ptr(),
                                                    None if resd is None else resd.data_ptr(), int(relu), out.data_ptr(),
                                                    int(nhwc), _lib.current_stream_ptr()))
    torch.cuda.synchronize()
    if nhwc:
        got = out
    else:
        got, border = gu.p8_to_nhwc(out)
        assert border == 0.0, "P8 border must be written as zeros"
    got = got.double().cpu()
    assert torch.isfinite(got).all(), "unwritten (NaN) outputs"
    err = (got - want).abs()
    bad = err > 2.0 ** -10 * want.abs() + 2e-3
    assert not bad.any(), f"{int(bad.sum())}/{bad.numel()} off; max err {float(err.max()):.5f}; first {bad.nonzero()[:4].tolist()}"


@pytest.mark.parametrize("C,B", [(3, 1), (3, 5), (4, 3), (3, 40)])
def test_x2_stem_pool_vs_operand_matched_reference(C, B):
    """conv1+bn1+relu+maxpool, fp16 frames (k * 2^-12, d * 2^-4), hi + lo weight sets in shared memory."""
    from perseus_b200 import _lib

    model, sd = gu.native_model(C, "fp16x2", max_batch=64)
    x = golden_input(C, B=B, seed=21)
    want = do.fp16x2_emulated_forward(sd, x, upto="pool").permute(0, 2, 3, 1)  # (B, 64, 64, 64) NHWC
    xd = x.cuda()
    plan = model._plan(xd.device, B)
    model(xd)  # the frames buffer sits at the start of the workspace after a forward through the repack path
    frames = plan.workspace[: plan.frames_bytes(B)]
    out = torch.full((B, 8, 66, 66, 8), float("nan"), dtype=torch.float16, device="cuda")
    pad = torch.empty(4096, dtype=torch.uint8, device="cuda")  # noqa: F841
    _lib.check(_lib.lib().pb200_debug_stem_pool(plan.handle, frames.data_ptr(), out.data_ptr(), B,
                                                _lib.current_stream_ptr()))
    torch.cuda.synchronize()
    got, border = gu.p8_to_nhwc(out)
    got = got.float().cpu()
    assert torch.isfinite(got).all(), "unwritten (NaN) outputs"
    assert border == 0.0, "P8 border must be written as zeros"
    err = (got - want).abs()
    bad = err > 2.0 ** -10 * want.abs() + 2e-3
    assert not bad.any(), f"{int(bad.sum())}/{bad.numel()} off; max err {float(err.max()):.5f}; first {bad.nonzero()[:4].tolist()}"


def test_x2_ingest_frames_are_exact_scaled_values():
    """fp16 frames: RGB = k * 2^-12 exactly, depth = fp16(d * 2^-4); zero margins."""
    model, _ = gu.native_model(4, "fp16x2")
    B = 2
    bgra = gu.synth_u8_frames(B, 376, 672, 4, seed=1)
    depth = gu.synth_depth_m(B, 376, 672, seed=1)
    buf = model.ingest(bgra.cuda(), depth.cuda(), bgr=True, depth_divisor=0.035)
    fr = gu.frames_view(model, buf, B).float().cpu()
    assert fr[:, :3].abs().max() == 0 and fr[:, -3:].abs().max() == 0
    assert fr[:, :, :4].abs().max() == 0 and fr[:, :, -4:].abs().max() == 0
    inner = fr[:, 3:-3, 4:-4]
    for b in range(B):
        want = do.streaming_ingest(bgra[b].numpy(), depth[b].numpy(), 4).permute(1, 2, 0)
        assert torch.equal(inner[b][..., :3], torch.round(want[..., :3] * 255.0) / 4096.0)
        assert torch.equal(inner[b][..., 3], (want[..., 3] / 16.0).half().float())


def test_forward_fp16x2_rgb_meets_the_bf16_mode_tolerance():
    """north_star: tensor-core-mode outputs within 0.05 px of the reference PyTorch detector on identical inputs and
    random-init weights.  Golden fixture = the REAL reference class (oracle/make_golden.py), plain and randomised BN."""
    g = np.load(GOLDEN)
    for bn in (0, 1):
        model, sd = gu.native_model(3, "fp16x2", seed=0, randomize_bn=bool(bn))
        x = golden_input(3)
        y = model(x.cuda())
        assert y.shape == (2, 16) and y.dtype == torch.float32
        mx32, mean32 = gu.px_err(y, torch.from_numpy(g[f"c3_bn{bn}_out32"]))
        mx_emu, _ = gu.px_err(y, do.fp16x2_emulated_forward(sd, x))
        print(f"[fp16x2 RGB bn={bn}] vs reference fp32: max {mx32:.4f} mean {mean32:.4f} px | vs operand-matched emulation {mx_emu:.4f} px")
        assert mx_emu <= 0.04, f"kernel vs its operand-matched emulation: {mx_emu:.4f} px"
        assert mx32 <= 0.05, f"bn={bn}: {mx32:.4f} px vs the reference (bound 0.05 px)"


def test_forward_fp16x2_rgbd_bound_not_met_reports_plain_pixels():
    """RGBD on random-init weights: outputs reach +-24.5 normalised (3100 px) because depth enters at 3-14 data units, and
    0.05 px is 1.6e-5 of that range: NOT met by this mode (measured on B200: 0.40 px; emulation: 0.25-0.30 px; the
    split-bf16 mode: 0.10 px; only fp32 mode is below 0.05 px on RGBD).  Asserted in plain pixels, no range scaling:
    measured value + 50 % against the reference and against the emulation."""
    g = np.load(GOLDEN)
    for bn in (0, 1):
        model, sd = gu.native_model(4, "fp16x2", seed=0, randomize_bn=bool(bn))
        x = golden_input(4)
        y = model(x.cuda())
        ref = torch.from_numpy(g[f"c4_bn{bn}_out32"])
        emu = do.fp16x2_emulated_forward(sd, x)
        mx32, _ = gu.px_err(y, ref)
        mx_emu, _ = gu.px_err(y, emu)
        emu32, _ = gu.px_err(emu, ref)
        print(f"[fp16x2 RGBD bn={bn}] vs reference fp32 {mx32:.4f} px (NOT within 0.05) | vs emulation {mx_emu:.4f} px | "
              f"emulation vs reference {emu32:.4f} px | max |y| {float(ref.abs().max()):.1f}")
        assert mx_emu <= 0.6
        assert mx32 <= 0.6


@pytest.mark.parametrize("C", [3, 4])
def test_fp16x2_full_size_batch256_sampled_frames_vs_emulation(C):
    """BASELINE.json configs[1] / configs[2] at FULL size (bs 256, u8 frames [+ depth with NaN/Inf] through ingest ->
    backbone -> head -> decode): 8 sampled frames against the operand-matched emulation and the fp32 reference, plus
    the size-independent property that a frame's result does not depend on the batch it rides in."""
    model, sd = gu.native_model(C, "fp16x2", max_batch=256)
    B = 256
    rgb = gu.synth_u8_frames(B, seed=13)
    depth = gu.synth_depth_m(B, seed=13) if C == 4 else None
    px = model.forward_frames(rgb.cuda(), None if depth is None else depth.cuda(), depth_divisor=0.035)
    assert px.shape == (B, 8, 2) and torch.isfinite(px).all()
    idx = [0, 1, 63, 64, 127, 128, 254, 255]
    x = torch.stack([do.streaming_ingest(
        torch.cat([rgb[b].flip(-1), torch.zeros(256, 256, 1, dtype=torch.uint8)], -1).numpy(),
        None if depth is None else depth[b].numpy(), C) for b in idx])
    emu = do.denormalize_pixel_coordinates(do.fp16x2_emulated_forward(sd, x).reshape(len(idx), 8, 2), 256, 256)
    ref = do.denormalize_pixel_coordinates(do.reference_forward(sd, x).reshape(len(idx), 8, 2), 256, 256)
    got = px.cpu()[idx]
    e_emu, e_ref = float((got - emu).abs().max()), float((got - ref).abs().max())
    print(f"[fp16x2 C={C} bs256, 8 sampled frames] vs emulation {e_emu:.4f} px | vs reference fp32 {e_ref:.4f} px")
    if C == 3:
        assert e_ref <= 0.05, "north_star bound, RGB, at the full BASELINE batch size"
        assert e_emu <= 0.05
    else:  # outputs up to +-3100 px; the 0.05 px bound is not met on RGBD (see the test above): measured 0.28 / 0.41 px
        assert e_ref <= 0.6 and e_emu <= 0.8
    for i in (0, 127, 254):
        sub = model.forward_frames(rgb[i:i + 2].cuda(), None if depth is None else depth[i:i + 2].cuda(), depth_divisor=0.035)
        assert torch.equal(sub, px[i:i + 2])


def test_fp16x2_batch_edge_cases_and_default_precision():
    from perseus_b200.detector import KeypointCNN

    assert KeypointCNN().precision == "fp16x2", "the drop-in default must be the mode that meets the tolerance"
    model, sd = gu.native_model(3, "fp16x2", max_batch=8)
    x = golden_input(3, B=5, seed=9).cuda()
    y5 = model(x)
    y1 = torch.cat([model(x[i:i + 1]) for i in range(5)])
    assert torch.equal(y5, y1), "per-frame results must not depend on the batch they ride in"
    assert torch.equal(model(x), y5), "run-to-run determinism"
    xb = golden_input(3, B=11, seed=9).cuda()
    assert torch.equal(model(xb)[:5], y5)


def test_streaming_detector_fp16x2_private_plan_survives_replan_and_weight_reload():
    """The graph-captured streaming path owns its plan: a later larger-batch forward() on the same module (which frees
    and regrows the module's shared plan) must not disturb it, and load_state_dict must reach the graph's weights."""
    from perseus_b200.detector import KeypointCNN, StreamingDetector

    sd = do.seeded_state_dict(3, 8, seed=0)
    model = KeypointCNN(precision="fp16x2", max_batch=1)
    model.load_state_dict(sd)
    model.eval().cuda()
    det = StreamingDetector(model, 376, 672, 4, batch=1)
    bgra = gu.synth_u8_frames(1, 376, 672, 4, seed=3)
    first = det.infer(bgra.numpy()).copy()
    model(golden_input(3, B=9, seed=1).cuda())  # regrows the module's plan (max_batch 1 -> 9)
    torch.cuda.synchronize()
    assert np.array_equal(det.infer(bgra.numpy()), first)
    want = model.forward_frames(bgra.cuda(), bgr=True)
    assert torch.equal(torch.from_numpy(first), want.cpu())
    sd2 = do.seeded_state_dict(3, 8, seed=5)
    model.load_state_dict(sd2)
    second = det.infer(bgra.numpy())
    assert not np.array_equal(second, first)
    assert torch.equal(torch.from_numpy(second), model.forward_frames(bgra.cuda(), bgr=True).cpu())
