"""-m gpu: detector hot path through the C ABI vs the oracle.

Tolerances (stated per test):
  * ingest: bit-exact (byte/float-division work)
  * fp32 mode vs the reference fp32 forward: <= 1e-3 px RGB (north_star), <= 4e-3 px RGBD (the
    reference's own fp32-vs-fp64 error is 1.1e-3 px there)
  * bf16 mode vs the operand-matched bf16 emulation of the reference: <= 0.5 px max (only fp32
    accumulation order + rare 1-ulp bf16 rounding flips differ); vs the fp32 reference it is
    REPORTED, not asserted at 0.05 px (single-pass bf16 cannot meet that on random-init weights,
    BASELINE.md §4: PyTorch's own bf16 path is 0.79 px off).
"""
import ctypes

import numpy as np
import pytest
import torch

from oracle import detector_oracle as do
from oracle.make_golden import golden_input
from tests import gpu_util as gu

pytestmark = pytest.mark.gpu


# ---- ingest -------------------------------------------------------------------------------------
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_ingest_streaming_bgra_depth_bit_exact(precision):
    """ZED-shaped frames (376x672 BGRA + metres with NaN/Inf), centre crop: streaming.py:68-80."""
    model, _ = gu.native_model(4, precision)
    B = 2
    bgra = gu.synth_u8_frames(B, 376, 672, 4, seed=1)
    depth = gu.synth_depth_m(B, 376, 672, seed=1)
    buf = model.ingest(bgra.cuda(), depth.cuda(), bgr=True, depth_divisor=0.035)
    fr = gu.frames_view(model, buf, B).float().cpu()
    # zero margins
    assert fr[:, :3].abs().max() == 0 and fr[:, -3:].abs().max() == 0
    assert fr[:, :, :4].abs().max() == 0 and fr[:, :, -4:].abs().max() == 0
    inner = fr[:, 3:-3, 4:-4]  # (B, 256, 256, 4)
    for b in range(B):
        want = do.streaming_ingest(bgra[b].numpy(), depth[b].numpy(), 4).permute(1, 2, 0)  # HWC f32
        if precision == "fp32":
            assert torch.equal(inner[b], want), "fp32 ingest must equal the reference bit for bit"
        else:
            # bf16 frames carry RGB as exact integers (1/255 lives in the stem weights), depth as bf16
            assert torch.equal(inner[b][..., :3], torch.round(want[..., :3] * 255.0))
            assert torch.equal(inner[b][..., 3], want[..., 3].bfloat16().float())


def test_ingest_dataset_rgb_and_clip_bit_exact():
    """data.py:84-85 + validate.py:122-124 + val-mode DepthPlaneAugmentation with explicit jitter."""
    model, _ = gu.native_model(4, "fp32")
    B = 3
    rgb = gu.synth_u8_frames(B, 256, 256, 3, seed=2)
    g = torch.Generator().manual_seed(5)
    depth_units = torch.rand(B, 256, 256, generator=g) * 20.0  # 0..0.7 m scaled: both planes bite
    jitter = (torch.rand(B, 2, 256, 256, generator=g) - 0.5) * 0.1
    for jit in (None, jitter):
        buf = model.ingest(rgb.cuda(), depth_units.cuda(), clip=(0.1, 0.5), jitter=None if jit is None else jit.cuda())
        inner = gu.frames_view(model, buf, B).cpu()[:, 3:-3, 4:-4]
        for b in range(B):
            x = do.dataset_ingest(rgb[b].numpy(), depth_units[b].numpy())  # (4,H,W)
            d = do.depth_plane_clip(x[3], 0.1, 0.5, None if jit is None else jit[b, 0], None if jit is None else jit[b, 1])
            assert torch.equal(inner[b][..., :3], x[:3].permute(1, 2, 0))
            assert torch.equal(inner[b][..., 3], d)
    # RGB-only model: depth channel stays zero
    m3, _ = gu.native_model(3, "fp32")
    buf = m3.ingest(rgb.cuda())
    assert gu.frames_view(m3, buf, B)[..., 3].abs().max().item() == 0


def test_decode_matches_kornia_definition():
    from perseus_b200.detector import denormalize_pixel_coordinates

    g = torch.Generator().manual_seed(0)
    xy = torch.rand(1000, 8, 2, generator=g) * 2.4 - 1.2
    got = denormalize_pixel_coordinates(xy.cuda(), 256, 256).cpu()
    assert torch.equal(got, do.denormalize_pixel_coordinates(xy, 256, 256))
    got = denormalize_pixel_coordinates(xy.cuda(), 376, 672).cpu()
    assert torch.equal(got, do.denormalize_pixel_coordinates(xy, 376, 672))


# ---- one conv layer on the tensor cores --------------------------------------------------------
CONV_CASES = [
    # (B, H, W, Cin, Cout, R, stride, pad, residual, relu)
    (2, 64, 64, 64, 64, 3, 1, 1, False, True),     # layer1 shape
    (2, 64, 64, 64, 64, 3, 1, 1, True, True),      # + residual
    (3, 64, 64, 64, 128, 3, 2, 1, False, True),    # layer2.0.conv1 (stride 2, parity maps)
    (3, 64, 64, 64, 128, 1, 2, 0, False, False),   # layer2.0.downsample
    (2, 32, 32, 128, 128, 3, 1, 1, True, True),    # layer2
    (2, 16, 16, 256, 256, 3, 1, 1, True, True),    # layer3, N tile 256
    (3, 16, 16, 256, 512, 3, 2, 1, False, True),   # layer4.0.conv1: 8x8 images, 2 per tile, ragged tail
    (1, 8, 8, 512, 512, 3, 1, 1, True, True),      # layer4 at batch 1: half-empty tile
]


@pytest.mark.parametrize("cluster", ["1", "p", "2", "4"])
@pytest.mark.parametrize("case", CONV_CASES, ids=[str(c) for c in CONV_CASES])
def test_umma_conv_layer_vs_exact_operand_reference(case, cluster, monkeypatch):
    """tcgen05 implicit GEMM vs F.conv2d in fp64 on the SAME bf16-rounded operands: only fp32
    accumulation order and the final bf16 rounding differ (tolerance 2 bf16 ulps of the value)."""
    from perseus_b200 import _lib

    # cluster p: CTA pair (cta_group::2, M = 256); 2 / 4: the weight tile is split across the CTAs of a thread-block cluster and TMA-multicast
    # (ragged M-tile counts leave some CTAs of the last cluster without a valid tile)
    if True:
        monkeypatch.setenv("PB200_FORCE_CLUSTER", cluster)
    B, H, W, Cin, Cout, R, stride, pad, use_res, relu = case
    g = torch.Generator().manual_seed(hash(case) % 1000)
    x = torch.randn(B, H, W, Cin, generator=g).bfloat16()
    w = (torch.randn(Cout, Cin, R, R, generator=g) / (R * R * Cin) ** 0.5).bfloat16().float()
    bias = torch.randn(Cout, generator=g)
    Ho, Wo = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - R) // stride + 1
    res = torch.randn(B, Ho, Wo, Cout, generator=g).bfloat16() if use_res else None
    want = torch.nn.functional.conv2d(x.double().permute(0, 3, 1, 2), w.double(), bias.double(), stride, pad)
    want = want.permute(0, 2, 3, 1)
    if use_res:
        want = want + res.double()
    if relu:
        want = want.relu()
    xd, outd = x.cuda(), torch.full((B, Ho, Wo, Cout), float("nan"), dtype=torch.bfloat16, device="cuda")
    resd = res.cuda() if use_res else None
    wc, bc = w.contiguous(), bias.contiguous()
    _lib.check(
        _lib.lib().pb200_debug_conv_bf16(
            xd.data_ptr(), B, H, W, Cin, wc.data_ptr(), bc.data_ptr(), Cout, R, stride, pad,
            None if resd is None else resd.data_ptr(), int(relu), outd.data_ptr(), _lib.current_stream_ptr(),
        )
    )
    torch.cuda.synchronize()
    got = outd.double().cpu()
    assert torch.isfinite(got).all(), "unwritten (NaN) outputs"
    err = (got - want).abs()
    tol = 2.0 ** -7 * want.abs() + 1e-2
    bad = err > tol
    assert not bad.any(), f"{int(bad.sum())} / {bad.numel()} outputs off; max err {float(err.max()):.4f}; first bad idx {bad.nonzero()[:4].tolist()}"


DS_CASES = [
    # (B, H, W, Cin = Cout, ds_Cin): conv2 of layer2.0 / layer3.0 / layer4.0 with the block's 1x1/2 projection fused in
    (3, 32, 32, 128, 64),
    (2, 16, 16, 256, 128),
    (3, 8, 8, 512, 256),
]


@pytest.mark.parametrize("cluster", ["1", "p"])
@pytest.mark.parametrize("case", DS_CASES, ids=[str(c) for c in DS_CASES])
def test_umma_conv_with_fused_projection_vs_exact_operand_reference(case, cluster, monkeypatch):
    """conv3x3(y1) + conv1x1/2(x) accumulated in ONE TMEM tile (extra K columns through the parity-(0,0) map) vs the two
    F.conv2d in fp64 on the same bf16 operands: the projection is never rounded, so the only error is the final rounding."""
    from perseus_b200 import _lib

    monkeypatch.setenv("PB200_FORCE_CLUSTER", cluster)
    B, H, W, C, Cds = case
    g = torch.Generator().manual_seed(sum(case))
    y1 = torch.randn(B, H, W, C, generator=g).bfloat16()
    x = torch.randn(B, 2 * H, 2 * W, Cds, generator=g).bfloat16()
    w = (torch.randn(C, C, 3, 3, generator=g) / (9 * C) ** 0.5).bfloat16().float()
    wds = (torch.randn(C, Cds, generator=g) / Cds ** 0.5).bfloat16().float()
    bias = torch.randn(C, generator=g)
    want = torch.nn.functional.conv2d(y1.double().permute(0, 3, 1, 2), w.double(), bias.double(), 1, 1)
    want = want + torch.nn.functional.conv2d(x.double().permute(0, 3, 1, 2), wds.double()[:, :, None, None], None, 2, 0)
    want = want.permute(0, 2, 3, 1).relu()
    out = torch.full((B, H, W, C), float("nan"), dtype=torch.bfloat16, device="cuda")
    y1d, xd = y1.cuda(), x.cuda()
    _lib.check(_lib.lib().pb200_debug_conv_ds_bf16(y1d.data_ptr(), B, H, W, C, w.contiguous().data_ptr(), bias.data_ptr(), C,
                                                   xd.data_ptr(), Cds, wds.contiguous().data_ptr(), 1, out.data_ptr(),
                                                   _lib.current_stream_ptr()))
    torch.cuda.synchronize()
    got = out.double().cpu()
    assert torch.isfinite(got).all(), "unwritten (NaN) outputs"
    err = (got - want).abs()
    bad = err > 2.0 ** -7 * want.abs() + 1e-2
    assert not bad.any(), f"{int(bad.sum())} / {bad.numel()} outputs off; max err {float(err.max()):.4f}"


def _split(t: torch.Tensor):
    hi = t.float().bfloat16()
    lo = (t.float() - hi.float()).bfloat16()
    return hi, lo


SPLIT_CASES = [
    # B, H, W, Cin, Cout, R, stride, pad, residual, relu
    (2, 64, 64, 64, 64, 3, 1, 1, True, True),
    (3, 64, 64, 64, 128, 3, 2, 1, False, True),
    (3, 64, 64, 64, 128, 1, 2, 0, False, False),
    (5, 16, 16, 256, 256, 3, 1, 1, True, True),
    (9, 8, 8, 512, 512, 3, 1, 1, True, True),
]


@pytest.mark.parametrize("case", SPLIT_CASES, ids=[str(c) for c in SPLIT_CASES])
def test_umma_conv_split_layer_vs_fp64(case):
    """split-bf16 ("bf16x3") conv: operands a = hi + lo, w = hi + lo on tcgen05 (3 MMAs per k-block) vs F.conv2d in
    fp64 on the fp32 operands; the only losses are the dropped lo*lo term and the 16-bit split of the operands
    (~2^-16 relative each) plus fp32 accumulation."""
    from perseus_b200 import _lib

    B, H, W, Cin, Cout, R, stride, pad, use_res, relu = case
    g = torch.Generator().manual_seed(sum(case))
    x = torch.randn(B, H, W, Cin, generator=g)
    xh, xl = _split(x)
    xv = xh.float() + xl.float()  # the value the split tensor carries
    w = torch.randn(Cout, Cin, R, R, generator=g) / (R * R * Cin) ** 0.5
    bias = torch.randn(Cout, generator=g)
    Ho, Wo = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - R) // stride + 1
    want = torch.nn.functional.conv2d(xv.double().permute(0, 3, 1, 2), w.double(), bias.double(), stride, pad)
    want = want.permute(0, 2, 3, 1)
    resd = None
    if use_res:
        res = torch.randn(B, Ho, Wo, Cout, generator=g)
        rh, rl = _split(res)
        want = want + rh.double() + rl.double()
        resd = torch.cat([rh, rl], -1).contiguous().cuda()
    if relu:
        want = want.relu()
    xd = torch.cat([xh, xl], -1).contiguous().cuda()
    outd = torch.full((B, Ho, Wo, 2 * Cout), float("nan"), dtype=torch.bfloat16, device="cuda")
    wc, bc = w.contiguous(), bias.contiguous()
    _lib.check(
        _lib.lib().pb200_debug_conv_split(
            xd.data_ptr(), B, H, W, Cin, wc.data_ptr(), bc.data_ptr(), Cout, R, stride, pad,
            None if resd is None else resd.data_ptr(), int(relu), outd.data_ptr(), _lib.current_stream_ptr(),
        )
    )
    torch.cuda.synchronize()
    o = outd.double().cpu()
    assert torch.isfinite(o).all(), "unwritten (NaN) outputs"
    got = o[..., :Cout] + o[..., Cout:]
    err = (got - want).abs()
    tol = 2.0 ** -14 * want.abs() + 2e-4
    bad = err > tol
    assert not bad.any(), f"{int(bad.sum())} / {bad.numel()} outputs off; max err {float(err.max()):.6f}; first bad idx {bad.nonzero()[:4].tolist()}"


def _bf16x3_errors(C, bn):
    g = np.load(gu.os.path.join(gu.os.path.dirname(__file__), "golden", "detector_golden.npz"))
    model, sd = gu.native_model(C, "bf16x3", seed=0, randomize_bn=bool(bn))
    x = golden_input(C)
    y = model(x.cuda())
    assert y.shape == (2, 16) and y.dtype == torch.float32
    mx32, _ = gu.px_err(y, torch.from_numpy(g[f"c{C}_bn{bn}_out32"]))
    mx_emu, _ = gu.px_err(y, do.bf16x3_emulated_forward(sd, x))
    print(f"[bf16x3 C={C} bn={bn}] vs reference fp32 {mx32:.5f} px | vs split emulation {mx_emu:.5f} px | "
          f"max |y| {float(np.abs(g[f'c{C}_bn{bn}_out32']).max()):.1f}")
    return mx32, mx_emu


def _frames_path_error(C, precision):
    """u8 frames path (ingest -> forward -> decode) on a ragged batch vs the fp32 reference, in plain pixels"""
    model, sd = gu.native_model(C, precision, max_batch=8)
    B = 5
    rgb = gu.synth_u8_frames(B, 256, 256, 3, seed=11)
    depth = gu.synth_depth_m(B, 256, 256, seed=11) if C == 4 else None
    px = model.forward_frames(rgb.cuda(), None if depth is None else depth.cuda(), depth_divisor=0.035)
    x = torch.stack([do.streaming_ingest(
        torch.cat([rgb[b].flip(-1), torch.zeros(256, 256, 1, dtype=torch.uint8)], -1).numpy(),
        None if depth is None else depth[b].numpy(), C) for b in range(B)])
    want = do.denormalize_pixel_coordinates(do.reference_forward(sd, x).reshape(B, 8, 2), 256, 256)
    return float((px.cpu() - want).abs().max())


def test_forward_bf16x3_rgb_meets_the_bf16_mode_tolerance():
    """north_star: tensor-core-mode outputs within 0.05 px of the reference.  Single-pass bf16 cannot (0.6-0.9 px on
    random-init weights, like PyTorch's own bf16); the split-bf16 tcgen05 mode does on RGB (measured ~0.006 px)."""
    for bn in (0, 1):
        mx32, _ = _bf16x3_errors(3, bn)
        assert mx32 <= 0.05, f"bn={bn}: {mx32:.4f} px vs the reference"
    err = _frames_path_error(3, "bf16x3")
    print(f"[bf16x3 C=3] forward_frames: {err:.5f} px vs the reference")
    assert err <= 0.05


def test_forward_bf16x3_rgbd_bound_not_met_reports_plain_pixels():
    """RGBD, random-init weights: outputs reach +-24.5 normalised = 3100 px (depth enters at 3-14 data units) and the
    0.05 px bound is NOT met: 0.11 px measured on B200 (r01; the emulation with exact accumulation gives 0.014 px, the
    rest is fp32 accumulation in the tensor core at that output range).  Asserted: the measured value + 20 %, in plain
    pixels, no range scaling."""
    for bn in (0, 1):
        mx32, _ = _bf16x3_errors(4, bn)
        assert mx32 <= 0.132, f"bn={bn}: {mx32:.4f} px vs the reference"
    err = _frames_path_error(4, "bf16x3")
    print(f"[bf16x3 C=4] forward_frames: {err:.5f} px vs the reference (bound 0.05 px NOT met on RGBD)")
    assert err <= 0.2


# ---- P8-layout kernels: fused stem + maxpool, resident-weight 3x3 conv ---------------------------------
@pytest.mark.parametrize("C,B", [(3, 1), (3, 5), (4, 3), (3, 40)])
def test_stem_pool_fused_vs_operand_matched_reference(C, B):
    """conv1+bn1+relu+maxpool in one tcgen05 kernel vs the bf16-operand emulation of the reference stem
    (exact accumulation): only fp32 accumulation order + the final bf16 rounding differ.  B = 1 / 5 / 40
    exercise band sizes 1 / 2 / 8 pooled rows (the overlap row between bands) and multi-unit CTAs."""
    from perseus_b200 import _lib

    model, sd = gu.native_model(C, "bf16", max_batch=64)
    x = golden_input(C, B=B, seed=21)
    want = do.bf16_emulated_forward(sd, x, upto="pool").permute(0, 2, 3, 1)  # (B, 64, 64, 64) NHWC
    xd = x.cuda()
    plan = model._plan(xd.device, B)
    # frames via the public repack path: run a forward, the frames buffer sits at the start of the workspace
    model(xd)
    frames = plan.workspace[: plan.frames_bytes(B)]
    out = torch.full((B, 8, 66, 66, 8), float("nan"), dtype=torch.bfloat16, device="cuda")
    pad = torch.empty(4096, dtype=torch.uint8, device="cuda")  # noqa: F841  (keeps slack after `out` mapped)
    _lib.check(_lib.lib().pb200_debug_stem_pool(plan.handle, frames.data_ptr(), out.data_ptr(), B,
                                                _lib.current_stream_ptr()))
    torch.cuda.synchronize()
    got, border = gu.p8_to_nhwc(out)
    got = got.float().cpu()
    assert torch.isfinite(got).all(), "unwritten (NaN) outputs"
    assert border == 0.0, "P8 border must be written as zeros"
    err = (got - want).abs()
    tol = 2.0 ** -7 * want.abs() + 1e-2
    bad = err > tol
    assert not bad.any(), f"{int(bad.sum())}/{bad.numel()} off; max err {float(err.max()):.4f}; first {bad.nonzero()[:4].tolist()}"


@pytest.mark.parametrize("reverse", [False, True], ids=["fwd", "rev"])
@pytest.mark.parametrize("single", [False, True], ids=["pair", "single"])
@pytest.mark.parametrize("B,use_res,relu,nhwc", [(1, False, True, False), (3, True, True, False), (2, True, True, True),
                                                 (5, False, False, False), (11, True, True, False)])
def test_conv_p8_vs_exact_operand_reference(B, use_res, relu, nhwc, single, reverse, monkeypatch):
    """layer1 conv (3x3/1, 64->64) on the P8 layout vs F.conv2d in fp64 on the same bf16 operands."""
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
    g = torch.Generator().manual_seed(100 + B)
    x = torch.randn(B, H, W, 64, generator=g).bfloat16()
    w = (torch.randn(64, 64, 3, 3, generator=g) / 24.0).bfloat16().float()
    bias = torch.randn(64, generator=g)
    res = torch.randn(B, H, W, 64, generator=g).bfloat16() if use_res else None
    want = torch.nn.functional.conv2d(x.double().permute(0, 3, 1, 2), w.double(), bias.double(), 1, 1).permute(0, 2, 3, 1)
    if use_res:
        want = want + res.double()
    if relu:
        want = want.relu()
    slack = torch.zeros(4096, dtype=torch.uint8, device="cuda")  # noqa: F841
    xin = torch.cat([gu.nhwc_to_p8(x).cuda().flatten(), torch.zeros(2048, dtype=torch.bfloat16, device="cuda")])
    resd = gu.nhwc_to_p8(res).cuda() if use_res else None
    if nhwc:
        out = torch.full((B, H, W, 64), float("nan"), dtype=torch.bfloat16, device="cuda")
    else:
        out = torch.full((B, 8, H + 2, W + 2, 8), float("nan"), dtype=torch.bfloat16, device="cuda")
    wc, bc = w.contiguous(), bias.contiguous()
    _lib.check(_lib.lib().pb200_debug_conv_p8(xin.data_ptr(), B, H, W, wc.data_ptr(), bc.data_ptr(),
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
    tol = 2.0 ** -7 * want.abs() + 1e-2
    bad = err > tol
    assert not bad.any(), f"{int(bad.sum())}/{bad.numel()} off; max err {float(err.max()):.4f}; first {bad.nonzero()[:4].tolist()}"

# ---- whole forward --------------------------------------------------------------------------------
def _fp32_errors(C, bn):
    g = np.load(gu.os.path.join(gu.os.path.dirname(__file__), "golden", "detector_golden.npz"))
    model, sd = gu.native_model(C, "fp32", seed=0, randomize_bn=bool(bn))
    x = golden_input(C)
    y = model(x.cuda())
    assert y.shape == (2, 16) and y.dtype == torch.float32
    mx32, _ = gu.px_err(y, torch.from_numpy(g[f"c{C}_bn{bn}_out32"]))
    mx64, _ = gu.px_err(y, torch.from_numpy(g[f"c{C}_bn{bn}_out64"]))
    ref_own, _ = gu.px_err(torch.from_numpy(g[f"c{C}_bn{bn}_out32"]), torch.from_numpy(g[f"c{C}_bn{bn}_out64"]))
    print(f"[fp32 C={C} bn={bn}] vs reference fp32 {mx32:.2e} px | vs fp64 truth {mx64:.2e} px | reference fp32 vs its "
          f"own fp64 run {ref_own:.2e} px")
    return mx32, mx64, ref_own


def test_forward_fp32_mode_rgb_matches_reference_to_1e3_px():
    """BASELINE.json configs[0] through forward(x): golden fixture generated by the real reference; north_star's
    1e-3 px fp32-mode bound, RGB."""
    for bn in (0, 1):
        mx32, mx64, _ = _fp32_errors(3, bn)
        assert mx32 <= 1e-3 and mx64 <= 1e-3, f"bn={bn}: {mx32:.2e} px vs reference fp32 (vs fp64 truth {mx64:.2e})"


def test_forward_fp32_mode_rgbd_1e3_px_not_met_is_the_references_own_fp32_noise():
    """RGBD, random-init weights (outputs up to 3100 px): the 1e-3 px bound is NOT met against the reference's fp32
    run, because that run is itself 1.1e-3 px away from its own fp64 evaluation -- two fp32
    summation orders at this output range.  Asserted in plain pixels: within 4e-3 px of the reference and no further
    from fp64 truth than the reference's own fp32 run is + 1e-3 px."""
    for bn in (0, 1):
        mx32, mx64, ref_own = _fp32_errors(4, bn)
        assert mx32 <= 4e-3, f"bn={bn}: {mx32:.2e} px vs reference fp32"
        assert mx64 <= ref_own + 1e-3, f"bn={bn}: {mx64:.2e} px vs fp64 truth (reference's own: {ref_own:.2e})"


@pytest.mark.parametrize("mode", ["0", "1", "2"])
@pytest.mark.parametrize("C", [3, 4])
def test_stem_f32_kernel_bit_identical_to_generic_simt_conv(C, mode, monkeypatch):
    """The dedicated fp32 stem (conv_simt.cu::stem_f32_kernel, scalar / FFMA2 / mixed builds) keeps the generic
    kernel's K order, so the whole fp32 forward must not change by a single bit; odd batch = ragged last wave."""
    model, _ = gu.native_model(C, "fp32", seed=0, randomize_bn=True)
    g = torch.Generator().manual_seed(11)
    x = torch.rand(5, C, 256, 256, generator=g).cuda()
    monkeypatch.setenv("PB200_STEM_F32_MODE", "-1")
    want = model(x).clone()
    monkeypatch.setenv("PB200_STEM_F32_MODE", mode)
    got = model(x).clone()
    assert torch.isfinite(want).all()
    assert torch.equal(got, want)


@pytest.mark.parametrize("C", [3, 4])
def test_forward_bf16_mode_matches_operand_matched_reference(C):
    model, sd = gu.native_model(C, "bf16")
    x = golden_input(C, B=3, seed=3)
    y = model(x.cuda())
    emu = do.bf16_emulated_forward(sd, x)
    ref = do.reference_forward(sd, x, torch.float32)
    mx, mean = gu.px_err(y, emu)
    mx_ref, mean_ref = gu.px_err(y, ref)
    mx_emu_ref, _ = gu.px_err(emu, ref)
    print(f"[bf16 C={C}] vs emulation max {mx:.4f} mean {mean:.4f} px | vs fp32 ref max {mx_ref:.3f} mean "
          f"{mean_ref:.3f} px | emulation vs ref {mx_emu_ref:.3f} px")
    # plain pixels, no range scaling.  RGB: measured 0.21 px on B200 (r01).  RGBD outputs reach +-3100 px on random-init
    # weights, one bf16 ulp there is 8-16 px: the kernel-vs-emulation distance is a handful of 1-ulp activation flips.
    bound = 0.3 if C == 3 else 3.0
    assert mx <= bound, f"bf16 path vs operand-matched emulation: {mx:.3f} px (bound {bound} px)"
    # vs the fp32 reference single-pass bf16 is NOT within north_star's 0.05 px (0.6 px RGB): reported above; the
    # kernel must not be further from the reference than the emulation's own distance + the bound above
    assert mx_ref <= mx_emu_ref + bound


def test_forward_frames_equals_forward_on_ingested_tensor():
    """ingest -> forward (u8 frames) == forward(x) on the reference-style float tensor (fp32 mode)."""
    model, sd = gu.native_model(4, "fp32")
    B = 2
    rgb = gu.synth_u8_frames(B, 256, 256, 3, seed=4)
    depth = gu.synth_depth_m(B, 256, 256, seed=4)
    px = model.forward_frames(rgb.cuda(), depth.cuda(), depth_divisor=0.035)
    x = torch.stack([do.streaming_ingest(
        torch.cat([rgb[b].flip(-1), torch.zeros(256, 256, 1, dtype=torch.uint8)], -1).numpy(), depth[b].numpy(), 4)
        for b in range(B)])
    norm, px2 = model.forward_pixels(x.cuda())
    assert torch.equal(px, px2)
    want = do.denormalize_pixel_coordinates(do.reference_forward(sd, x).reshape(B, 8, 2), 256, 256)
    assert (px.cpu() - want).abs().max() <= 4e-3
    assert px.shape == (B, 8, 2)


def test_batch_edge_cases_and_determinism():
    model, sd = gu.native_model(3, "bf16", max_batch=8)
    x = golden_input(3, B=5, seed=9).cuda()
    y5 = model(x)
    y1 = torch.cat([model(x[i:i + 1]) for i in range(5)])
    assert torch.equal(y5, y1), "per-frame results must not depend on the batch they ride in"
    assert torch.equal(model(x), y5), "run-to-run determinism"
    # growing past max_batch re-plans transparently
    xb = golden_input(3, B=11, seed=9).cuda()
    assert torch.equal(model(xb)[:5], y5)
    # 3-D input (the streaming caller adds the batch dim itself; accept both)
    assert torch.equal(model(x[0]), y5[0])
    with pytest.raises(ValueError):
        model(torch.rand(1, 4, 256, 256).cuda())


def test_load_state_dict_refreshes_device_weights():
    from perseus_b200.detector import KeypointCNN

    m = KeypointCNN(precision="fp32", max_batch=2).eval().cuda()
    x = golden_input(3, B=1).cuda()
    y0 = m(x)
    sd = do.seeded_state_dict(3, 8, seed=5)
    m.load_state_dict(sd)
    y1 = m(x)
    assert not torch.equal(y0, y1)
    assert gu.px_err(y1, do.reference_forward(sd, x.cpu()))[0] <= 1e-3


@pytest.mark.parametrize("C", [3, 4])
def test_full_size_batch256_bf16_sampled_frames_vs_emulation(C):
    """BASELINE.json configs[1] (RGB) and configs[2] (RGBD) at FULL size: batch 256 bf16 through ingest -> backbone ->
    head -> decode.  8 sampled frames against the operand-matched bf16 emulation of the reference (plain pixels), plus
    the size-independent property that every frame equals the same frame evaluated in a batch of 2; all outputs finite."""
    model, sd = gu.native_model(C, "bf16", max_batch=256)
    rgb = gu.synth_u8_frames(256, seed=11)
    depth = gu.synth_depth_m(256, seed=11) if C == 4 else None
    dd = None if depth is None else depth.cuda()
    px = model.forward_frames(rgb.cuda(), dd, depth_divisor=0.035)
    assert px.shape == (256, 8, 2) and torch.isfinite(px).all()
    for i in (0, 127, 254):
        assert torch.equal(model.forward_frames(rgb[i:i + 2].cuda(), None if dd is None else dd[i:i + 2], depth_divisor=0.035),
                           px[i:i + 2])
    idx = [0, 1, 63, 64, 127, 128, 254, 255]
    x = torch.stack([do.streaming_ingest(
        torch.cat([rgb[b].flip(-1), torch.zeros(256, 256, 1, dtype=torch.uint8)], -1).numpy(),
        None if depth is None else depth[b].numpy(), C) for b in idx])
    emu = do.denormalize_pixel_coordinates(do.bf16_emulated_forward(sd, x).reshape(len(idx), 8, 2), 256, 256)
    ref = do.denormalize_pixel_coordinates(do.reference_forward(sd, x).reshape(len(idx), 8, 2), 256, 256)
    got = px.cpu()[idx]
    e_emu, e_ref = float((got - emu).abs().max()), float((got - ref).abs().max())
    print(f"[bf16 C={C} bs256, 8 sampled frames] vs emulation {e_emu:.4f} px | vs reference fp32 {e_ref:.4f} px "
          "(single-pass bf16 does not meet 0.05 px)")
    # RGBD on random-init weights: outputs span +-3100 px, so single-pass bf16 rounding noise is px-sized; kernel and
    # emulation accumulate in different orders and are two draws of that noise (a different K walk in layer1 alone moved
    # this figure from 2.4 to 3.17 px in r02): bound = measured + margin
    assert e_emu <= (0.3 if C == 3 else 3.0)


def test_streaming_detector_graph_matches_eager_path():
    """BASELINE.json configs[4] shape: ZED-shaped BGRA + depth frames, bs 1 and 2, one CUDA-graph replay per step
    (H2D + ingest + forward + decode + D2H) == the eager forward_frames result, bit for bit, for changing inputs."""
    from perseus_b200.detector import StreamingDetector

    for C, B in ((4, 1), (3, 2)):
        model, sd = gu.native_model(C, "bf16")
        det = StreamingDetector(model, 376, 672, 4, batch=B)
        for seed in (1, 2, 3):
            bgra = gu.synth_u8_frames(B, 376, 672, 4, seed=seed)
            depth = gu.synth_depth_m(B, 376, 672, seed=seed)
            got = det.infer(bgra.numpy(), depth.numpy() if C == 4 else None)
            want = model.forward_frames(bgra.cuda(), depth.cuda() if C == 4 else None, bgr=True, depth_divisor=0.035)
            assert got.shape == (B, 8, 2)
            assert torch.equal(torch.from_numpy(got), want.cpu())
    # fp32 parity mode against the reference on the streaming ingest
    model, sd = gu.native_model(3, "fp32")
    det = StreamingDetector(model, 376, 672, 4, batch=1)
    bgra = gu.synth_u8_frames(1, 376, 672, 4, seed=7)
    got = torch.from_numpy(det.infer(bgra.numpy()))
    x = do.streaming_ingest(bgra[0].numpy(), None, 3)[None]
    want = do.denormalize_pixel_coordinates(do.reference_forward(sd, x).reshape(1, 8, 2), 256, 256)
    assert (got - want).abs().max() <= 1e-3


@pytest.mark.parametrize("case", [CONV_CASES[1], CONV_CASES[2], CONV_CASES[4]], ids=str)
def test_pair_conv_resident_weights_variant(case, monkeypatch):
    """conv_umma2_kernel<.., RESB>: the CTA pair keeps its halves of the whole weight matrix in shared memory and streams
    activations only (single-N-tile convs; opt-in with PB200_RESB=1 after it measured slower on layer2)."""
    monkeypatch.setenv("PB200_RESB", "1")
    test_umma_conv_layer_vs_exact_operand_reference(case, "p", monkeypatch)
    if case is CONV_CASES[4]:
        test_umma_conv_with_fused_projection_vs_exact_operand_reference(DS_CASES[0], "p", monkeypatch)


HALO_CASES = [
    # (B, H, W, Cin, residual, relu, ds_Cin)
    (2, 32, 32, 128, True, True, 0),     # layer2.1.conv2
    (4, 32, 32, 128, False, True, 0),    # layer2.1.conv1
    (3, 32, 32, 128, False, True, 64),   # layer2.0.conv2 + fused projection, odd batch (the pair's second image is missing)
    (2, 64, 64, 64, True, False, 0),     # a different width: 33 tiles per image, patch rows of 66 pixels
]


@pytest.mark.parametrize("case", HALO_CASES, ids=[str(c) for c in HALO_CASES])
def test_conv_halo_vs_exact_operand_reference(case):
    """csrc/conv_halo.cu (halo patch + resident KT weights, CTA pairs over image pairs) vs F.conv2d in fp64 on the same bf16
    operands; every output pixel must be written (NaN-initialised output) and nothing outside the tensor."""
    from perseus_b200 import _lib

    B, H, W, Cin, use_res, relu, Cds = case
    Cout = 128
    g = torch.Generator().manual_seed(sum(case) + 3)
    x = torch.randn(B, H, W, Cin, generator=g).bfloat16()
    w = (torch.randn(Cout, Cin, 3, 3, generator=g) / (9 * Cin) ** 0.5).bfloat16().float()
    bias = torch.randn(Cout, generator=g)
    res = torch.randn(B, H, W, Cout, generator=g).bfloat16() if use_res else None
    want = torch.nn.functional.conv2d(x.double().permute(0, 3, 1, 2), w.double(), bias.double(), 1, 1)
    xds = wds = None
    if Cds:
        xds = torch.randn(B, 2 * H, 2 * W, Cds, generator=g).bfloat16()
        wds = (torch.randn(Cout, Cds, generator=g) / Cds ** 0.5).bfloat16().float()
        want = want + torch.nn.functional.conv2d(xds.double().permute(0, 3, 1, 2), wds.double()[:, :, None, None], None, 2, 0)
    want = want.permute(0, 2, 3, 1)
    if use_res:
        want = want + res.double()
    if relu:
        want = want.relu()
    guard = 4096
    buf = torch.full((B * H * W * Cout + 2 * guard,), float("nan"), dtype=torch.bfloat16, device="cuda")
    outd = buf[guard:-guard].view(B, H, W, Cout)
    xd = x.cuda()
    resd = res.cuda() if use_res else None
    xdsd = xds.cuda() if Cds else None
    wc, bc = w.contiguous(), bias.contiguous()
    wdsc = wds.contiguous() if Cds else None
    _lib.check(_lib.lib().pb200_debug_conv_halo(
        xd.data_ptr(), B, H, W, Cin, wc.data_ptr(), bc.data_ptr(), Cout, None if resd is None else resd.data_ptr(), int(relu),
        outd.data_ptr(), None if xdsd is None else xdsd.data_ptr(), Cds, None if wdsc is None else wdsc.data_ptr(),
        _lib.current_stream_ptr()))
    torch.cuda.synchronize()
    assert torch.isnan(buf[:guard]).all() and torch.isnan(buf[-guard:]).all(), "wrote outside the output tensor"
    # widths whose 128-position tile would need more than the 7-row halo box are refused, not mis-computed
    assert _lib.lib().pb200_debug_conv_halo(xd.data_ptr(), B, 16, 16, Cin, wc.data_ptr(), bc.data_ptr(), Cout, None, 1,
                                            outd.data_ptr(), None, 0, None, _lib.current_stream_ptr()) == 1
    got = outd.double().cpu()
    assert torch.isfinite(got).all(), "unwritten (NaN) outputs"
    err = (got - want).abs()
    bad = err > 2.0 ** -7 * want.abs() + 1e-2
    assert not bad.any(), f"{int(bad.sum())} / {bad.numel()} outputs off; max err {float(err.max()):.4f}; first bad idx {bad.nonzero()[:4].tolist()}"


@pytest.mark.parametrize("precision", ["bf16", "fp16x2"])
@pytest.mark.parametrize("B", [256, 37, 131, 255])
def test_chained_layers34_bit_equal_to_one_launch_per_conv(precision, B, monkeypatch):
    """Layers 3-4 run as ONE persistent launch whose tiles wait on in-kernel dependency counters (conv_umma2.cu CHAIN).
    The arithmetic of a tile is unchanged, so the keypoints must equal the one-launch-per-conv plan bit for bit -- at a full
    batch, at odd batches (dead peer tiles, partial last M-supers), and on every repeat (the counters re-arm themselves)."""
    from perseus_b200 import _lib

    from perseus_b200.detector import KeypointCNN

    sd = do.seeded_state_dict(3, 8, seed=0)

    def fresh():  # the plan reads PB200_NO_CHAIN when it is created
        m = KeypointCNN(num_channels=3, precision=precision, max_batch=B)
        m.load_state_dict(sd)
        return m.eval().to(gu.dev())

    monkeypatch.delenv("PB200_NO_CHAIN", raising=False)
    chained = fresh()
    x = gu.synth_u8_frames(B, seed=21).cuda()
    got = [chained.forward_frames(x).clone() for _ in range(3)]
    n_chain = _lib.lib().pb200_plan_launches_at(chained._plan(gu.dev(), B).handle, B)
    monkeypatch.setenv("PB200_NO_CHAIN", "1")
    plain = fresh()
    want = plain.forward_frames(x)
    n_plain = _lib.lib().pb200_plan_launches_at(plain._plan(gu.dev(), B).handle, B)
    print(f"[chain {precision} B={B}] {n_chain} launches chained vs {n_plain} one per conv")
    if B >= 255:  # smaller batches may select other tile shapes (no pair<256> run to chain)
        assert n_chain < n_plain, f"no chain was built at B = {B} ({n_chain} vs {n_plain} launches)"
    assert torch.isfinite(want).all()
    for g in got:
        assert torch.equal(g, want)
