"""-m gpu: training-step operators (pb200_train_*) vs torch autograd on the same bf16 operands.

Tolerances: tensor-core GEMMs on identical bf16 operands differ from the fp64 result only by fp32 accumulation
(relative ~1e-4 of the sum of |terms|); outputs stored in bf16 add 2^-8 relative.
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _lib():
    from perseus_b200 import _lib as L

    return L


def _stream():
    return _lib().current_stream_ptr()


WGRAD_CASES = [
    # B, H, W, Cin, Cout, R, stride, pad
    (3, 64, 64, 64, 64, 3, 1, 1),     # layer1: 9 units -> 5 pairs (odd tail), N = 64
    (4, 64, 64, 64, 128, 3, 2, 1),    # layer2.0.conv1: stride 2 parity views
    (4, 64, 64, 64, 128, 1, 2, 0),    # layer2.0.downsample: 1 unit, half-empty pair
    (5, 32, 32, 128, 128, 3, 1, 1),
    (6, 16, 16, 256, 256, 3, 1, 1),   # N = 256
    (9, 8, 8, 512, 512, 3, 1, 1),     # image smaller than a k-block (2 images per tile, ragged batch)
    (3, 16, 16, 256, 512, 3, 2, 1),
]


@pytest.mark.parametrize("case", WGRAD_CASES, ids=[str(c) for c in WGRAD_CASES])
def test_wgrad_tcgen05_vs_autograd(case):
    L = _lib()
    B, H, W, Cin, Cout, R, stride, pad = case
    g = torch.Generator().manual_seed(sum(case))
    Ho, Wo = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - R) // stride + 1
    x = torch.randn(B, H, W, Cin, generator=g).bfloat16()
    dz = torch.randn(B, Ho, Wo, Cout, generator=g).bfloat16()
    want = torch.nn.grad.conv2d_weight(x.double().permute(0, 3, 1, 2), (Cout, Cin, R, R),
                                       dz.double().permute(0, 3, 1, 2), stride=stride, padding=pad)
    xd, dzd = x.cuda(), dz.cuda()
    scratch = torch.zeros(R * R, Cin, Cout, dtype=torch.float32, device="cuda")
    dw = torch.full((Cout, Cin, R, R), float("nan"), dtype=torch.float32, device="cuda")
    L.check(L.lib().pb200_train_wgrad(xd.data_ptr(), dzd.data_ptr(), B, H, W, Cin, Cout, R, stride, pad,
                                      scratch.data_ptr(), _stream()))
    L.check(L.lib().pb200_train_unpack_wgrad(scratch.data_ptr(), Cout, Cin, R, Cin, dw.data_ptr(), _stream()))
    torch.cuda.synchronize()
    got = dw.double().cpu()
    assert torch.isfinite(got).all()
    err = (got - want).abs()
    tol = 2e-3 * (B * Ho * Wo) ** 0.5 + 1e-4 * want.abs()
    assert (err <= tol).all(), f"max err {float(err.max()):.4f} (|want| max {float(want.abs().max()):.2f}); " \
                               f"bad {int((err > tol).sum())}/{err.numel()} first {(err > tol).nonzero()[:4].tolist()}"


DGRAD_CASES = [
    (2, 64, 64, 64, 64, 3),
    (3, 16, 16, 256, 256, 3),
    (5, 8, 8, 512, 512, 3),
]


@pytest.mark.parametrize("case", DGRAD_CASES, ids=[str(c) for c in DGRAD_CASES])
def test_dgrad_stride1_is_a_conv_with_the_flipped_transposed_filter(case):
    """data gradient of a 3x3/1 conv = pb200_train_conv on dz with weights packed in mode 1."""
    L = _lib()
    B, H, W, Cin, Cout, R = case
    g = torch.Generator().manual_seed(sum(case))
    w = (torch.randn(Cout, Cin, R, R, generator=g) / (R * R * Cin) ** 0.5)
    dz = torch.randn(B, H, W, Cout, generator=g).bfloat16()
    wb = w.bfloat16().double()
    want = torch.nn.grad.conv2d_input((B, Cin, H, W), wb, dz.double().permute(0, 3, 1, 2), stride=1, padding=1)
    want = want.permute(0, 2, 3, 1)
    wd = w.cuda().contiguous()
    packed = torch.empty(Cin * R * R * Cout, dtype=torch.bfloat16, device="cuda")
    L.check(L.lib().pb200_train_pack_weights(wd.data_ptr(), Cout, Cin, R, 1, packed.data_ptr(), _stream()))
    zero_bias = torch.zeros(512, dtype=torch.float32, device="cuda")
    out = torch.full((B, H, W, Cin), float("nan"), dtype=torch.bfloat16, device="cuda")
    dzd = dz.cuda()
    # dgrad: input channels = Cout, output channels = Cin
    L.check(L.lib().pb200_train_conv(dzd.data_ptr(), B, H, W, Cout, packed.data_ptr(), zero_bias.data_ptr(), Cin, R, 1, 1,
                                     None, 0, out.data_ptr(), _stream()))
    torch.cuda.synchronize()
    got = out.double().cpu()
    assert torch.isfinite(got).all()
    err = (got - want).abs()
    tol = 2.0 ** -7 * want.abs() + 1e-2
    assert (err <= tol).all(), f"max err {float(err.max()):.4f}"
