"""pb200_forward_rgb8: the ingest step fused into the stem kernel (csrc/stem_pool.cu, U8 producer).

The stem converts the uint8 rows itself, so the result must equal pb200_ingest + pb200_forward BIT FOR BIT: the frames
encoding of the 16-bit modes is the exact integer pixel value (bf16 k, fp16 k * 2^-12), the rest of the network is the same
launches.  Covered: both 16-bit modes, RGB and BGR sources, batch 1 / odd / full, batches whose bands start mid-image, and
inputs built to expose a wrong row, column or channel (single bright pixels at the image corners and edges)."""
import pytest
import torch

from oracle import detector_oracle as do
from tests import gpu_util as gu

pytestmark = pytest.mark.gpu


def _model(precision, B):
    from perseus_b200.detector import KeypointCNN

    m = KeypointCNN(num_channels=3, precision=precision, max_batch=B)
    m.load_state_dict(do.seeded_state_dict(3, 8, seed=0))
    return m.eval().to(gu.dev())


@pytest.mark.parametrize("precision", ["bf16", "fp16x2"])
@pytest.mark.parametrize("B,bgr", [(1, False), (5, True), (37, False), (256, False), (256, True)])
def test_fused_ingest_bit_equal_to_ingest_plus_forward(precision, B, bgr, monkeypatch):
    m = _model(precision, B)
    x = gu.synth_u8_frames(B, seed=31 + B).cuda()
    assert m.fused_rgb8_ok(x, {"bgr": bgr})
    norm, px = m.forward_rgb8(x, bgr=bgr)
    frames = m.ingest(x, None, bgr=bgr)
    want_norm, want_px = m.forward_ingested(frames, B)
    assert torch.isfinite(want_px).all()
    assert torch.equal(norm, want_norm) and torch.equal(px, want_px)
    # forward_frames takes the fused path when asked to (opt-in); repeat calls are stable
    monkeypatch.setenv("PB200_FUSED_INGEST", "1")
    assert torch.equal(m.forward_frames(x, bgr=bgr), want_px)
    assert torch.equal(m.forward_frames(x, bgr=bgr), want_px)
    monkeypatch.delenv("PB200_FUSED_INGEST")
    assert torch.equal(m.forward_frames(x, bgr=bgr), want_px)


@pytest.mark.parametrize("precision", ["bf16", "fp16x2"])
def test_fused_ingest_places_every_row_column_and_channel(precision):
    """One bright pixel per image at corners / edges / band boundaries, one channel each: a transposed, shifted or
    channel-swapped conversion changes the keypoints of that image."""
    spots = [(0, 0, 0), (0, 255, 1), (255, 0, 2), (255, 255, 0), (1, 1, 1), (127, 128, 2), (63, 64, 0), (64, 63, 1),
             (3, 250, 2), (252, 5, 0), (31, 32, 1), (32, 31, 2), (128, 0, 0), (0, 128, 1), (200, 17, 2), (17, 200, 0)]
    B = len(spots)
    m = _model(precision, B)
    x = torch.full((B, 256, 256, 3), 7, dtype=torch.uint8)
    for i, (r, c, ch) in enumerate(spots):
        x[i, r, c, ch] = 255
    x = x.cuda()
    _, px = m.forward_rgb8(x)
    _, want = m.forward_ingested(m.ingest(x, None), B)
    assert torch.equal(px, want)
    assert len({tuple(px[i].flatten().tolist()) for i in range(B)}) == B, "every spot must change the output"


def test_fused_ingest_refuses_what_it_cannot_do():
    from perseus_b200 import _lib
    from perseus_b200.detector import KeypointCNN

    m = _model("bf16", 2)
    x = gu.synth_u8_frames(2, seed=3).cuda()
    assert not m.fused_rgb8_ok(x[:, :, :, :2])                      # not 3 channels
    assert not m.fused_rgb8_ok(x.permute(0, 2, 1, 3))               # not contiguous
    assert not m.fused_rgb8_ok(x.float())                           # not uint8
    plan = m._plan(gu.dev(), 2)
    out = torch.empty((2, 16), dtype=torch.float32, device="cuda")
    odd = torch.empty(2 * 256 * 256 * 3 + 16, dtype=torch.uint8, device="cuda")[1:]  # misaligned base
    rc = plan.lib.pb200_forward_rgb8(plan.handle, odd.data_ptr(), 0, plan.workspace.data_ptr(), out.data_ptr(), None, 2,
                                     _lib.current_stream_ptr())
    assert rc != 0 and "aligned" in plan.lib.pb200_last_error().decode()
    m32 = KeypointCNN(num_channels=3, precision="fp32", max_batch=2)
    m32.load_state_dict(do.seeded_state_dict(3, 8, seed=0))
    m32 = m32.eval().to(gu.dev())
    assert not m32.fused_rgb8_ok(x)
    p32 = m32._plan(gu.dev(), 2)
    rc = p32.lib.pb200_forward_rgb8(p32.handle, x.data_ptr(), 0, p32.workspace.data_ptr(), out.data_ptr(), None, 2,
                                    _lib.current_stream_ptr())
    assert rc != 0 and "16-bit" in p32.lib.pb200_last_error().decode()
