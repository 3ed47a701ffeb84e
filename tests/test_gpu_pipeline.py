"""-m gpu: BatchPipeline (pinned host batch -> H2D on a copy stream -> one CUDA-graph replay -> pinned keypoints) returns,
bit for bit, what the eager public path `forward_frames` returns, for every in-flight batch, with one or two copy
streams, RGB and RGBD, and keeps doing so when the module's own plan is regrown or its weights are reloaded."""
import numpy as np
import pytest
import torch

from oracle import detector_oracle as do
from tests import gpu_util as gu

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("C,precision,copy_streams", [(3, "fp16x2", 1), (3, "bf16", 2), (4, "fp16x2", 2)])
def test_pipeline_matches_eager_forward_frames(C, precision, copy_streams):
    from perseus_b200.detector import BatchPipeline

    model, sd = gu.native_model(C, precision, max_batch=8)
    B, n_batches = 6, 7
    pipe = BatchPipeline(model, B, 256, 256, 3, depth_divisor=0.035, copy_streams=copy_streams)
    rgbs = [gu.synth_u8_frames(B, seed=40 + i).pin_memory() for i in range(n_batches)]
    deps = [gu.synth_depth_m(B, seed=40 + i).pin_memory() if C == 4 else None for i in range(n_batches)]
    want = [model.forward_frames(r.cuda(), None if d is None else d.cuda(), depth_divisor=0.035).cpu() for r, d in zip(rgbs, deps)]
    got = {}
    for i in range(n_batches):  # two batches in flight; results read with a lag of two
        t = pipe.submit(rgbs[i], deps[i])
        assert t == i
        if i >= 2:
            got[i - 2] = pipe.result(i - 2).clone()
    got[n_batches - 2] = pipe.result(n_batches - 2).clone()
    got[n_batches - 1] = pipe.result(n_batches - 1).clone()
    for i in range(n_batches):
        assert torch.equal(got[i], want[i]), f"batch {i}"
    with pytest.raises(ValueError):
        pipe.result(0)  # fell out of the result ring
    with pytest.raises(ValueError):
        pipe.submit(rgbs[0][:3], None if deps[0] is None else deps[0][:3])


def test_pipeline_survives_replan_and_sees_new_weights():
    from perseus_b200.detector import BatchPipeline, KeypointCNN

    sd = do.seeded_state_dict(3, 8, seed=0)
    model = KeypointCNN(max_batch=2)
    model.load_state_dict(sd)
    model.eval().cuda()
    pipe = BatchPipeline(model, 2)
    rgb = gu.synth_u8_frames(2, seed=5).pin_memory()
    first = pipe.result(pipe.submit(rgb)).clone()
    model.forward_frames(gu.synth_u8_frames(9, seed=6).cuda())  # regrows the module's shared plan
    torch.cuda.synchronize()
    assert torch.equal(pipe.result(pipe.submit(rgb)), first)
    model.load_state_dict(do.seeded_state_dict(3, 8, seed=9))
    second = pipe.result(pipe.submit(rgb)).clone()
    assert not torch.equal(second, first)
    assert torch.equal(second, model.forward_frames(rgb.cuda()).cpu())
