"""Small invocations of every kernel added in round 2, for compute-sanitizer --tool memcheck (racecheck / initcheck are not
run: one tool per call, B200_PROFILING.md).  python tools/sanitize_smoke.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np
import torch

from oracle import smoother_oracle as so
from perseus_b200 import _lib
from perseus_b200.detector import BatchPipeline, KeypointCNN
from perseus_b200.detector.synthetic import random_init_state_dict
from perseus_b200.smoother import FixedLagSmoother
from perseus_b200.smoother.linearize import chain_factors

dev = torch.device("cuda", 0)
lib = _lib.lib()
s = _lib.current_stream_ptr
g = torch.Generator().manual_seed(0)
# conv_halo (with residual + fused projection, odd batch)
B, H, W, C = 3, 32, 32, 128
x = torch.randn(B, H, W, C, generator=g).bfloat16().to(dev)
xds = torch.randn(B, 2 * H, 2 * W, 64, generator=g).bfloat16().to(dev)
res = torch.randn(B, H, W, 128, generator=g).bfloat16().to(dev)
w = (torch.randn(128, C, 3, 3, generator=g) / 34).contiguous()
wds = (torch.randn(128, 64, generator=g) / 8).contiguous()
bias = torch.randn(128, generator=g)
out = torch.empty(B, H, W, 128, dtype=torch.bfloat16, device=dev)
_lib.check(lib.pb200_debug_conv_halo(x.data_ptr(), B, H, W, C, w.data_ptr(), bias.data_ptr(), 128, res.data_ptr(), 1, out.data_ptr(),
                                     xds.data_ptr(), 64, wds.data_ptr(), s()))
# fp16x2 pair conv with fused projection, KT weights
x16, xds16 = x.half(), xds.half()
out16 = torch.empty(B, H, W, 128, dtype=torch.float16, device=dev)
os.environ["PB200_FORCE_CLUSTER"] = "p"
_lib.check(lib.pb200_debug_conv_f16x2(x16.data_ptr(), B, H, W, C, w.data_ptr(), bias.data_ptr(), 128, 3, 1, 1, None, 1, out16.data_ptr(),
                                      xds16.data_ptr(), 64, wds.data_ptr(), s()))
os.environ.pop("PB200_FORCE_CLUSTER")
# whole forwards (stem x2, conv_p8 x2, pair<256,x2>, halo inside the bf16 plan), through the pipeline + smoother chain
for precision in ("fp16x2", "bf16"):
    m = KeypointCNN(precision=precision, max_batch=4)
    m.load_state_dict(random_init_state_dict(3, 8, seed=0))
    m.eval().to(dev)
    rgb = torch.randint(0, 256, (4, 256, 256, 3), dtype=torch.uint8, generator=g)
    px = m.forward_frames(rgb.to(dev))
    assert torch.isfinite(px).all()
c = so.synthetic_chain(6, seed=1)
f64 = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
sm = FixedLagSmoother(2, 6, f64(c["pts_body"]), c["K5"], c["dt"], iters=2, lam=1e-3, prior=1.0)
pipe = BatchPipeline(m, 2, 376, 672, 4, bgr=True, post=sm.push)
sm.poses[..., 11] = 0.45
fr = torch.randint(0, 256, (2, 376, 672, 4), dtype=torch.uint8, generator=g).pin_memory()
for _ in range(3):
    pose = pipe.result_post(pipe.submit(fr))
assert torch.isfinite(pose).all()
# fused chain factors, both launch shapes
for Wn, Tn in ((5, 7), (300, 2)):
    R = torch.eye(3, dtype=torch.float64).reshape(1, 1, 9).repeat(Wn, Tn, 1)
    t = torch.zeros(Wn, Tn, 3, dtype=torch.float64)
    t[..., 2] = 0.45
    poses = torch.cat([R, t], -1).to(dev)
    z = torch.randn(Wn, Tn, 3, dtype=torch.float64, generator=g).to(dev)
    meas = (torch.rand(Wn, Tn, 8, 2, dtype=torch.float64, generator=g) * 255).to(dev)
    f = chain_factors(poses, z, z, meas, f64(c["pts_body"]), f64(c["K5"]), 0.1)
    assert torch.isfinite(f["dyn_r"]).all()
torch.cuda.synchronize()
print("sanitize smoke ok")
