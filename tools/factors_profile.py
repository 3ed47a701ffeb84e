"""Runs the smoother factor kernels on N poses (default 1M) for ncu captures / CUDA-event timing per kernel."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from perseus_b200.smoother import constant_velocity_batch, keypoint_projection_batch, pose_dynamics_batch

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
dev = torch.device("cuda", 0)
g = torch.Generator().manual_seed(0)
qn = torch.nn.functional.normalize(torch.randn(2, n, 4, generator=g, dtype=torch.float64), dim=-1)
w, x, y, z = qn.unbind(-1)
R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w), 2 * (x * y + z * w), 1 - 2 * (x * x + z * z),
                 2 * (y * z - x * w), 2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1)
t = torch.randn(2, n, 3, generator=g, dtype=torch.float64) * 0.05
t[..., 2] += 0.45
pose = torch.cat([R, t], -1).to(dev)
wv, vv = (torch.randn(n, 3, generator=g, dtype=torch.float64).to(dev) for _ in range(2))
K5 = torch.tensor([221.7, 221.7, 0.0, 128.0, 128.0], dtype=torch.float64, device=dev)
corners = torch.tensor([[a, b, c] for a in (-1, 1) for b in (-1, 1) for c in (-1, 1)], dtype=torch.float64, device=dev) * 0.0175
meas = (torch.rand(n, 8, 2, generator=g, dtype=torch.float64) * 255).to(dev)
fns = {"projection": lambda: keypoint_projection_batch(pose[0], K5, meas, corners),
       "dynamics": lambda: pose_dynamics_batch(pose[0], wv, vv, pose[1], 0.1, "world"),
       "constvel": lambda: constant_velocity_batch(vv, wv)}
nbytes = {"projection": 1120, "dynamics": 1152, "constvel": 72 + 144}
for name, fn in fns.items():
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"{name}: {ms:.3f} ms for {n} poses, {n * nbytes[name] / ms / 1e6:.0f} GB/s algorithmic (incl. output allocation)")
