"""A/B of the dynamics kernel's register cap at 1 M pairs: python tools/factors_ab.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from perseus_b200.smoother import pose_dynamics_batch
dev = torch.device("cuda", 0)
n = 1_000_000
g = torch.Generator().manual_seed(0)
q = torch.nn.functional.normalize(torch.randn(2, n, 4, generator=g, dtype=torch.float64), dim=-1)
w, x, y, z = q.unbind(-1)
R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w), 2 * (x * y + z * w), 1 - 2 * (x * x + z * z),
                 2 * (y * z - x * w), 2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1)
t = torch.randn(2, n, 3, generator=g, dtype=torch.float64) * 0.05
pose = torch.cat([R, t], -1).to(dev)
wv, vv = torch.randn(n, 3, generator=g, dtype=torch.float64).to(dev), torch.randn(n, 3, generator=g, dtype=torch.float64).to(dev)
res = {}
ref = None
for minb in (sys.argv[1:] or ("1", "3", "4", "5", "6")):
    os.environ["PB200_DYN_MINB"] = minb
    for _ in range(3):
        r, H = pose_dynamics_batch(pose[0], wv, vv, pose[1], 0.1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        r, H = pose_dynamics_batch(pose[0], wv, vv, pose[1], 0.1)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    if ref is None:
        ref = (r.clone(), H[0].clone())
    res[f"minb{minb}"] = {"ms": ms, "GBps": n * 1152 / (ms * 1e-3) / 1e9, "same": bool(torch.allclose(r, ref[0], atol=1e-12) and torch.allclose(H[0], ref[1], atol=1e-12))}
print(json.dumps(res))
