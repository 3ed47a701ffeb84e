"""Per-operator CUDA-event times of one forward in any precision mode:
python tools/profile_ops.py [precision=bf16x3] [B=256] [C=3]."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

from perseus_b200.detector import KeypointCNN
from perseus_b200.detector.synthetic import random_init_state_dict

precision = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
C = int(sys.argv[3]) if len(sys.argv) > 3 else 3
dev = torch.device("cuda", 0)
m = KeypointCNN(num_channels=C, precision=precision, max_batch=B)
m.load_state_dict(random_init_state_dict(C, 8, seed=0))
m.eval().to(dev)
g = torch.Generator().manual_seed(0)
rgb = torch.randint(0, 256, (B, 256, 256, 3), dtype=torch.uint8, generator=g).to(dev)
depth = torch.rand(B, 256, 256, generator=g).to(dev) * 11 + 3 if C == 4 else None
for _ in range(3):
    m.forward_frames(rgb, depth)
frames = m.ingest(rgb, depth)
best = None
for _ in range(5):
    rows = m.profile_ops(frames, B)
    if best is None or sum(r[1] for r in rows) < sum(r[1] for r in best):
        best = rows
for _ in range(5):
    m.forward_frames(rgb, depth)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
torch.cuda.synchronize()
ev[0].record()
for _ in range(30):
    m.forward_frames(rgb, depth)
ev[1].record()
torch.cuda.synchronize()
print(json.dumps({"precision": precision, "B": B, "C": C, "ms_per_step": ev[0].elapsed_time(ev[1]) / 30,
                  "sum_op_ms": sum(r[1] for r in best),
                  "ops": [{"op": r[0], "ms": round(r[1], 4), "kernel": r[4],
                           "TFLOPs": round(r[2] / r[1] / 1e9, 1) if r[1] > 0 else None} for r in best]}, indent=1))
