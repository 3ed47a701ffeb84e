"""Minimal forward loop for ncu captures: python tools/profile_step.py [B] [steps] [C]."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

from perseus_b200.detector import KeypointCNN
from perseus_b200.detector.synthetic import random_init_state_dict

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
C = int(sys.argv[3]) if len(sys.argv) > 3 else 3
dev = torch.device("cuda", 0)
m = KeypointCNN(num_channels=C, precision="bf16", max_batch=B)
m.load_state_dict(random_init_state_dict(C, 8, seed=0))
m.eval().to(dev)
g = torch.Generator().manual_seed(0)
rgb = torch.randint(0, 256, (B, 256, 256, 3), dtype=torch.uint8, generator=g).to(dev)
depth = torch.rand(B, 256, 256, generator=g).to(dev) * 11 + 3 if C == 4 else None
for _ in range(steps):
    px = m.forward_frames(rgb, depth)
torch.cuda.synchronize()
print("ok", float(px.abs().sum()))
