"""bs-1 / bs-2 streaming latency only (bench.py's measure_latency) - for A/B runs of kernel variants."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

import bench

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
model = bench.build_model(dev, 3, "bf16", 256) if hasattr(bench, "build_model") else None
if model is None:
    from perseus_b200.detector import KeypointCNN
    from perseus_b200.detector.synthetic import random_init_state_dict

    model = KeypointCNN(num_channels=3, precision="bf16", max_batch=256)
    model.load_state_dict(random_init_state_dict(3, 8, seed=0))
    model.eval().to(dev)
print(json.dumps(bench.measure_latency(model, dev)))
