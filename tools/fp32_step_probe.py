"""Where the fp32-mode step time goes: ingest alone, forward alone, both (CUDA events, 30 iterations each)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from perseus_b200.detector import KeypointCNN
from perseus_b200.detector.synthetic import random_init_state_dict

precision = sys.argv[1] if len(sys.argv) > 1 else "fp32"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
dev = torch.device("cuda", 0)
m = KeypointCNN(num_channels=3, precision=precision, max_batch=B)
m.load_state_dict(random_init_state_dict(3, 8, seed=0))
m.eval().to(dev)
rgb = torch.randint(0, 256, (B, 256, 256, 3), dtype=torch.uint8).to(dev)


def timed(fn, n=30):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


frames = m.ingest(rgb)
print(precision, "B", B)
print("ingest           ms", timed(lambda: m.ingest(rgb)))
print("forward_ingested ms", timed(lambda: m.forward_ingested(frames, B)))
print("forward_frames   ms", timed(lambda: m.forward_frames(rgb)))
rows = m.profile_ops(frames, B)
print("profile_ops sum  ms", sum(r[1] for r in rows))
rows = m.profile_ops(frames, B)
print("profile_ops sum  ms (2nd)", sum(r[1] for r in rows))
