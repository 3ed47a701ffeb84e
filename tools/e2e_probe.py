"""Where the e2e step loses time against the resident step: eager forward vs CUDA-graph replay of the same launches,
with and without a concurrent H2D copy stream.  python tools/e2e_probe.py [precision=bf16]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

from perseus_b200.detector import KeypointCNN
from perseus_b200.detector.synthetic import random_init_state_dict

precision = sys.argv[1] if len(sys.argv) > 1 else "bf16"
B = 256
dev = torch.device("cuda", 0)
m = KeypointCNN(num_channels=3, precision=precision, max_batch=B)
m.load_state_dict(random_init_state_dict(3, 8, seed=0))
m.eval().to(dev)
g = torch.Generator().manual_seed(0)
host = [torch.randint(0, 256, (B, 256, 256, 3), dtype=torch.uint8, generator=g).pin_memory() for _ in range(2)]
rgb = [h.to(dev) for h in host]
plan = m.private_plan(dev, B)
frames = torch.empty(plan.frames_bytes(B), dtype=torch.uint8, device=dev)
on, op = torch.empty(B, 16, device=dev), torch.empty(B, 8, 2, device=dev)
st = torch.cuda.Stream(device=dev)
cp = torch.cuda.Stream(device=dev)


def fwd():
    m.forward_ingested(frames, B, plan=plan, out_norm=on, out_px=op)


def timeit(fn, n=100):
    with torch.cuda.stream(st):
        for _ in range(5):
            fn()
        st.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


res = {}
with torch.cuda.stream(st):
    m.ingest(rgb[0], out=frames, plan=plan)
    fwd()
st.synchronize()
res["eager_forward_ms"] = timeit(fwd)
gr = torch.cuda.CUDAGraph()
with torch.cuda.graph(gr, stream=st):
    fwd()
res["graph_forward_ms"] = timeit(gr.replay)
gr2 = torch.cuda.CUDAGraph()
with torch.cuda.graph(gr2, stream=st):
    m.ingest(rgb[0], out=frames, plan=plan)
    fwd()
res["graph_ingest_forward_ms"] = timeit(gr2.replay)


def eager_both():
    m.ingest(rgb[0], out=frames, plan=plan)
    fwd()


res["eager_ingest_forward_ms"] = timeit(eager_both)
# the same with a copy stream streaming pinned batches host -> device all the time
dst = [torch.empty_like(rgb[0]) for _ in range(2)]
stop = False


def with_copies(fn, n=100):
    with torch.cuda.stream(cp):
        for i in range(int(n * 1.6) + 20):
            dst[i % 2].copy_(host[i % 2], non_blocking=True)
    return timeit(fn, n)


res["graph_forward_ms_with_h2d_stream"] = with_copies(gr.replay)
torch.cuda.synchronize()
res["eager_forward_ms_with_h2d_stream"] = with_copies(fwd)
torch.cuda.synchronize()
os.environ["PB200_NO_PDL"] = "1"
print(json.dumps(res, indent=1))
