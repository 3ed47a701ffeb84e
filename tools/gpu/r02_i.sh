#!/bin/bash
# fused ingest (stem reads uint8): correctness, then A/B against ingest + forward
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_fused_ingest.py -q -m gpu -x > gpurun_out/i_tests.log 2>&1; rc=$?; echo "fused tests exit $rc"; tail -6 gpurun_out/i_tests.log
if [ $rc -ne 0 ]; then exit 0; fi
python - <<'PY'
import torch, os, sys
sys.path.insert(0, os.getcwd())
from perseus_b200.detector import KeypointCNN
from perseus_b200.detector.synthetic import random_init_state_dict
dev = torch.device("cuda", 0)
for prec in ("bf16", "fp16x2"):
    m = KeypointCNN(num_channels=3, precision=prec, max_batch=256); m.load_state_dict(random_init_state_dict(3, 8, seed=0)); m.eval().to(dev)
    g = torch.Generator().manual_seed(0)
    bufs = [torch.randint(0, 256, (256, 256, 256, 3), dtype=torch.uint8, generator=g).to(dev) for _ in range(4)]  # 201 MB > L2
    def run(fused, n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for i in range(5):
            (m.forward_rgb8(bufs[i % 4]) if fused else m.forward_ingested(m.ingest(bufs[i % 4], None), 256))
        torch.cuda.synchronize(); e0.record()
        for i in range(n):
            (m.forward_rgb8(bufs[i % 4]) if fused else m.forward_ingested(m.ingest(bufs[i % 4], None), 256))
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    for rep in range(2):
        a, b = run(False, 40), run(True, 40)
        print(f"{prec}: ingest + forward {a:.4f} ms ({256 / a:.1f} k frames/s) | fused {b:.4f} ms ({256 / b:.1f} k frames/s)")
PY
