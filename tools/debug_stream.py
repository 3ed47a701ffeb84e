import torch, time, sys
sys.path.insert(0, '.')
from oracle import detector_oracle as do
from perseus_b200.detector import KeypointCNN, StreamingDetector
dev = torch.device('cuda', 0)
m = KeypointCNN(num_channels=3, precision='bf16', max_batch=2)
m.load_state_dict(do.seeded_state_dict(3, 8, seed=0)); m.eval().to(dev)
det = StreamingDetector(m, 376, 672, 4, batch=1)
g = torch.Generator().manual_seed(1)
bgra = torch.randint(0, 256, (1, 376, 672, 4), dtype=torch.uint8, generator=g)
got = det.infer(bgra.numpy())
want = m.forward_frames(bgra.cuda(), None, bgr=True, depth_divisor=0.035).cpu()
print('got', got[0, :2], 'want', want[0, :2])
det.graph.debug_dump if hasattr(det.graph, 'debug_dump') else None
t0 = time.perf_counter(); det.infer_pinned(); print('replay ms', (time.perf_counter() - t0) * 1e3)
# eager step on the stream for comparison
with torch.cuda.stream(det.stream):
    t0 = time.perf_counter(); det._step(); det.stream.synchronize(); print('eager ms', (time.perf_counter() - t0) * 1e3)
print('host_px after eager', det.host_px[0, :2])
