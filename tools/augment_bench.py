"""Times the fused train-time augmentation (reference perseus/detector/augmentations.py:434-488 as
train.py:270-284 calls it) on a synthetic (B, 5, 256, 256) batch: plasma map + transplant count + fused kernel + key
points, default AugmentationConfig, host-side parameter sampling included in the `e2e` figure.

    python tools/augment_bench.py [--batch 256] [--steps 50] [--cpu-sample 8]

Algorithmic bytes per pixel: 5 planes x 4 B of the acceptor + 5 x 4 B of the donor read, 4 x 4 B written = 56 B
(3.67 MB / frame at 256 x 256); the warp's 4-tap gather re-reads come from L1/L2.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--cpu-sample", type=int, default=8)
    args = ap.parse_args()
    from oracle.make_golden_augment import synthetic_batch
    from perseus_b200.detector import augmentations as host

    dev = torch.device("cuda", 0)
    B, H, W = args.batch, 256, 256
    base = synthetic_batch(16, H, W, 0)
    xs = [base[torch.randperm(16)].repeat(B // 16 + 1, 1, 1, 1)[:B].contiguous().to(dev) for _ in range(2)]  # 2 x 336 MB > L2
    kp = (torch.rand(B, 16) * 255).to(dev)
    cfg = host.AugmentationConfig()
    kc = host.kernel_config(cfg, True)
    g = torch.Generator().manual_seed(0)
    # ---- device-only: parameters resident, plasma + augment launches
    ip, fp, rough = host.sample_params(cfg, True, B, 5, H, W, g)
    ip_d, fp_d, rough_d = torch.from_numpy(ip).to(dev), torch.from_numpy(fp).to(dev), torch.from_numpy(rough).to(dev)

    def device_step(i):
        plasma = host.plasma_map(rough_d, i, H, W)
        return host.augment_batch(xs[i % 2], ip_d, fp_d, plasma, kc, i, kp, c_out=4)

    for i in range(args.warmup):
        device_step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        device_step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    # ---- the fused kernel alone
    plasma = host.plasma_map(rough_d, 0, H, W)
    torch.cuda.synchronize()
    e0.record()
    for i in range(args.steps):
        host.augment_batch(xs[i % 2], ip_d, fp_d, plasma, kc, i, kp, c_out=4)
    e1.record()
    torch.cuda.synchronize()
    ms_k = e0.elapsed_time(e1) / args.steps
    # ---- through the drop-in module (host sampling + table upload every call)
    aug = host.KeypointAugmentation(cfg, train=True, seed=0)
    for i in range(3):
        aug(xs[i % 2], kp)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(args.steps):
        out, coords = aug(xs[i % 2], kp)
    torch.cuda.synchronize()
    ms_mod = (time.perf_counter() - t0) * 1e3 / args.steps
    # ---- CPU baseline: the oracle (torch CPU restatement of the reference pipeline) on a bounded sample
    from oracle import augment_oracle as ao

    n = args.cpu_sample
    xc = xs[0][:n].cpu()
    ipc, fpc, rc = host.sample_params(cfg, True, n, 5, H, W, g)
    t0 = time.perf_counter()
    pl = torch.from_numpy(ao.plasma(rc, 0, H, W))
    ao.augment(xc, ipc, fpc, pl, kc, 0, kp[:n].cpu(), c_out=4)
    cpu_s = time.perf_counter() - t0
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    peak = float(peaks.get("hbm_gbs", 6553.0))
    alg = B * H * W * 56
    print(json.dumps({
        "metric": "augmented_frames_per_sec (default AugmentationConfig, 5-channel 256x256 batch -> 4 channels + key points)",
        "value": B / (ms * 1e-3), "unit": "frames/s", "batch": B, "ms_per_batch": ms, "ms_fused_kernel_only": ms_k,
        "ms_through_module_incl_host_sampling": ms_mod,
        "roofline": {"bound": "hbm", "achieved": alg / (ms_k * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                     "frac": alg / (ms_k * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_frame": H * W * 56,
                     "kernel": "augment_kernel (+ transplant_count_kernel, aug_keypoints_kernel)"},
        "cpu_baseline": {"value": n / cpu_s, "unit": "frames/s", "cores": torch.get_num_threads(), "kind": "port",
                         "sample": f"{n} frames, oracle/augment_oracle.py (torch CPU restatement), {cpu_s:.2f} s"},
        "finite": bool(torch.isfinite(out).all()) and bool(torch.isfinite(coords).all()),
    }), flush=True)


if __name__ == "__main__":
    main()
