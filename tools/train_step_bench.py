"""BASELINE.json configs[3]: data-parallel training step (forward + backward + clip + AdamW) on synthetic augmented
frames at N GPUs with NCCL gradient all-reduce.  Mirrors perseus/detector/train.py:259-304 (autocast forward,
SmoothL1Loss(beta=1), backward through DDP, clip_grad_norm_(1.0), AdamW(lr=1e-3)); north_star asks for bf16
(the reference uses fp16 + GradScaler; bf16 needs no scaler).

STAGE A: the training-mode forward/backward run on the torch module tree that the drop-in
KeypointCNN exposes (library autograd / cuDNN) wrapped in DistributedDataParallel; the eval-mode forward used for
validation inside the loop is the native sm_100a path.  This script measures the NCCL scaling of that step; it is
NOT the headline metric and says so in its output.

  python tools/train_step_bench.py --steps 20                       # 1 GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
      tools/train_step_bench.py --steps 20
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run_native(args, rank, world, local, dev):
    """STAGE B: the whole step on this repo's kernels (perseus_b200/detector/training.py)."""
    import time

    from perseus_b200.detector import KeypointCNN
    from perseus_b200.detector.training import NativeTrainer

    torch.manual_seed(0)
    model = KeypointCNN(num_channels=args.channels).to(dev)
    tr = NativeTrainer(model, args.batch)
    g = torch.Generator(device=dev).manual_seed(rank)
    xs = [torch.rand(args.batch, args.channels, 256, 256, device=dev, generator=g) for _ in range(2)]
    y = torch.rand(args.batch, 8, 2, device=dev, generator=g) * 2 - 1
    aug = None
    if args.augment:  # train.py:270-284: 5-channel batch (RGB, depth, segmentation) -> augmentation -> model channels
        from perseus_b200.detector.augmentations import AugmentationConfig, KeypointAugmentation

        aug = KeypointAugmentation(AugmentationConfig(), train=True, seed=rank, out_channels=args.channels)
        seg = (torch.rand(args.batch, 1, 256, 256, device=dev, generator=g) < 0.2).float()
        xs5 = [torch.cat([x[:, :3], 4 + 10 * x[:, 3:4] if args.channels > 3 else 4 + 10 * x[:, :1], seg], dim=1).contiguous()
               for x in xs]
        kp_px = (y + 1) * 127.5
        step_plain = tr.step

        def step_aug(x, _y):
            xa, ya = aug(xs5[0] if x is xs[0] else xs5[1], kp_px)
            return step_plain(xa, ya.reshape(args.batch, 8, 2))

        tr.step = step_aug
    for i in range(args.warmup):
        tr.step(xs[i % 2], y)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for i in range(args.steps):
        loss = tr.step(xs[i % 2], y)
    e1.record()
    host_ms = (time.perf_counter() - t0) * 1e3
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    prof = None
    if args.profile and rank == 0:
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ev[0].record(); tr.forward(xs[0]); ev[1].record(); tr.backward(y); ev[2].record(); tr.optimizer_step(); ev[3].record()
        torch.cuda.synchronize()
        prof = {"forward_ms": ev[0].elapsed_time(ev[1]), "backward_ms": ev[1].elapsed_time(ev[2]),
                "allreduce_clip_adamw_ms": ev[2].elapsed_time(ev[3])}
        ops = tr.profile_step(xs[0], y)
        prof["per_operator_ms"] = {k: round(v, 3) for k, v in sorted(ops.items(), key=lambda kv: -kv[1])}
    model.eval()
    with torch.no_grad():
        out = model(xs[0][:8])
    if rank == 0:
        print(json.dumps({
            "metric": "train_step_frames_per_sec (stage B: native sm_100a kernels, fwd+loss+bwd+allreduce+clip+AdamW)",
            "value": world * args.batch * args.steps / (ms * 1e-3), "unit": "frames/s", "n_gpus": world,
            "ms_per_step": ms / args.steps, "host_enqueue_ms_per_step": host_ms / args.steps, "per_gpu_batch": args.batch,
            "dtype": "bf16 activations, fp32 params/grads/optimizer", "channels": args.channels,
            "allreduce_bytes_per_step": 4 * tr.n_params, "loss": float(loss), "scaling": "weak", "phases": prof, "augmentation_in_step": bool(args.augment),
            "native_eval_forward_ok": bool(torch.isfinite(out).all()),
        }), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=256)  # per GPU, train.py:39
    ap.add_argument("--channels", type=int, default=4)  # train.py default is RGBD
    ap.add_argument("--impl", default="native", choices=["native", "library"])
    ap.add_argument("--augment", action="store_true", help="native: fused train-time augmentation inside the timed step")
    ap.add_argument("--profile", action="store_true", help="native: per-kernel-class CUDA-event times of one step")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)

    from perseus_b200.detector import KeypointCNN

    if args.impl == "native":
        return run_native(args, rank, world, local, dev)

    torch.manual_seed(0)
    model = KeypointCNN(num_channels=args.channels).to(dev).to(memory_format=torch.channels_last)
    model.train()
    net = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local]) if world > 1 else model
    opt = torch.optim.AdamW(net.parameters(), lr=1e-3)  # train.py:199
    loss_fn = torch.nn.SmoothL1Loss(beta=1.0)  # train.py:119
    g = torch.Generator(device=dev).manual_seed(rank)
    x = torch.rand(args.batch, args.channels, 256, 256, device=dev, generator=g).contiguous(memory_format=torch.channels_last)
    y = torch.rand(args.batch, 8, 2, device=dev, generator=g) * 2 - 1

    def step():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            pred = net(x).reshape(-1, 8, 2)  # train.py:287-289
            loss = loss_fn(pred, y)
        opt.zero_grad(set_to_none=True)
        loss.backward()  # DDP all-reduces the 11.19 M gradients in buckets, overlapped with backward
        torch.nn.utils.clip_grad_norm_(net.parameters(), 1.0)  # train.py:302
        opt.step()
        return loss

    for _ in range(args.warmup):
        step()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    # the eval-mode forward of the SAME module (validation inside the training loop, train.py:334) is the native path
    model.eval()
    with torch.no_grad():
        out = model(x[:8].float().contiguous())
    model.train()
    if rank == 0:
        n_params = sum(p.numel() for p in model.parameters())
        print(json.dumps({
            "metric": "train_step_frames_per_sec (stage A: library autograd + DDP/NCCL; not the headline)",
            "value": world * args.batch * args.steps / (ms * 1e-3), "unit": "frames/s", "n_gpus": world,
            "ms_per_step": ms / args.steps, "per_gpu_batch": args.batch, "dtype": "bf16 autocast",
            "allreduce_bytes_per_step": 4 * n_params, "loss": float(loss), "scaling": "weak",
            "native_eval_forward_ok": bool(torch.isfinite(out).all()),
        }), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
