"""Batch sharding across the GPUs of one box (one process per GPU, no data-path collective).

BASELINE.json north_star: "Batched inference shards by batch across the 8 GPUs of one box with no
collective".  The reference itself only runs single-device inference (validate.py:109-129); its one
multi-process pattern is train.py:122-152 (one process per GPU, NCCL, DistributedSampler), which is
the launch model mirrored here.  Frames (and factor batches) are independent, so rank r simply takes
the contiguous slice `shard_range(n, r, world)`; the only collectives are the optional result gather and
the max-over-ranks of a timing, both off the data path.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(n: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous [lo, hi) slice of n items for `rank`: sizes differ by at most one, earlier ranks get
    the extra item (same convention as torch.tensor_split)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard(t: torch.Tensor, rank: int | None = None, world: int | None = None) -> torch.Tensor:
    """This rank's slice of dim 0 (a view, no copy)."""
    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    lo, hi = shard_range(t.shape[0], rank, world)
    return t[lo:hi]


def gather_shards(local: torch.Tensor, n_total: int, group=None) -> torch.Tensor:
    """All-gather per-rank results (ragged in dim 0) back into global order.  Convenience for callers that
    want every keypoint on every rank; NOT used inside the timed region of bench.py."""
    world = dist.get_world_size(group)
    sizes = [shard_range(n_total, r, world) for r in range(world)]
    mx = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros((mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad, group=group)
    return torch.cat([o[: hi - lo] for o, (lo, hi) in zip(out, sizes)], dim=0)


def max_over_ranks(value: float, device: torch.device | str = "cpu", group=None) -> float:
    """The slowest rank's time: every multi-GPU number is the max over ranks of a device-side timing."""
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def sharded_forward_frames(model, rgb_u8: torch.Tensor, depth: torch.Tensor | None = None, **ingest_kwargs):
    """Rank-local part of a batch-sharded `forward_frames`: slices the GLOBAL batch (already on this rank's
    device or pinned host memory) and runs the native path on the slice.  Returns (pixels, (lo, hi))."""
    rank, world = dist.get_rank(), dist.get_world_size()
    lo, hi = shard_range(rgb_u8.shape[0], rank, world)
    if hi == lo:
        return torch.empty((0, model.n_keypoints, 2), dtype=torch.float32, device=rgb_u8.device), (lo, hi)
    d = None if depth is None else depth[lo:hi]
    return model.forward_frames(rgb_u8[lo:hi], d, **ingest_kwargs), (lo, hi)
