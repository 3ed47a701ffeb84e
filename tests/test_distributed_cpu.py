"""World-size-2 gloo tests (CPU) of the multi-GPU host logic: batch sharding, result gather, max-over-ranks
timing, and the `bench.py --impl reference` rank gating.  The data path itself has no collective."""
import json
import os
import subprocess
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from perseus_b200 import sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_range_partitions_every_size():
    for n in (0, 1, 2, 7, 255, 256, 257, 10_000):
        for world in (1, 2, 3, 4, 8):
            spans = [sharding.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (a, b), (c, d) in zip(spans, spans[1:]):
                assert b == c and a <= b
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(4, 2, 2)


def _worker(rank, world, port, n_total, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(0)
        full = torch.rand(n_total, 8, 2, generator=g)  # stands in for the keypoints of the global batch
        mine = sharding.shard(full)
        lo, hi = sharding.shard_range(n_total, rank, world)
        assert mine.shape[0] == hi - lo
        # the "computation": an independent per-frame function, so sharded == unsharded
        local = mine * 2.0 + 1.0
        gathered = sharding.gather_shards(local, n_total)
        assert torch.equal(gathered, full * 2.0 + 1.0)
        slow = sharding.max_over_ranks(10.0 + rank)
        assert slow == 10.0 + world - 1
        torch.save({"lo": lo, "hi": hi}, os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_total", [256, 7, 1])
def test_world2_gloo_shard_gather_and_timing(tmp_path, n_total):
    port = 29500 + (os.getpid() + n_total) % 400
    mp.spawn(_worker, args=(2, port, n_total, str(tmp_path)), nprocs=2, join=True)
    spans = [torch.load(os.path.join(tmp_path, f"r{r}.pt")) for r in range(2)]
    assert spans[0]["lo"] == 0 and spans[0]["hi"] == spans[1]["lo"] and spans[1]["hi"] == n_total


def test_bench_reference_arm_runs_on_rank0_only():
    """Under torchrun (N > 1) rank 0 alone runs and prints the reference line; other ranks exit 0 silently."""
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                        "--steps", "1", "--warmup", "0"], env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""
    env = dict(os.environ, RANK="0", WORLD_SIZE="2", LOCAL_RANK="0", PB200_REF_CHUNK="4")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                        "--steps", "1", "--warmup", "0"], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["value"] > 0
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["cpu_baseline"]["kind"] == "port"
