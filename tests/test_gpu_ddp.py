"""-m gpu: what DistributedDataParallel guarantees for the reference's training step (train.py:130,152,300), checked on
the native trainer with TWO ranks: identical parameters on every rank after each step, reduced gradient = sum of the
per-rank gradients (1/world is folded into the optimizer), BatchNorm running statistics broadcast from rank 0 before
every forward.  Both ranks share cuda:0 (the GPU test tier has one device; NCCL refuses two ranks on one device), so the
process group is gloo over CUDA tensors: the host-side logic under test -- bucket boundaries, stream ordering of the
overlapped all-reduce around the two CUDA graphs, buffer broadcast -- is the same code path NCCL runs at N = 2/4/8
(bench.py's training_step section repeats the checks there)."""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, overlap, out):
    import torch.distributed as dist

    from perseus_b200.detector import KeypointCNN
    from perseus_b200.detector.training import NativeTrainer

    torch.cuda.set_device(0)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        torch.manual_seed(100 + rank)  # DIFFERENT initial weights and BN buffers per rank: the broadcast must fix that
        model = KeypointCNN(num_channels=4).cuda()
        with torch.no_grad():
            for b in model.buffers():
                if b.dtype.is_floating_point:
                    b.add_(0.01 * rank)
        B = 4
        tr = NativeTrainer(model, B, overlap_allreduce=overlap)
        g = torch.Generator(device="cuda").manual_seed(7 + rank)  # different data per rank (DistributedSampler)
        res = {}
        for step in range(5):  # steps 0-1 eager, step 2 captures the graph(s), 3-4 replay
            x = torch.rand(B, 4, 256, 256, device="cuda", generator=g)
            x[:, 3] = 3 + 11 * x[:, 3]
            y = torch.rand(B, 8, 2, device="cuda", generator=g) * 2 - 1
            chk = tr.verify_next_step(1 << 18) if step in (0, 2, 4) else None
            loss = tr.step(x, y)
            assert torch.isfinite(loss).all()
            if chk is not None:
                res[step] = dict(chk)
        res["graphs"] = (tr._graph is not None, tr._graph2 is not None)
        res["tail_off"], res["n_params"] = tr.tail_off, tr.n_params
        if rank == 0:
            torch.save(res, out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("overlap", [True, False])
def test_two_rank_step_keeps_replicas_identical(tmp_path, overlap):
    import torch.multiprocessing as mp

    out = str(tmp_path / "res.pt")
    mp.spawn(_worker, args=(2, _free_port(), overlap, out), nprocs=2, join=True)
    res = torch.load(out)
    assert res["graphs"] == (True, overlap), "overlap = two CUDA graphs with the [layer4, fc] all-reduce between them"
    assert 0 < res["tail_off"] < res["n_params"]
    for step in (0, 2, 4):
        chk = res[step]
        assert chk["world"] == 2
        assert chk["params_identical"], f"step {step}: ranks diverged"
        assert chk["bn_buffers_follow_rank0"], f"step {step}: BN buffers not broadcast"
        assert chk["grad_allreduce_max_rel_err"] <= 1e-5, (step, chk)
