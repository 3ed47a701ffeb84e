"""Host -> device ceiling of the e2e step at N ranks (torchrun): every rank streams 50.3 MB pinned batches with one
cudaMemcpyAsync per batch, no compute.  Variants: torch pin_memory (cudaHostRegister'd pageable pages), cudaHostAlloc
default, cudaHostAlloc write-combined, and two copy streams.  Prints one JSON line from rank 0."""
import ctypes
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
NB = 256 * 256 * 256 * 3
cudart = ctypes.CDLL("libcudart.so.12")


def host_alloc(flags):
    p = ctypes.c_void_p()
    rc = cudart.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(NB), ctypes.c_uint(flags))
    assert rc == 0, rc
    buf = (ctypes.c_uint8 * NB).from_address(p.value)
    t = torch.frombuffer(buf, dtype=torch.uint8)
    t.fill_(7)
    return t


def bench(hosts, streams, iters=40):
    dst = [torch.empty(NB, dtype=torch.uint8, device=dev) for _ in range(2)]
    ss = [torch.cuda.Stream(device=dev) for _ in range(streams)]
    part = NB // streams

    def run(n):
        for i in range(n):
            for c, s in enumerate(ss):
                with torch.cuda.stream(s):
                    dst[i % 2][c * part:(c + 1) * part].copy_(hosts[i % len(hosts)][c * part:(c + 1) * part], non_blocking=True)

    run(3)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(ss[0]):
        e0.record()
    run(iters)
    for s in ss[1:]:
        ss[0].wait_stream(s)
    with torch.cuda.stream(ss[0]):
        e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return NB / (float(t.item()) * 1e-3) / 1e9


res = {"ranks": world}
torch_pinned = [torch.full((NB,), 7, dtype=torch.uint8).pin_memory() for _ in range(3)]
res["torch_pin_memory_GBps_per_gpu"] = bench(torch_pinned, 1)
res["torch_pin_memory_2streams_GBps_per_gpu"] = bench(torch_pinned, 2)
del torch_pinned
ha = [host_alloc(0) for _ in range(3)]
res["cudaHostAlloc_default_GBps_per_gpu"] = bench(ha, 1)
wc = [host_alloc(4) for _ in range(3)]  # cudaHostAllocWriteCombined
res["cudaHostAlloc_writecombined_GBps_per_gpu"] = bench(wc, 1)
res["is_pinned"] = [bool(ha[0].is_pinned()), bool(wc[0].is_pinned())]
if rank == 0:
    res["aggregate_GBps_best"] = world * max(v for k, v in res.items() if k.endswith("per_gpu"))
    print(json.dumps(res))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
