#!/bin/bash
# N-GPU bench exactly as the driver launches it
N=${1:-8}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 100 --warmup 10 > gpurun_out/bench_n$N.log 2> gpurun_out/bench_n$N.err
echo "bench n$N exit $?"; tail -c 1200 gpurun_out/bench_n$N.log; tail -3 gpurun_out/bench_n$N.err
NCCL_DEBUG=INFO timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 tools/train_step_bench.py --steps 20 > gpurun_out/train_native_n$N.log 2>&1
echo "train n$N exit $?"; grep -E "NVLS|via P2P|NET/|Channel 00.*via" gpurun_out/train_native_n$N.log | head -5; tail -1 gpurun_out/train_native_n$N.log | cut -c1-400
