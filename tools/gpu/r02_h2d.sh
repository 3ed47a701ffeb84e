#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/h2d_probe.py > gpurun_out/h2d_n${N}.json 2> gpurun_out/h2d_n${N}.err; echo "exit $?"; cat gpurun_out/h2d_n${N}.json; tail -3 gpurun_out/h2d_n${N}.err
lscpu | head -20 > gpurun_out/h2d_lscpu.txt; free -g >> gpurun_out/h2d_lscpu.txt
