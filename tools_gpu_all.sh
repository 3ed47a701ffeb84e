#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_detector.py -q -m gpu -x 2>&1 | tail -30 > gpurun_out/dettests.log
tail -8 gpurun_out/dettests.log
timeout 1200 python bench.py --no-training > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"; head -c 400 gpurun_out/bench.log; tail -3 gpurun_out/bench.err
PB200_NO_DS_FUSION=1 timeout 1200 python bench.py --no-training > gpurun_out/bench_nofuse.log 2> gpurun_out/bench_nofuse.err; echo "bench(nofuse) exit $?"; head -c 400 gpurun_out/bench_nofuse.log
