#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_detector.py tests/test_gpu_training.py -q -m gpu -x 2>&1 | tail -30 > gpurun_out/dettests.log
tail -4 gpurun_out/dettests.log
timeout 1200 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"; head -c 300 gpurun_out/bench.log; echo; tail -3 gpurun_out/bench.err
