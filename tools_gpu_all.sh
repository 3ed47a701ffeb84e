#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_smoother.py -q -m gpu 2>&1 | tail -40 > gpurun_out/smoothertests.log
tail -30 gpurun_out/smoothertests.log
timeout 1200 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"; tail -c 1500 gpurun_out/bench.log; tail -5 gpurun_out/bench.err
