#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_factors.py tests/test_gpu_smoother.py -q -m gpu 2>&1 | tail -5 > gpurun_out/factests.log; tail -2 gpurun_out/factests.log
timeout 600 python tools/factors_profile.py > gpurun_out/factors_plain.log 2>&1; echo "factors exit $?"; cat gpurun_out/factors_plain.log
timeout 600 python tools/factors_profile.py 10000 >> gpurun_out/factors_plain.log 2>&1; tail -3 gpurun_out/factors_plain.log
