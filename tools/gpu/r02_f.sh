#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_factors.py tests/test_gpu_gtsam_boundary.py tests/test_gpu_fixed_lag.py tests/test_gpu_smoother.py -q -m gpu -x > gpurun_out/f_tests.log 2>&1; echo "factor tests exit $?"; tail -4 gpurun_out/f_tests.log
timeout 300 python tools/factors_ab.py > gpurun_out/f_ab.json 2> gpurun_out/f_ab.err; echo "ab exit $?"; cat gpurun_out/f_ab.json
