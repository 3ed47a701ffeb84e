#!/bin/bash
# r02 call A: the new fp16x2 mode -- parity tests, per-operator profile, and the re-bounded detector tests
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv,noheader > gpurun_out/a_gpu.txt
timeout 900 python -m pytest tests/test_gpu_fp16x2.py -q -s -m gpu > gpurun_out/a_fp16x2_tests.log 2>&1; echo "fp16x2 tests exit $?"; tail -15 gpurun_out/a_fp16x2_tests.log
timeout 300 python tools/profile_ops.py fp16x2 256 3 > gpurun_out/a_ops_fp16x2_rgb.json 2> gpurun_out/a_ops_fp16x2_rgb.err; echo "ops fp16x2 exit $?"; head -c 300 gpurun_out/a_ops_fp16x2_rgb.json
timeout 300 python tools/profile_ops.py fp16x2 256 4 > gpurun_out/a_ops_fp16x2_rgbd.json 2> gpurun_out/a_ops_fp16x2_rgbd.err; echo "ops fp16x2 rgbd exit $?"
timeout 300 python tools/profile_ops.py bf16 256 3 > gpurun_out/a_ops_bf16_rgb.json 2> gpurun_out/a_ops_bf16_rgb.err; echo "ops bf16 exit $?"
timeout 900 python -m pytest tests/test_gpu_detector.py -q -s -m gpu > gpurun_out/a_detector_tests.log 2>&1; echo "detector tests exit $?"; tail -8 gpurun_out/a_detector_tests.log
grep -h "^\[" gpurun_out/a_fp16x2_tests.log gpurun_out/a_detector_tests.log | head -60
