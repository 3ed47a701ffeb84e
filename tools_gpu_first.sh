#!/bin/bash
# first GPU bring-up: each stage in its own process with a timeout, logs under gpurun_out/
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
run() { name=$1; shift; echo "=== $name" ; timeout 240 "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary.txt; tail -5 gpurun_out/$name.log; }
run factors python -m pytest tests/test_gpu_factors.py -q -m gpu
run ingest python -m pytest tests/test_gpu_detector.py -q -m gpu -k "ingest or decode"
run fp32 python -m pytest tests/test_gpu_detector.py -q -m gpu -k "fp32_mode"
run conv python -m pytest tests/test_gpu_detector.py -q -m gpu -k "umma_conv"
run bf16 python -m pytest tests/test_gpu_detector.py -q -m gpu -s -k "bf16_mode"
run rest python -m pytest tests/test_gpu_detector.py -q -m gpu -k "not ingest and not decode and not fp32_mode and not umma_conv and not bf16_mode"
