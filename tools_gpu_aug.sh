#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_augment.py -q -m gpu 2>&1 | tail -150 > gpurun_out/augtests.log
cat gpurun_out/augtests.log | tail -150
