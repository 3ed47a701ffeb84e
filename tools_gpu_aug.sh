#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_augment.py -q -m gpu 2>&1 | tail -30 > gpurun_out/augtests.log
tail -3 gpurun_out/augtests.log
timeout 600 python tools/augment_bench.py > gpurun_out/aug_bench.log 2>&1; echo "bench exit $?"; tail -1 gpurun_out/aug_bench.log
timeout 600 python tools/train_step_bench.py --steps 20 --augment > gpurun_out/train_native_aug_n1.log 2>&1; echo "train+aug exit $?"; tail -2 gpurun_out/train_native_aug_n1.log
