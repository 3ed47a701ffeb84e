#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_training.py tests/test_gpu_detector.py -q -m gpu -x 2>&1 | tail -30 > gpurun_out/traintests.log
tail -6 gpurun_out/traintests.log
timeout 600 python tools/train_step_bench.py --steps 20 --profile > gpurun_out/train_native_n1.log 2>&1; echo "train exit $?"; tail -1 gpurun_out/train_native_n1.log
PB200_NO_FUSED_STATS=1 timeout 600 python tools/train_step_bench.py --steps 20 > gpurun_out/train_native_nofstats.log 2>&1; echo "train(no fused stats) exit $?"; tail -1 gpurun_out/train_native_nofstats.log | cut -c1-300
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -4 gpurun_out/smoke.log
