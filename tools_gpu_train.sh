#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; timeout 900 "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary_train.txt; tail -${TAILN:-6} gpurun_out/$name.log; }
run traintests python -m pytest tests/test_gpu_training.py -q -m gpu -x
run train_native_n1 python tools/train_step_bench.py --steps 20 --profile
