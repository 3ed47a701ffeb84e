#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; timeout 240 "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary_cl.txt; tail -4 gpurun_out/$name.log; }
run convcl python -m pytest tests/test_gpu_detector.py -q -m gpu -x -k "umma_conv"
run all python -m pytest tests -q -m gpu -x
run bench python bench.py --steps 20 --warmup 5
