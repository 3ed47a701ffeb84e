#!/bin/bash
# bring-up of the P8 kernels: unit tests first (each in its own process with a timeout), then everything
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; timeout 240 "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary_p8.txt; tail -6 gpurun_out/$name.log; }
run convp8 python -m pytest tests/test_gpu_detector.py -q -m gpu -x -k "conv_p8"
run stempool python -m pytest tests/test_gpu_detector.py -q -m gpu -x -k "stem_pool"
run all python -m pytest tests -q -m gpu -x
run bench python bench.py --steps 20 --warmup 5
