#!/bin/bash
# round-1 GPU pass: split-bf16 (bf16x3) tests first, then the whole -m gpu suite, smoke, bench (both arms)
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; timeout 900 "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary_x3.txt; tail -6 gpurun_out/$name.log; }
run x3tests python -m pytest tests/test_gpu_detector.py -q -m gpu -s -k "split or bf16x3"
run alltests python -m pytest tests -q -m gpu -x
run smoke python __graft_entry__.py --smoke
run bench python bench.py --steps 100 --warmup 10
run benchref python bench.py --impl reference --steps 3 --warmup 1
