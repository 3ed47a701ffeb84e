#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift; echo "=== $name" ; timeout 600 "$@" > gpurun_out/$name.log 2>&1; echo "$name exit $?" | tee -a gpurun_out/summary2.txt; tail -4 gpurun_out/$name.log; }
run factors2 python -m pytest tests/test_gpu_factors.py -q -m gpu
run smoke python __graft_entry__.py --smoke
run bench python bench.py --steps 20 --warmup 5
run benchref python bench.py --impl reference --steps 3 --warmup 1
python bench.py --steps 3 --warmup 3 > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches exit $?" | tee -a gpurun_out/summary2.txt
