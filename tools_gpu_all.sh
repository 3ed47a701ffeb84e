#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | tail -30 > gpurun_out/alltests.log
tail -8 gpurun_out/alltests.log
timeout 1200 python bench.py --no-training > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"; head -c 300 gpurun_out/bench.log; echo; tail -3 gpurun_out/bench.err
PB200_NO_PDL=1 timeout 1200 python bench.py --no-training > gpurun_out/bench_nopdl.log 2> gpurun_out/bench_nopdl.err; echo "bench(nopdl) exit $?"; head -c 300 gpurun_out/bench_nopdl.log; echo
