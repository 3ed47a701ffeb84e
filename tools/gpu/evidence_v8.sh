#!/bin/bash
# round-end evidence (v8 = v7 forward kernels + reworked training kernels): GPU tests, smoke, both bench arms,
# ncu launch list of the bench command, per-operator training profile
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu 2>&1 | tail -5 > gpurun_out/alltests.log; tail -2 gpurun_out/alltests.log
timeout 400 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/benchref.log 2>&1; echo "benchref exit $?"
timeout 900 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"; head -c 200 gpurun_out/bench.log; echo
timeout 400 python tools/train_step_bench.py --steps 20 --profile > gpurun_out/train_profile.log 2>&1; echo "train profile exit $?"; tail -c 400 gpurun_out/train_profile.log; echo
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed --clock-control none -s 285 -c 114 --csv --log-file gpurun_out/launches.csv python bench.py --steps 100 --warmup 10 --no-training > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches exit $?"
