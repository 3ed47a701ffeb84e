#!/bin/bash
# round-end evidence: full GPU test suite, smoke, bench (both arms), ncu launch list + full capture of the top kernels
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu 2>&1 | tail -5 > gpurun_out/alltests.log; tail -2 gpurun_out/alltests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/benchref.log 2>&1; echo "benchref exit $?"
timeout 1200 python bench.py > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench exit $?"; head -c 200 gpurun_out/bench.log; echo
timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed --clock-control none -s 285 -c 114 --csv --log-file gpurun_out/launches.csv python bench.py --steps 100 --warmup 10 --no-training > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"conv_umma|conv_p8|stem_pool" -s 19 -c 18 -o gpurun_out/prof_v6 -f python tools/profile_step.py 256 3 > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"
