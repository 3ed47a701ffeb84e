#!/bin/bash
# round-2 evidence: full GPU suite, smoke, both bench arms, per-op profiles of every mode, ncu launch list of the bench
# command (traffic.json is regenerated from it), ncu --set full of the fp16x2 kernels
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/ev_alltests.log 2>&1; echo "gpu tests exit $?"; tail -3 gpurun_out/ev_alltests.log
timeout 400 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/ev_smoke.log 2>&1; echo "smoke exit $?"; tail -8 gpurun_out/ev_smoke.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/ev_benchref.json 2> gpurun_out/ev_benchref.err; echo "benchref exit $?"
timeout 900 python bench.py > gpurun_out/ev_bench.json 2> gpurun_out/ev_bench.err; echo "bench exit $?"; head -c 300 gpurun_out/ev_bench.json; echo
for m in "bf16 3" "fp16x2 3" "fp16x2 4" "bf16 4" "bf16x3 3"; do set -- $m; timeout 300 python tools/profile_ops.py $1 256 $2 > gpurun_out/ev_ops_$1_c$2.json 2>/dev/null; echo "ops $1 c$2 exit $?"; done
timeout 300 python bench.py --steps 20 --warmup 5 --no-training --no-compile > gpurun_out/ev_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed --clock-control none -s 40 -c 160 --csv --log-file gpurun_out/ev_launches.csv python bench.py --steps 20 --warmup 5 --no-training --no-compile > gpurun_out/ev_ncu_launch.log 2>&1; echo "ncu launches exit $?"
