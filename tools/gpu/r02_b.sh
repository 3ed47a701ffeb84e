#!/bin/bash
# r02 call B: whole GPU suite, smoke, bench N=1 (both arms), torch.compile probe timing
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/b_alltests.log 2>&1; echo "gpu tests exit $?"; tail -6 gpurun_out/b_alltests.log
timeout 400 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/b_smoke.log 2>&1; echo "smoke exit $?"; tail -8 gpurun_out/b_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/b_bench.json 2> gpurun_out/b_bench.err; echo "bench exit $?"; head -c 300 gpurun_out/b_bench.json; echo; tail -5 gpurun_out/b_bench.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/b_benchref.json 2>&1; echo "benchref exit $?"
