#!/bin/bash
mkdir -p gpurun_out
for rep in 1 2; do
for mode in graph eager; do
if [ $mode = eager ]; then export PB200_PIPE_EAGER=1; else unset PB200_PIPE_EAGER; fi
timeout 600 python bench.py --steps 50 --warmup 5 --no-training --no-compile > gpurun_out/x_bench_$mode.json 2> gpurun_out/x_bench_$mode.err
python - <<PY
import json
d = json.loads([l for l in open("gpurun_out/x_bench_$mode.json") if l.startswith("{")][-1])
print("$mode: value %.0f (%.4f ms) e2e %.0f (%.4f ms) | fp16x2 %.0f e2e %.0f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["parity_mode"]["value"], d["parity_mode"]["e2e"]["value"]))
PY
done
done
