#!/bin/bash
# r02 call F: full GPU suite (fixed-lag chain, fused factors split launch, RESB test), bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/f_alltests.log 2>&1; echo "gpu tests exit $?"; tail -12 gpurun_out/f_alltests.log
timeout 900 python bench.py --steps 50 --warmup 5 > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; echo "bench exit $?"; tail -3 gpurun_out/f_bench.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/f_bench.json"))
print("value %.0f ms %.4f | e2e %.0f ms %.4f | h2d ceil %.0f | fp16x2 %.0f e2e %.0f | sustained %.0f @%s" % (
    d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"]["h2d_ceiling_frames_per_s"],
    d["parity_mode"]["value"], d["parity_mode"]["e2e"]["value"], d["sustained"]["value"], d["sustained"]["clocks"]["sm_mhz"]))
print({k: v for k, v in d["factors"].items() if k.startswith("chain")})
print(d["latency"])
PY
