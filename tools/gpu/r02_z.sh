#!/bin/bash
# the no-allocation plasma entry point: alone in a fresh process, then after the whole GPU suite (collected last)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -x > gpurun_out/z_alone.log 2>&1; echo "alone exit $?"; tail -4 gpurun_out/z_alone.log
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/z_all.log 2>&1; echo "all gpu tests exit $?"; tail -3 gpurun_out/z_all.log
timeout 300 python tools/profile_ops.py bf16 256 3 > gpurun_out/z_ops_bf16.json 2>/dev/null; echo "ops exit $?"; python - <<PY
import json
d = json.load(open("gpurun_out/z_ops_bf16.json"))
print("step %.4f sum %.4f" % (d["ms_per_step"], d["sum_op_ms"]))
for o in d["ops"]: print("  %-70s %.4f %s %s" % (o["op"], o["ms"], o["kernel"], o["TFLOPs"]))
PY
