#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fp16x2.py tests/test_gpu_factors.py -q -m gpu > gpurun_out/i_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/i_tests.log
timeout 300 python tools/profile_ops.py fp16x2 256 3 > gpurun_out/i_ops_fp16x2.json 2> gpurun_out/i_ops.err; python - <<'PY'
import json
d = json.load(open("gpurun_out/i_ops_fp16x2.json")); fr = d["ops"]
g = lambda p: sum(o["ms"] for o in fr if o["op"].startswith(p))
print("fp16x2: step %.4f sum %.4f | stem %.4f l1 %.4f l2 %.4f l3 %.4f l4 %.4f" % (d["ms_per_step"], d["sum_op_ms"], fr[0]["ms"], g("layer1"), g("layer2"), g("layer3"), g("layer4")))
PY
timeout 300 python tools/factors_ab.py > gpurun_out/i_factors_ab.json 2> gpurun_out/i_factors_ab.err; cat gpurun_out/i_factors_ab.json; tail -2 gpurun_out/i_factors_ab.err
