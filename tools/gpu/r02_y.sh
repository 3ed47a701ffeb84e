#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_detector.py tests/test_gpu_fp16x2.py tests/test_gpu_pipeline.py -q -m gpu -x > gpurun_out/y_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/y_tests.log
for prec in bf16 fp16x2; do
timeout 300 python tools/profile_ops.py $prec 256 3 > gpurun_out/y_ops_$prec.json 2>/dev/null
python - <<PY
import json
d = json.load(open("gpurun_out/y_ops_$prec.json")); fr = d["ops"]
print("$prec: step %.4f ms sum %.4f | head %.4f" % (d["ms_per_step"], d["sum_op_ms"], fr[-1]["ms"]))
PY
done
timeout 200 python tools/latency_bench.py 2>/dev/null | tail -3
