#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_detector.py tests/test_gpu_fp16x2.py -q -m gpu -k "stem or forward or batch" > gpurun_out/m_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/m_tests.log
for prec in bf16 fp16x2; do
timeout 300 python tools/profile_ops.py $prec 256 3 > gpurun_out/m_ops_$prec.json 2>/dev/null
python - <<PY
import json
d = json.load(open("gpurun_out/m_ops_$prec.json")); fr = d["ops"]
g = lambda p: sum(o["ms"] for o in fr if o["op"].startswith(p))
print("$prec: step %.4f ms  sum_ops %.4f | stem %.4f l1 %.4f l2 %.4f l3 %.4f l4 %.4f" % (d["ms_per_step"], d["sum_op_ms"], fr[0]["ms"], g("layer1"), g("layer2"), g("layer3"), g("layer4")))
PY
done
