#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/t_all.log 2>&1; echo "all gpu tests exit $?"; tail -3 gpurun_out/t_all.log
for prec in bf16 fp16x2; do
timeout 300 python tools/profile_ops.py $prec 256 3 > gpurun_out/t_ops_$prec.json 2>/dev/null
python - <<PY
import json
d = json.load(open("gpurun_out/t_ops_$prec.json")); fr = d["ops"]
g = lambda p: sum(o["ms"] for o in fr if o["op"].startswith(p))
print("$prec: step %.4f ms  sum_ops %.4f | stem %.4f l1 %.4f l2 %.4f l3+4 %.4f" % (d["ms_per_step"], d["sum_op_ms"], fr[0]["ms"], g("layer1"), g("layer2"), g("layer3")))
PY
done
