#!/bin/bash
mkdir -p gpurun_out
for rep in 1 2; do
for v in old new; do
if [ $v = old ]; then d=_wt_old; else d=.; fi
(cd $d && PB200_NO_P8CHAIN=1 timeout 300 python tools/profile_ops.py bf16 256 3 2>/dev/null) > gpurun_out/u_$v.json
python - <<PY
import json
d = json.load(open("gpurun_out/u_$v.json")); fr = d["ops"]
print("$v: step %.4f | l1 %s | l2 %s" % (d["ms_per_step"], " ".join("%.4f" % o["ms"] for o in fr if o["op"].startswith("layer1")), " ".join("%.4f" % o["ms"] for o in fr if o["op"].startswith("layer2"))))
PY
done
done
