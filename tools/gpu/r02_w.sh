#!/bin/bash
# layer1 tile walk direction alternating launch by launch (L2 reuse across launches): tests, then A/B
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_detector.py tests/test_gpu_fp16x2.py tests/test_gpu_fused_ingest.py -q -m gpu -x -k "conv_p8 or batch or forward or fused or stem" > gpurun_out/w_tests.log 2>&1; rc=$?; echo "tests exit $rc"; tail -3 gpurun_out/w_tests.log
if [ $rc -ne 0 ]; then exit 0; fi
for rep in 1 2; do
for prec in bf16 fp16x2; do
for mode in alt fwd; do
if [ $mode = fwd ]; then export PB200_P8_NOREVERSE=1; else unset PB200_P8_NOREVERSE; fi
timeout 300 python tools/profile_ops.py $prec 256 3 > gpurun_out/w_ops_${prec}_$mode.json 2>/dev/null
python - <<PY
import json
d = json.load(open("gpurun_out/w_ops_${prec}_$mode.json")); fr = d["ops"]
g = lambda p: sum(o["ms"] for o in fr if o["op"].startswith(p))
print("$prec $mode: step %.4f ms sum %.4f | stem %.4f l1 %.4f (%s) l2 %.4f (%s) l3+4 %.4f" % (d["ms_per_step"], d["sum_op_ms"], fr[0]["ms"], g("layer1"), " ".join("%.4f" % o["ms"] for o in fr if o["op"].startswith("layer1")), g("layer2"), " ".join("%.4f" % o["ms"] for o in fr if o["op"].startswith("layer2")), g("layer3")))
PY
done
done
done
