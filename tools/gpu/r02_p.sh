#!/bin/bash
# layer1 CTA-pair kernels: tests (bounded: a cluster kernel that deadlocks must not hang the box), then per-op A/B
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_detector.py tests/test_gpu_fp16x2.py -q -m gpu -x -k "conv_p8" > gpurun_out/p_tests.log 2>&1; rc=$?; echo "p8 tests exit $rc"; tail -5 gpurun_out/p_tests.log
if [ $rc -ne 0 ]; then exit 0; fi
for prec in bf16 fp16x2; do
for mode in pair single; do
if [ $mode = single ]; then export PB200_P8_NOPAIR=1; else unset PB200_P8_NOPAIR; fi
timeout 300 python tools/profile_ops.py $prec 256 3 > gpurun_out/p_ops_${prec}_$mode.json 2>/dev/null
python - <<PY
import json
d = json.load(open("gpurun_out/p_ops_${prec}_$mode.json")); fr = d["ops"]
g = lambda p: sum(o["ms"] for o in fr if o["op"].startswith(p))
print("$prec $mode: step %.4f ms  sum_ops %.4f | stem %.4f l1 %.4f (%s) l2 %.4f l3 %.4f l4 %.4f" % (d["ms_per_step"], d["sum_op_ms"], fr[0]["ms"], g("layer1"), " ".join("%.4f" % o["ms"] for o in fr if o["op"].startswith("layer1")), g("layer2"), g("layer3"), g("layer4")))
PY
done
done
unset PB200_P8_NOPAIR
timeout 900 python -m pytest tests/test_gpu_detector.py tests/test_gpu_fp16x2.py -q -m gpu -x > gpurun_out/p_tests_all.log 2>&1; echo "detector tests exit $?"; tail -3 gpurun_out/p_tests_all.log
