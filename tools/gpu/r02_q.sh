#!/bin/bash
# chained layers 3-4: correctness (bounded: the kernel traps after ~1 s of spinning, the test has its own timeout), then A/B
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_detector.py -q -m gpu -x -k "chained" -s > gpurun_out/q_tests.log 2>&1; rc=$?; echo "chain tests exit $rc"; grep "^\[chain" gpurun_out/q_tests.log; tail -5 gpurun_out/q_tests.log
if [ $rc -ne 0 ]; then exit 0; fi
for prec in bf16 fp16x2; do
for mode in chain plain; do
if [ $mode = plain ]; then export PB200_NO_CHAIN=1; else unset PB200_NO_CHAIN; fi
timeout 300 python tools/profile_ops.py $prec 256 3 > gpurun_out/q_ops_${prec}_$mode.json 2>/dev/null
python - <<PY
import json
d = json.load(open("gpurun_out/q_ops_${prec}_$mode.json")); fr = d["ops"]
g = lambda p: sum(o["ms"] for o in fr if o["op"].startswith(p))
print("$prec $mode: step %.4f ms  sum_ops %.4f | stem %.4f l1 %.4f l2 %.4f l3 %.4f l4 %.4f" % (d["ms_per_step"], d["sum_op_ms"], fr[0]["ms"], g("layer1"), g("layer2"), g("layer3"), g("layer4")))
PY
done
done
unset PB200_NO_CHAIN
timeout 1200 python -m pytest tests -q -m gpu -x > gpurun_out/q_tests_all.log 2>&1; echo "all gpu tests exit $?"; tail -3 gpurun_out/q_tests_all.log
