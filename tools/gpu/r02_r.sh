#!/bin/bash
# chained layer1: correctness (the kernel traps after ~1 s of spinning; the test run is bounded), then A/B over group sizes
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_detector.py -q -m gpu -x -k "chained_layer1" > gpurun_out/r_tests.log 2>&1; rc=$?; echo "p8 chain tests exit $rc"; tail -6 gpurun_out/r_tests.log
if [ $rc -ne 0 ]; then exit 0; fi
run() {
timeout 300 python tools/profile_ops.py bf16 256 3 > gpurun_out/r_ops_$1.json 2>/dev/null
python - <<PY
import json
d = json.load(open("gpurun_out/r_ops_$1.json")); fr = d["ops"]
g = lambda p: sum(o["ms"] for o in fr if o["op"].startswith(p))
print("$1: step %.4f ms  sum_ops %.4f | stem %.4f l1 %.4f l2 %.4f l3+4 %.4f" % (d["ms_per_step"], d["sum_op_ms"], fr[0]["ms"], g("layer1"), g("layer2"), g("layer3")))
PY
}
PB200_NO_P8CHAIN=1 run plain
for G in 16 32 64 256; do PB200_P8_GROUP=$G run group$G; done
timeout 1200 python -m pytest tests/test_gpu_detector.py tests/test_gpu_fp16x2.py tests/test_gpu_pipeline.py -q -m gpu -x > gpurun_out/r_tests_all.log 2>&1; echo "detector tests exit $?"; tail -3 gpurun_out/r_tests_all.log
