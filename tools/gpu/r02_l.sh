#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_detector.py tests/test_gpu_fp16x2.py tests/test_gpu_training.py tests/test_gpu_pipeline.py -q -m gpu > gpurun_out/l_tests.log 2>&1; echo "tests exit $?"; tail -12 gpurun_out/l_tests.log
run() { name=$1; prec=$2; shift 2
  env "$@" timeout 300 python tools/profile_ops.py $prec 256 3 > gpurun_out/l_ops_${name}.json 2> gpurun_out/l_ops_${name}.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/l_ops_${name}.json")); fr = d["ops"]
    g = lambda p: sum(o["ms"] for o in fr if o["op"].startswith(p))
    print("${name}: step %.4f ms  sum_ops %.4f | stem %.4f l1 %.4f l2 %.4f l3 %.4f l4 %.4f" % (d["ms_per_step"], d["sum_op_ms"], fr[0]["ms"], g("layer1"), g("layer2"), g("layer3"), g("layer4")))
except Exception as e:
    print("${name}: FAILED", e)
PY
}
run bf16_halo bf16 A=1
run bf16_nohalo bf16 PB200_NO_HALO=1
run x2 fp16x2 A=1
