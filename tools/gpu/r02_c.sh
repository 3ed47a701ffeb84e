#!/bin/bash
# r02 call C: A/B of the tile cost model and the L2 chunking of stem + layer1; pipeline tests; bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pipeline.py tests/test_gpu_detector.py tests/test_gpu_fp16x2.py -q -m gpu > gpurun_out/c_tests.log 2>&1; echo "tests exit $?"; tail -4 gpurun_out/c_tests.log
run() { # name precision env...
  name=$1; prec=$2; shift 2
  env "$@" timeout 300 python tools/profile_ops.py $prec 256 3 > gpurun_out/c_ops_${name}.json 2> gpurun_out/c_ops_${name}.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/c_ops_${name}.json"))
    fr = [o for o in d["ops"]]
    l1 = sum(o["ms"] for o in fr if o["op"].startswith("layer1"))
    l2 = sum(o["ms"] for o in fr if o["op"].startswith("layer2"))
    l3 = sum(o["ms"] for o in fr if o["op"].startswith("layer3"))
    l4 = sum(o["ms"] for o in fr if o["op"].startswith("layer4"))
    print("${name}: step %.4f ms  sum_ops %.4f | stem %.4f l1 %.4f l2 %.4f l3 %.4f l4 %.4f | %s" % (d["ms_per_step"], d["sum_op_ms"], fr[0]["ms"], l1, l2, l3, l4, fr[-2]["kernel"]))
except Exception as e:
    print("${name}: FAILED", e)
PY
}
run bf16_r01 bf16 PB200_R01_TILES=1 PB200_L1_CHUNK=0
run bf16_tiles bf16 PB200_L1_CHUNK=0
run bf16_tiles_c40 bf16 PB200_L1_CHUNK=40
run bf16_tiles_c31 bf16 PB200_L1_CHUNK=31
run bf16_tiles_c64 bf16 PB200_L1_CHUNK=64
run bf16_tiles_c86 bf16 PB200_L1_CHUNK=86
run x2_c0 fp16x2 PB200_L1_CHUNK=0
run x2_c40 fp16x2 PB200_L1_CHUNK=40
run x2_c64 fp16x2 PB200_L1_CHUNK=64
timeout 900 python bench.py --steps 20 --warmup 5 --no-training > gpurun_out/c_bench.json 2> gpurun_out/c_bench.err; echo "bench exit $?"; tail -3 gpurun_out/c_bench.err
python - <<PY
import json
d = json.load(open("gpurun_out/c_bench.json"))
print("value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["value"], d["e2e"]["ms_per_step"], "h2d ceil", d["e2e"]["h2d_ceiling_frames_per_s"])
print("parity_mode", d["parity_mode"]["value"], d["parity_mode"]["e2e"]["value"], d["parity_mode"]["max_px_vs_reference"])
print("sustained", d["sustained"]["value"], d["sustained"]["clocks"])
PY
