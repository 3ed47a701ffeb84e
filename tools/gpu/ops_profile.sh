mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_detector.py -q -m gpu -x 2>&1 | tail -3
timeout 300 python tools/profile_ops.py bf16x3 256 3 > gpurun_out/ops_bf16x3.json 2> gpurun_out/ops_bf16x3.err; echo "x3 exit $?"
PB200_STEM_F32_SCALAR=1 timeout 300 python tools/profile_ops.py bf16x3 256 3 > gpurun_out/ops_bf16x3_scalar.json 2> gpurun_out/ops_bf16x3_scalar.err; echo "x3 scalar exit $?"
timeout 300 python tools/profile_ops.py bf16x3 256 4 > gpurun_out/ops_bf16x3_rgbd.json 2> gpurun_out/ops_bf16x3_rgbd.err; echo "x3 rgbd exit $?"
python - <<'PY'
import json
for f in ("ops_bf16x3","ops_bf16x3_scalar","ops_bf16x3_rgbd"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json")); print(f, d["ms_per_step"], d["sum_op_ms"], [(o["op"][:22],o["ms"]) for o in d["ops"][:3]])
    except Exception as e: print(f, "ERR", e)
PY
