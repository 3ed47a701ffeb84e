mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_detector.py -q -m gpu -x 2>&1 | tail -3
for mode in 0 1 2; do
PB200_STEM_F32_MODE=$mode timeout 300 python tools/profile_ops.py bf16x3 256 3 > gpurun_out/ops_bf16x3_m$mode.json 2> gpurun_out/ops_bf16x3_m$mode.err; echo "x3 mode $mode exit $?"
PB200_STEM_F32_MODE=$mode timeout 300 python tools/profile_ops.py bf16x3 256 4 > gpurun_out/ops_bf16x3_rgbd_m$mode.json 2> gpurun_out/ops_bf16x3_rgbd_m$mode.err; echo "x3 rgbd mode $mode exit $?"
done
python - <<'PY'
import json
for f in ("ops_bf16x3_m0","ops_bf16x3_m1","ops_bf16x3_m2","ops_bf16x3_rgbd_m0","ops_bf16x3_rgbd_m1","ops_bf16x3_rgbd_m2"):
    try:
        d=json.load(open(f"gpurun_out/{f}.json")); print(f, d["ms_per_step"], d["sum_op_ms"], [(o["op"][:22],o["ms"]) for o in d["ops"][:2]])
    except Exception as e: print(f, "ERR", e)
PY
