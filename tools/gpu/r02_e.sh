#!/bin/bash
# r02 call E: full GPU suite with resident-weights layer2 + fused chain factors, A/B of resB, bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/e_alltests.log 2>&1; echo "gpu tests exit $?"; tail -4 gpurun_out/e_alltests.log
run() { name=$1; prec=$2; shift 2
  env "$@" timeout 300 python tools/profile_ops.py $prec 256 3 > gpurun_out/e_ops_${name}.json 2> gpurun_out/e_ops_${name}.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/e_ops_${name}.json"))
    fr = d["ops"]
    g = lambda p: sum(o["ms"] for o in fr if o["op"].startswith(p))
    print("${name}: step %.4f ms  sum_ops %.4f | stem %.4f l1 %.4f l2 %.4f l3 %.4f l4 %.4f | %s" % (d["ms_per_step"], d["sum_op_ms"], fr[0]["ms"], g("layer1"), g("layer2"), g("layer3"), g("layer4"), [o["kernel"] for o in fr if o["op"].startswith("layer2")]))
except Exception as e:
    print("${name}: FAILED", e)
PY
}
run bf16_resb bf16 A=1
run bf16_noresb bf16 PB200_NO_RESB=1
timeout 900 python bench.py --steps 50 --warmup 5 > gpurun_out/e_bench.json 2> gpurun_out/e_bench.err; echo "bench exit $?"; tail -3 gpurun_out/e_bench.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/e_bench.json"))
print("value %.0f ms %.4f | e2e %.0f ms %.4f | h2d ceil %.0f | fp16x2 %.0f e2e %.0f | sustained %.0f @%s" % (
    d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"]["h2d_ceiling_frames_per_s"],
    d["parity_mode"]["value"], d["parity_mode"]["e2e"]["value"], d["sustained"]["value"], d["sustained"]["clocks"]["sm_mhz"]))
print({k: v for k, v in d["factors"].items() if k.startswith("chain") or k.startswith("poses")})
print(d["training_step"]["step"], d["latency"])
PY
