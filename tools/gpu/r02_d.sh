#!/bin/bash
# r02 call D: pipelined ingest A/B in bench, pipeline tests, ncu --set full of the front kernels (bf16)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_pipeline.py tests/test_gpu_detector.py -q -m gpu > gpurun_out/d_tests.log 2>&1; echo "tests exit $?"; tail -3 gpurun_out/d_tests.log
summ() { python - "$1" <<'PY'
import json, sys
d = json.load(open(sys.argv[1]))
print(sys.argv[1], "value %.0f ms %.4f | e2e %.0f ms %.4f | h2d ceil %.0f | fp16x2 %.0f e2e %.0f | sustained %.0f @%s" % (
    d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"]["h2d_ceiling_frames_per_s"],
    d["parity_mode"]["value"], d["parity_mode"]["e2e"]["value"], d["sustained"]["value"], d["sustained"]["clocks"]["sm_mhz"]))
PY
}
timeout 900 python bench.py --steps 50 --warmup 5 --no-training --no-compile > gpurun_out/d_bench_pipe.json 2> gpurun_out/d_bench_pipe.err; echo "bench exit $?"; tail -3 gpurun_out/d_bench_pipe.err; summ gpurun_out/d_bench_pipe.json
PB200_BENCH_SERIAL=1 timeout 900 python bench.py --steps 50 --warmup 5 --no-training --no-compile > gpurun_out/d_bench_serial.json 2> gpurun_out/d_bench_serial.err; echo "bench serial exit $?"; summ gpurun_out/d_bench_serial.json
PB200_COPY_STREAMS=2 timeout 900 python bench.py --steps 50 --warmup 5 --no-training --no-compile > gpurun_out/d_bench_copy2.json 2> gpurun_out/d_bench_copy2.err; echo "bench copy2 exit $?"; summ gpurun_out/d_bench_copy2.json
timeout 300 python tools/profile_step.py 256 3 > gpurun_out/d_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"stem_pool|conv_p8|conv_umma2" -s 19 -c 9 -o gpurun_out/prof_r02_front python tools/profile_step.py 256 3 > gpurun_out/d_ncu.log 2>&1; echo "ncu exit $?"; tail -3 gpurun_out/d_ncu.log
ls -la gpurun_out/*.ncu-rep
