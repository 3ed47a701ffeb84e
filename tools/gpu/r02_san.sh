#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/sanitize_smoke.py > gpurun_out/san_plain.log 2>&1 && timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 7 --log-file gpurun_out/san_memcheck.log python tools/sanitize_smoke.py > gpurun_out/san_run.log 2>&1; echo "sanitizer exit $?"; tail -3 gpurun_out/san_plain.log; tail -12 gpurun_out/san_memcheck.log
