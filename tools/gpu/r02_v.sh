#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/factors_profile.py > gpurun_out/v_plain.log 2>&1 || exit 0
cat gpurun_out/v_plain.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:projection_kernel -s 2 -c 1 -o gpurun_out/v_proj python tools/factors_profile.py > gpurun_out/v_ncu.log 2>&1; echo "ncu exit $?"
