#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/factors_ab.py 4 > gpurun_out/g_plain.json 2>&1 || exit 0
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dynamics_kernel -s 3 -c 1 -o gpurun_out/g_dyn python tools/factors_ab.py 4 > gpurun_out/g_ncu.log 2>&1; echo "ncu exit $?"; tail -3 gpurun_out/g_ncu.log
