#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/profile_step.py 256 2 3 > gpurun_out/st_plain.log 2>&1 || exit 0
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stem_pool_kernel16 -s 1 -c 1 -o gpurun_out/st_stem python tools/profile_step.py 256 2 3 > gpurun_out/st_ncu.log 2>&1; echo "ncu exit $?"
