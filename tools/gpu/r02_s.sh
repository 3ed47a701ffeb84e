#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/profile_step.py 256 2 3 > gpurun_out/s_plain.log 2>&1 || exit 0
PB200_P8_GROUP=256 timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_p8_chain -s 2 -c 2 -o gpurun_out/s_chain python tools/profile_step.py 256 2 3 > gpurun_out/s_ncu1.log 2>&1; echo "ncu chain exit $?"
PB200_NO_P8CHAIN=1 timeout 600 ncu --set full --clock-control none -k regex:conv_p8_pair -s 4 -c 4 -o gpurun_out/s_pair python tools/profile_step.py 256 2 3 > gpurun_out/s_ncu2.log 2>&1; echo "ncu pair exit $?"
