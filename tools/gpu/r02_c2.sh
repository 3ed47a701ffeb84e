#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/profile_step.py 256 2 3 > gpurun_out/c2_plain.log 2>&1 || exit 0
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"conv_chain_kernel|conv_p8_pair_kernel|conv_halo_kernel" -s 7 -c 7 -o gpurun_out/c2_top python tools/profile_step.py 256 2 3 > gpurun_out/c2_ncu.log 2>&1; echo "ncu exit $?"
