#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/e2e_probe.py bf16 > gpurun_out/g_e2e_probe.json 2> gpurun_out/g_e2e_probe.err; echo "probe exit $?"; cat gpurun_out/g_e2e_probe.json; tail -3 gpurun_out/g_e2e_probe.err
PB200_NO_PDL=1 timeout 300 python tools/e2e_probe.py bf16 > gpurun_out/g_e2e_probe_nopdl.json 2> gpurun_out/g_e2e_probe_nopdl.err; echo "probe nopdl exit $?"; cat gpurun_out/g_e2e_probe_nopdl.json
