mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:stem_f32 -s 2 -c 1 -o gpurun_out/prof_stem_f32 -f python tools/profile_ops.py bf16x3 256 3 > gpurun_out/ncu_stem_f32.log 2>&1; echo "ncu exit $?"
ncu -i gpurun_out/prof_stem_f32.ncu-rep --page raw --csv > gpurun_out/prof_stem_f32_raw.csv 2>/dev/null; wc -c gpurun_out/prof_stem_f32_raw.csv
