mkdir -p gpurun_out
python tools/prof_op.py stem 256 128 8 8 256 3 1 1 | tail -2
python tools/prof_op.py stem 256 64 16 16 128 3 1 1 | tail -1
python tools/prof_op.py stem 256 512 2 2 512 3 1 1 | tail -1
python tools/prof_op.py stem 256 512 1 1 512 1 1 0 | tail -1
ncu --set full --clock-control none --import-source on -k regex:igemm_fused -s 2 -c 1 -f -o gpurun_out/prof_fused python tools/prof_op.py stem 256 128 8 8 256 3 1 1 > gpurun_out/ncu_fused.log 2>&1; echo rc=$?
ncu -i gpurun_out/prof_fused.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/fused_src.csv 2>/dev/null
ncu -i gpurun_out/prof_fused.ncu-rep --page raw --csv > gpurun_out/fused_raw.csv 2>/dev/null
