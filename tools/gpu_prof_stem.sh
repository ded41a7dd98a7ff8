mkdir -p gpurun_out
python tools/prof_op.py stem && \
ncu --set full --clock-control none --import-source on -k regex:stem_fwd -s 2 -c 1 -f -o gpurun_out/prof_stem python tools/prof_op.py stem > gpurun_out/ncu_stem.log 2>&1; echo rc=$?
ncu -i gpurun_out/prof_stem.ncu-rep --page details --csv 2>/dev/null | grep -E "Duration|Executed Ipc|Warp Cycles Per Issued|Stall|DRAM Throughput|L2 Cache Throughput|Registers|Mem Busy" | cut -c1-200 | head -30
