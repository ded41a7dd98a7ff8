NITI_B200_AUTO_TMA2=1 python bench.py --no-cpu-baseline > gpurun_out/bench_tma2.json 2> gpurun_out/bench_tma2.err; echo rc=$?; cat gpurun_out/bench_tma2.json | cut -c1-3000
NITI_B200_AUTO_TMA2=1 timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests2.log 2>&1; echo rc=$?; tail -4 gpurun_out/gpu_tests2.log
