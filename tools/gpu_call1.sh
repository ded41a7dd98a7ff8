# round-1 evidence refresh: tests, bench (both arms), ncu launch list + full-set capture of the CTA-pair kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests3.log 2>&1; echo tests rc=$?; tail -3 gpurun_out/gpu_tests3.log
python bench.py > gpurun_out/bench_r1b.json 2> gpurun_out/bench_r1b.err; echo bench rc=$?
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_r1b_ref.json 2> gpurun_out/bench_r1b_ref.err; echo ref rc=$?
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain_b.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_r1_vgg_tma2.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_b.log 2>&1; echo ncu-list rc=$?
python tools/prof_step.py vgg11 256 3 > gpurun_out/plain_p.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:igemm_tma2 -s 20 -c 8 -f -o gpurun_out/prof_tma2 \
    python tools/prof_step.py vgg11 256 3 > gpurun_out/ncu_p.log 2>&1; echo ncu-full rc=$?
cat gpurun_out/bench_r1b.json | cut -c1-2500
