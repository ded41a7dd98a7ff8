mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_s3.log 2>&1; echo tests rc=$?; tail -5 gpurun_out/gpu_tests_s3.log
python bench.py --no-cpu-baseline > gpurun_out/bench_s3_vgg.json 2> gpurun_out/bench_s3_vgg.err; echo bench rc=$?; cut -c1-400 gpurun_out/bench_s3_vgg.json; tail -3 gpurun_out/bench_s3_vgg.err
