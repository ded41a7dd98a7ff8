# session re-entry baseline: parity tests + bench on all three workloads
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_s2.log 2>&1; echo tests rc=$?; tail -3 gpurun_out/gpu_tests_s2.log
python bench.py > gpurun_out/bench_s2_vgg.json 2> gpurun_out/bench_s2_vgg.err; echo bench rc=$?; cut -c1-1800 gpurun_out/bench_s2_vgg.json
python bench.py --workload resnet18 --steps 5 --no-cpu-baseline > gpurun_out/bench_s2_resnet.json 2> gpurun_out/bench_s2_resnet.err; echo bench rc=$?; cut -c1-2500 gpurun_out/bench_s2_resnet.json
python bench.py --workload lenet --no-cpu-baseline > gpurun_out/bench_s2_lenet.json 2> gpurun_out/bench_s2_lenet.err; echo bench rc=$?; cut -c1-1500 gpurun_out/bench_s2_lenet.json
