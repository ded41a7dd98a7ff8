#!/bin/bash
# Multi-GPU box call: data-parallel parity tests, then the bench at N ranks (weak and strong), then the N-rank timeline.
#   gpurun --gpus N --timeout 1500 -- 'bash tools/gpu_dp.sh N [tag]'
n=${1:-2}; tag=${2:-cur}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dp.py -x -q -m gpu > gpurun_out/${tag}_dp_pytest_${n}.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/${tag}_dp_pytest_${n}.log
for mode in weak strong; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus $n --scaling $mode --no-cpu-baseline \
      > gpurun_out/${tag}_bench_vgg_${n}gpu_${mode}.json 2> gpurun_out/${tag}_bench_vgg_${n}gpu_${mode}.err; echo bench $mode rc=$?
  tail -1 gpurun_out/${tag}_bench_vgg_${n}gpu_${mode}.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d.get('dp_check'), d.get('scaling'), d['config'])"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29573 tools/timeline.py vgg11 256 > gpurun_out/${tag}_timeline_vgg_${n}gpu.txt 2>&1; echo timeline rc=$?
head -45 gpurun_out/${tag}_timeline_vgg_${n}gpu.txt
