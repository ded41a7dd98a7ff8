#!/bin/bash
# N-GPU box call: DP parity test, then the bench at N ranks — weak (default env), weak with PDL kept on under DP, strong — and the timeline.
#   gpurun --gpus 8 --timeout 1500 -- 'bash tools/gpu_dp.sh 8 [tag]'
n=${1:-8}; tag=${2:-dp}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dp.py -x -q -m gpu > gpurun_out/${tag}_dp_pytest_${n}.log 2>&1; echo pytest rc=$?; tail -2 gpurun_out/${tag}_dp_pytest_${n}.log
run() {   # name, scaling, env...
  name=$1; mode=$2; shift; shift
  env "$@" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus $n --scaling $mode --no-cpu-baseline \
      > gpurun_out/${tag}_bench_vgg_${n}gpu_${name}.json 2> gpurun_out/${tag}_bench_vgg_${n}gpu_${name}.err; echo bench $name rc=$?
  tail -1 gpurun_out/${tag}_bench_vgg_${n}gpu_${name}.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$name', round(d['value']), round(d['ms_per_step']*1e3,1), 'us/step', d.get('dp_check'), d.get('scaling'), round(d['e2e']['value']))"
}
run weak weak X=0
run weak_pdl weak NITI_B200_PDL=1537
run strong strong X=0
run strong_pdl strong NITI_B200_PDL=1537
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29573 tools/timeline.py vgg11 256 > gpurun_out/${tag}_timeline_vgg_${n}gpu.txt 2>&1; echo timeline rc=$?
grep -v "^\*\|OMP_NUM" gpurun_out/${tag}_timeline_vgg_${n}gpu.txt | head -42
