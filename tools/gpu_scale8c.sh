mkdir -p gpurun_out
export NCCL_DEBUG=WARN
for n in 4 8 2; do
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + n)) bench.py --gpus $n --steps 200 --warmup 5 > gpurun_out/bench_vgg_${n}gpu.json 2> gpurun_out/bench_vgg_${n}gpu.err; echo bench n=$n rc=$?
python -c "
import json
d=json.loads([l for l in open('gpurun_out/bench_vgg_${n}gpu.json') if l.startswith('{')][-1]); print($n, d['value'], d['ms_per_step'], d['e2e']['value'])"
done
timeout 200 python -m pytest tests/test_gpu_dp.py -x -q 2>&1 | tail -2
