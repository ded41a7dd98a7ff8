mkdir -p gpurun_out
export NCCL_DEBUG=WARN
for n in 8 4 2; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + n)) bench.py --gpus $n --steps 200 --warmup 5 > gpurun_out/bench_vgg_${n}gpu.json 2> gpurun_out/bench_vgg_${n}gpu.err; echo bench n=$n rc=$?
python -c "
import json
d=json.loads([l for l in open('gpurun_out/bench_vgg_${n}gpu.json') if l.startswith('{')][-1]); print($n, d['value'], d['ms_per_step'], d['e2e']['value'])"
done
for cfg in "8 32" "16 16" "16 32"; do set -- $cfg
NITI_B200_DP_RESERVE_SMS=$1 NCCL_MAX_CTAS=$2 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29720 bench.py --gpus 8 --steps 100 --warmup 5 2> gpurun_out/dp.err | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith(chr(123))][0]); print('N=8 reserve $1 maxctas $2:', round(d['value']), round(d['ms_per_step'],4))"
done
