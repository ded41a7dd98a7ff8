mkdir -p gpurun_out
export NCCL_DEBUG=WARN
for n in 8 4; do for pdl in 769 513; do
NITI_B200_PDL=$pdl timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29730 + n)) bench.py --gpus $n --steps 100 --warmup 5 2> gpurun_out/dp.err | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith(chr(123))][0]); print('N=$n pdl $pdl:', round(d['value']), round(d['ms_per_step'],4))"
done; done
