mkdir -p gpurun_out
export NCCL_DEBUG=WARN
for r in 0 8 16; do
NITI_B200_DP_RESERVE_SMS=$r timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2971$((r % 10)) bench.py --gpus 2 2> gpurun_out/dp.err | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith(chr(123))][0]); print('reserve $r', d['value'], d['ms_per_step'])"
done
NCCL_MAX_CTAS=8 NITI_B200_DP_RESERVE_SMS=8 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29719 bench.py --gpus 2 2> gpurun_out/dp.err | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith(chr(123))][0]); print('reserve 8 maxctas 8', d['value'], d['ms_per_step'])"
