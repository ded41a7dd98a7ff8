mkdir -p gpurun_out
export NCCL_DEBUG=WARN
run() { # reserve maxctas pdl
NITI_B200_DP_RESERVE_SMS=$1 NCCL_MAX_CTAS=$2 NITI_B200_PDL=$3 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 --steps 100 --warmup 5 2> gpurun_out/dp.err | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith(chr(123))][0]); print('reserve $1 maxctas $2 pdl $3:', round(d['value']), round(d['ms_per_step'],4))"
}
run 16 16 513
run 16 16 769
run 16 16 513
run 16 16 769
