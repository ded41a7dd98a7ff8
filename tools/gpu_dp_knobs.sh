#!/bin/bash
# A/B of data-parallel launch knobs on the VGG-11 bench line at N ranks (short runs, dp_check included).
#   gpurun --gpus 2 --timeout 900 -- 'bash tools/gpu_dp_knobs.sh 2'
n=${1:-2}
mkdir -p gpurun_out
run() {   # tag, env assignments...
  tag=$1; shift
  env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29581 bench.py --gpus $n --steps 100 --no-cpu-baseline \
      > gpurun_out/dpk_${n}_${tag}.json 2> gpurun_out/dpk_${n}_${tag}.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/dpk_${n}_${tag}.json") if l.startswith("{")][-1])
    print("${tag}", "$*", round(d["value"]), round(d["ms_per_step"] * 1e3, 1), "us/step", d.get("dp_check"))
except Exception as e:
    print("${tag} unreadable", e)
PY
}
run default X=0
run pdl NITI_B200_PDL=769
run reserve8 NITI_B200_DP_RESERVE_SMS=8
run reserve0_small NITI_B200_DP_RESERVE_SMS=0 NITI_B200_DP_AR_CTAS=148 NITI_B200_DP_AR_THREADS=128
run reserve0_small_pdl NITI_B200_DP_RESERVE_SMS=0 NITI_B200_DP_AR_CTAS=148 NITI_B200_DP_AR_THREADS=128 NITI_B200_PDL=769
run reserve8_pdl NITI_B200_DP_RESERVE_SMS=8 NITI_B200_PDL=769
