#!/bin/bash
# A/B a library environment knob on the VGG-11 bench line (short runs).
#   gpurun --timeout 900 -- 'bash tools/gpu_knob.sh NITI_B200_DW_CTAS 48 74 96 148'
var=$1; shift
mkdir -p gpurun_out
for v in "$@"; do
  env $var=$v timeout 300 python bench.py --no-cpu-baseline --steps 100 > gpurun_out/knob_${var}_${v}.json 2> gpurun_out/knob_${var}_${v}.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/knob_${var}_${v}.json") if l.startswith("{")][-1])
    fam = (d.get("roofline") or {}).get("families") or {}
    print("${var}=${v}", round(d["value"]), round(d["ms_per_step"] * 1e3, 1), "us/step", {k: round(x["ms"] * 1e3, 1) for k, x in fam.items()}, d.get("exponent_prediction"))
except Exception as e:
    print("${var}=${v} unreadable", e)
PY
done
