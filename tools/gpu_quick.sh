#!/bin/bash
# Short GPU-box call for kernel iteration: the parity / step tests, then the VGG-11 bench line, timeline and phase trace.
#   gpurun --timeout 1200 -- 'bash tools/gpu_quick.sh [tag] [pytest -k expression]'
tag=${1:-q}; kexpr=${2:-}
mkdir -p gpurun_out
if [ -n "$kexpr" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q -k "$kexpr" > gpurun_out/${tag}_pytest.log 2>&1; echo pytest rc=$?
else
  timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo pytest rc=$?
fi
tail -5 gpurun_out/${tag}_pytest.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/${tag}_bench_vgg.json 2> gpurun_out/${tag}_bench_vgg.err; echo bench rc=$?
NITI_B200_SPECULATE=0 timeout 300 python bench.py --no-cpu-baseline --steps 100 > gpurun_out/${tag}_bench_vgg_nospec.json 2> gpurun_out/${tag}_bench_vgg_nospec.err; echo bench nospec rc=$?
python tools/timeline.py vgg11 256 > gpurun_out/${tag}_timeline_vgg.txt 2>&1; echo timeline rc=$?
python tools/phase_trace.py vgg11 256 > gpurun_out/${tag}_phase_vgg.txt 2>&1; echo phase rc=$?
for f in vgg vgg_nospec; do
python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/${tag}_bench_${f}.json") if l.startswith("{")][-1])
    print("${f}", d["value"], d["ms_per_step"], d.get("gpu_launches"), d.get("e2e", {}).get("value"), (d.get("roofline") or {}).get("frac"))
    for k, v in ((d.get("roofline") or {}).get("families") or {}).items(): print("   ", k, v)
except Exception as e:
    print("bench ${f} unreadable", e)
PY
done
head -40 gpurun_out/${tag}_timeline_vgg.txt
head -30 gpurun_out/${tag}_phase_vgg.txt
