#!/bin/bash
# One GPU-box call: parity tests, smoke, the bench line of each arm, the phase trace and the ncu launch list.
#   gpurun --timeout 1500 -- 'bash tools/gpu_check.sh [tag]'
# Everything lands in gpurun_out/<tag>_*; copy what should be judged into profiles/.
tag=${1:-cur}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/${tag}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo smoke rc=$?; tail -1 gpurun_out/${tag}_smoke.log
timeout 400 python bench.py --no-cpu-baseline > gpurun_out/${tag}_bench_vgg.json 2> gpurun_out/${tag}_bench_vgg.err; echo bench rc=$?
python tools/phase_trace.py vgg11 256 > gpurun_out/${tag}_phase_vgg.txt 2>&1; echo phase rc=$?
python tools/timeline.py vgg11 256 > gpurun_out/${tag}_timeline_vgg.txt 2>&1; echo timeline rc=$?
python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/${tag}_bench_vgg.json") if l.startswith("{")][-1])
    print("vgg", d["value"], d["ms_per_step"], d.get("gpu_launches"), d.get("e2e", {}).get("value"), (d.get("roofline") or {}).get("frac"))
    for k, v in ((d.get("roofline") or {}).get("families") or {}).items(): print("   ", k, v)
except Exception as e:
    print("bench unreadable", e)
PY
head -60 gpurun_out/${tag}_timeline_vgg.txt
