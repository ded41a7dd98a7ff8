#!/bin/bash
# One GPU-box call that collects the round's evidence: parity tests, smoke, bench lines (VGG-11 / LeNet / ResNet-18),
# step timeline + phase trace, config-5 shape sweep, then the ncu launch list and ONE ncu --set full pass over the
# step's hot kernels (run only after the same command exited 0 without ncu).
#   gpurun --timeout 2400 -- 'bash tools/gpu_evidence.sh [tag] [skip-tests]'
tag=${1:-cur}
mkdir -p gpurun_out
if [ "$2" != "skip-tests" ]; then
  timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/${tag}_pytest.log
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo smoke rc=$?; tail -1 gpurun_out/${tag}_smoke.log
fi
timeout 500 python bench.py > gpurun_out/${tag}_bench_vgg.json 2> gpurun_out/${tag}_bench_vgg.err; echo bench vgg rc=$?
timeout 300 python bench.py --workload lenet --no-cpu-baseline > gpurun_out/${tag}_bench_lenet.json 2> gpurun_out/${tag}_bench_lenet.err; echo bench lenet rc=$?
timeout 400 python bench.py --workload resnet18 --no-cpu-baseline --steps 30 > gpurun_out/${tag}_bench_resnet18.json 2> gpurun_out/${tag}_bench_resnet18.err; echo bench resnet rc=$?
timeout 400 python bench.py --workload lenet --via mnn > gpurun_out/${tag}_bench_lenet_via_mnn.json 2> gpurun_out/${tag}_bench_lenet_via_mnn.err; echo bench lenet via mnn rc=$?
python tools/phase_trace.py vgg11 256 > gpurun_out/${tag}_phase_vgg.txt 2>&1; echo phase rc=$?
python tools/timeline.py vgg11 256 > gpurun_out/${tag}_timeline_vgg.txt 2>&1; echo timeline rc=$?
timeout 600 python tools/sweep.py --iters 10 --out gpurun_out/${tag}_sweep.json > gpurun_out/${tag}_sweep.log 2>&1; echo sweep rc=$?
for f in vgg lenet resnet18; do
python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/${tag}_bench_${f}.json") if l.startswith("{")][-1])
    print("${f}", d["value"], d["ms_per_step"], d.get("gpu_launches"), d.get("e2e", {}).get("value"), (d.get("roofline") or {}).get("frac"), d.get("cpu_baseline"))
    for k, v in ((d.get("roofline") or {}).get("families") or {}).items(): print("   ", k, v)
except Exception as e:
    print("bench ${f} unreadable", e)
PY
done
# ncu: launch list of 3 eager steps, then the full set over the second step's hot kernels
python tools/prof_step.py vgg11 256 3 > gpurun_out/${tag}_prof_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${tag}_launches.csv python tools/prof_step.py vgg11 256 3 > gpurun_out/${tag}_ncu_list.log 2>&1
echo launch list rc=$?
python tools/prof_step.py vgg11 256 2 > gpurun_out/${tag}_prof_plain2.log 2>&1 &&
ncu --set full --clock-control none -k 'regex:chain_kernel|dw2_kernel|stem2|update_batched' -s 26 -c 26 -o gpurun_out/${tag}_prof \
    python tools/prof_step.py vgg11 256 2 > gpurun_out/${tag}_ncu_full.log 2>&1
echo ncu full rc=$?
ncu -i gpurun_out/${tag}_prof.ncu-rep --page raw --csv > gpurun_out/${tag}_prof_raw.csv 2>/dev/null; echo export rc=$?
ls -la gpurun_out | head -40
