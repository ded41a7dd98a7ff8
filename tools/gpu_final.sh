# final round-1 check: parity tests, smoke, bench (3 workloads + reference arm), launch list, ncu full of the dominant kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo smoke rc=$?; tail -1 gpurun_out/smoke.log
timeout 400 python bench.py > gpurun_out/bench_vgg.json 2> gpurun_out/bench_vgg.err; echo bench vgg rc=$?
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo bench ref rc=$?
for wl in resnet18 lenet; do
timeout 400 python bench.py --workload $wl --steps 50 --no-cpu-baseline > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo bench $wl rc=$?
done
python - <<'PY'
import json
for n in ("vgg", "resnet18", "lenet", "ref"):
    try:
        d = json.loads([l for l in open(f"gpurun_out/bench_{n}.json") if l.startswith("{")][-1])
        print(n, d["value"], d["ms_per_step"], d.get("gpu_launches"), d.get("e2e", {}).get("value"), (d.get("roofline") or {}).get("achieved"), (d.get("roofline") or {}).get("frac"), d.get("clocks"))
    except Exception as e:
        print(n, "unreadable", e)
PY
python tools/prof_step.py vgg11 256 3 > gpurun_out/plain_p.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_cur.csv python tools/prof_step.py vgg11 256 4 > gpurun_out/ncu_l.log 2>&1; echo launches rc=$?
python tools/launch_summary.py gpurun_out/launches_cur.csv | head -12
timeout 600 ncu --set full --clock-control none --import-source on -k regex:igemm_fused -s 12 -c 3 -f -o gpurun_out/prof_fused python tools/prof_step.py vgg11 256 2 > gpurun_out/ncu_fused.log 2>&1; echo ncu fused rc=$?
