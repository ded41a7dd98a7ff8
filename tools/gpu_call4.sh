mkdir -p gpurun_out
for wl in resnet18 lenet; do
python bench.py --workload $wl --steps 5 --no-cpu-baseline > gpurun_out/bench_s4_$wl.json 2> gpurun_out/bench_s4_$wl.err; echo bench $wl rc=$?
python - <<PY
import json
d=json.load(open('gpurun_out/bench_s4_$wl.json'))
print(d['value'], d['ms_per_step'], d['gpu_launches'], d['roofline']['achieved'])
for k,v in d['roofline']['families'].items(): print(' ', k, v)
PY
done
