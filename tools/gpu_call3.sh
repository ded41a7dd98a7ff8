mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests_s3.log 2>&1; echo tests rc=$?; tail -8 gpurun_out/gpu_tests_s3.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_s3_vgg.json 2> gpurun_out/bench_s3_vgg.err; echo bench rc=$?; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_s3_vgg.json'))
print(d['value'], d['ms_per_step'], d['gpu_launches'], d['roofline']['achieved'])
for k,v in d['roofline']['families'].items(): print(k, v)
PY
tail -3 gpurun_out/bench_s3_vgg.err
