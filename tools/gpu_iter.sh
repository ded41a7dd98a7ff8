# iteration call: parity tests + 1-GPU bench (+ optional 2-GPU bench when 2 GPUs are visible)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -5 gpurun_out/pytest_gpu.log
timeout 400 python bench.py --no-cpu-baseline > gpurun_out/bench_vgg.json 2> gpurun_out/bench_vgg.err; echo bench vgg rc=$?; tail -3 gpurun_out/bench_vgg.err
python - <<'PY'
import json
d = json.loads([l for l in open("gpurun_out/bench_vgg.json") if l.startswith("{")][-1])
print(d["value"], d["ms_per_step"], d.get("gpu_launches"), d["e2e"]["value"], d["roofline"]["achieved"], d["clocks"])
for k, v in d["roofline"]["families"].items(): print("   ", k, v)
PY
if [ "$(nvidia-smi -L | wc -l)" -ge 2 ]; then
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29702 bench.py --gpus 2 --warmup 5 > gpurun_out/bench_vgg_2gpu.json 2> gpurun_out/bench_vgg_2gpu.err; echo bench n=2 rc=$?
python -c "
import json
d=json.loads([l for l in open('gpurun_out/bench_vgg_2gpu.json') if l.startswith('{')][-1]); print(2, d['value'], d['ms_per_step'], d['e2e']['value'])"
fi
