mkdir -p gpurun_out
for pdl in 1 513 521 527 1; do
NITI_B200_PDL=$pdl timeout 60 python bench.py --no-cpu-baseline --steps 50 > gpurun_out/bench_pdl$pdl.json 2> gpurun_out/bench_pdl$pdl.err; echo bench pdl=$pdl rc=$?; tail -2 gpurun_out/bench_pdl$pdl.err
python -c "
import json
d=json.loads([l for l in open('gpurun_out/bench_pdl$pdl.json') if l.startswith('{')][-1]); print('pdl', $pdl, d['value'], d['ms_per_step'], d['e2e']['value'])"
nvidia-smi --query-gpu=utilization.gpu --format=csv,noheader
done
