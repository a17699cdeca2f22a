#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_all.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_gpu_all.log
timeout 900 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench rc=$?"
python - <<'P'
import json
d=json.loads(open('gpurun_out/bench_c3.json').read().strip().splitlines()[-1])
e=d['e2e']; print('ms', d['ms_per_step'], 'e2e', e['ms_per_plan'], 'h2d', e['h2d_bytes_per_step'], 'cold', d.get('e2e_cold'))
c=d.get('c4'); 
if c: print('c4', c['ms_per_step'], 'e2e', c['e2e']['ms_per_plan'], c['e2e']['h2d_bytes_per_step'])
P
tail -3 gpurun_out/bench_c3.err
