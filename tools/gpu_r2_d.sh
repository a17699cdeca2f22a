#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --no-cpu-baseline --no-cold > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench rc=$?"
python - <<'P'
import json
try:
    d=json.loads(open('gpurun_out/bench_c3.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('ms_per_step','stage_ms_rank0','gpu_launches_per_step','host_syncs_per_plan')})
    print('e2e', d.get('e2e',{}).get('ms_per_plan'), 'c4', d['c4']['ms_per_step'], d['c4']['stage_ms_rank0'])
except Exception as e: print('bench parse failed', e)
P
tail -3 gpurun_out/bench_c3.err
