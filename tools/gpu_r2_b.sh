#!/bin/bash
# Round 2 GPU visit: parity suite (all failures shown), bench, ncu launch list + full captures.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench rc=$?"
python - <<'P'
import json
try:
    d=json.loads(open('gpurun_out/bench_c3.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('ms_per_step','value','stage_ms_rank0','gpu_launches_per_step','host_syncs_per_plan')})
    print('e2e', d.get('e2e',{}).get('ms_per_plan'), 'cold', d.get('e2e_cold'), 'roofline', {k:d['roofline'][k] for k in ('achieved','frac','avg_launch_ms','launches_per_plan_rank0')}, d['roofline']['fp32'])
    print('c4', d.get('c4'))
except Exception as e: print('bench parse failed', e)
P
tail -5 gpurun_out/bench_c3.err
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-c4 --no-cold"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'ks2_kernel|idd_sigma_kernel|trace_bev_kernel|bev_to_grid_kernel' -s 14 -c 14 -o gpurun_out/prof_all -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "full rc=$?"
ls -la gpurun_out/*.ncu-rep
