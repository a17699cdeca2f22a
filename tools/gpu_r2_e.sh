#!/bin/bash
# Round 2 evidence run: full parity suite, both bench arms, ncu launch list + full captures (C3), C4 launch list.
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
for w in c1a c2 c4 c5; do timeout 600 python bench.py --workload $w --steps 5 --no-cpu-baseline --no-cold > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "$w rc=$?"; done
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-c4 --no-cold"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'ks2_kernel|idd_sigma_kernel|trace_bev_kernel|bev_to_grid_kernel' -s 14 -c 14 -o gpurun_out/prof_all -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "full rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_c4.csv python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-c4 --no-cold --workload c4 > gpurun_out/ncu_launches_c4.log 2>&1; echo "c4 list rc=$?"
python - <<'P'
import json
for n in ('c3','ref','c1a','c2','c4','c5'):
    try:
        d=json.loads(open('gpurun_out/bench_%s.json'%n).read().strip().splitlines()[-1])
        print(n, 'ms', d.get('ms_per_step'), 'e2e', (d.get('e2e') or {}).get('ms_per_plan'), 'cold', (d.get('e2e_cold') or {}).get('ms_per_plan'), 'eager', (d.get('reference_eager_module_loading') or {}).get('ms_per_step'))
    except Exception as e: print(n,'failed',e)
P
