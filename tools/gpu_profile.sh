#!/bin/bash
# ncu evidence for profiles/: launch list of one short bench run + a --set full capture of the superposition kernels.
set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 || { tail -5 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:ks2_kernel -s 9 -c 9 -o gpurun_out/prof_ks2 -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'idd_sigma_kernel|bev_to_grid_kernel|trace_bev_kernel' -s 9 -c 9 -o gpurun_out/prof_rest -f $CMD > gpurun_out/ncu_full2.log 2>&1
echo "full2 rc=$?"
ls -la gpurun_out/*.ncu-rep
# launch list of the large-radius workload (which superposition kernel carries it)
CMD4="python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline --workload c4"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_c4.csv $CMD4 > gpurun_out/ncu_launches_c4.log 2>&1
echo "launch list c4 rc=$?"
