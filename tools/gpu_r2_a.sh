#!/bin/bash
# Round 2, first GPU visit: parity suite, both bench arms, ncu launch list + full capture of the superposition kernels.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench rc=$?"
cut -c1-1500 gpurun_out/bench_c3.json; tail -5 gpurun_out/bench_c3.err
timeout 900 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cut -c1-600 gpurun_out/bench_ref.json
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-c4 --no-cold"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:ks2_kernel -s 12 -c 8 -o gpurun_out/prof_ks2 -f $CMD > gpurun_out/ncu_full.log 2>&1; echo "full rc=$?"
ls -la gpurun_out/*.ncu-rep
