#!/bin/bash
# Round-end style measurement: reference arm first, then ours (N=1).
mkdir -p gpurun_out
timeout 900 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cut -c1-1500 gpurun_out/bench_ref.json
timeout 600 python bench.py > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "ours rc=$?"
cut -c1-600 gpurun_out/bench_ours.json
for w in c4 c5 c1a c2; do timeout 600 python bench.py --workload $w --steps 5 --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "$w rc=$?"; done
