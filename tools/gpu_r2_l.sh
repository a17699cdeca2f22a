#!/bin/bash
mkdir -p gpurun_out
for v in r0off intree r0off intree; do
  lib=build_ab/librtdose_$v.so; [ $v = intree ] && lib=raytracedicom_b200/librtdose.so
  timeout 600 python bench.py --steps 20 --no-cpu-baseline --no-cold --lib $lib > gpurun_out/bench_$v.json 2> gpurun_out/bench_$v.err; echo "v=$v rc=$?"
  python - $v <<'P'
import json,sys
d=json.loads(open(f'gpurun_out/bench_{sys.argv[1]}.json').read().strip().splitlines()[-1])
print('  c3 ms', round(d['ms_per_step'],4), 'superp', round(d['stage_ms_rank0']['msSuperp'],4), 'idd', round(d['stage_ms_rank0']['msIddSigma'],4), 'checksum', d.get('dose_checksum'), '| c4 ms', round(d['c4']['ms_per_step'],3), 'superp', round(d['c4']['stage_ms_rank0']['msSuperp'],3), d['c4'].get('dose_checksum'))
P
done
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_all.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_all.log
