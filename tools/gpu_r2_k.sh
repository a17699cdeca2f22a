#!/bin/bash
mkdir -p gpurun_out
for v in 4m16 3m1 2m1 5m1; do
  timeout 600 python bench.py --steps 20 --no-cpu-baseline --no-cold --lib build_ab/librtdose_idd$v.so > gpurun_out/bench_idd$v.json 2> gpurun_out/bench_idd$v.err; echo "v=$v rc=$?"
  python - $v <<'P'
import json,sys
d=json.loads(open(f'gpurun_out/bench_idd{sys.argv[1]}.json').read().strip().splitlines()[-1])
print('  c3 ms', round(d['ms_per_step'],4), 'idd', round(d['stage_ms_rank0']['msIddSigma'],4), 'checksum', d.get('dose_checksum'), '| c4 ms', round(d['c4']['ms_per_step'],3), 'idd', round(d['c4']['stage_ms_rank0']['msIddSigma'],3), d['c4'].get('dose_checksum'))
P
done
