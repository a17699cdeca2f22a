#!/bin/bash
# A/B of two builds of librtdose.so on C3 / C4: build_ab/librtdose_<name>.so against the in-tree one. usage: gpu_ab.sh <name> [rounds]
mkdir -p gpurun_out
for i in $(seq ${2:-2}); do for v in $1 intree; do
  lib=build_ab/librtdose_$v.so; [ $v = intree ] && lib=raytracedicom_b200/librtdose.so
  timeout 600 python bench.py --steps 20 --no-cpu-baseline --no-cold --no-e2e --lib $lib > gpurun_out/bench_$v.json 2> gpurun_out/bench_$v.err; echo "v=$v rc=$?"
  python - $v <<'P'
import json,sys
d=json.loads(open(f'gpurun_out/bench_{sys.argv[1]}.json').read().strip().splitlines()[-1])
st=d['stage_ms_rank0']; c=d['c4']['stage_ms_rank0']
print('  c3 ms', round(d['ms_per_step'],4), 'idd', round(st['msIddSigma'],4), 'superp', round(st['msSuperp'],4), 'checksum', d.get('dose_checksum'), '| c4 ms', round(d['c4']['ms_per_step'],3), 'idd', round(c['msIddSigma'],3), 'superp', round(c['msSuperp'],3), d['c4'].get('dose_checksum'))
P
done; done
