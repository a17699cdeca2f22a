#!/bin/bash
# A/B of the depth scan's block size (RTD_IDD_BLOCK 0 = four-slot pipeline, 4, 8 = in-tree build, 16): C3 and C4 stage times + parity of the default
mkdir -p gpurun_out
for nb in 0 4 8 16; do
  lib=build_ab/librtdose_idd$nb.so; [ $nb = 8 ] && lib=raytracedicom_b200/librtdose.so
  timeout 600 python bench.py --steps 20 --no-cpu-baseline --no-cold --lib $lib > gpurun_out/bench_idd$nb.json 2> gpurun_out/bench_idd$nb.err; echo "nb=$nb rc=$?"
  python - $nb <<'P'
import json,sys
d=json.loads(open(f'gpurun_out/bench_idd{sys.argv[1]}.json').read().strip().splitlines()[-1])
print('  c3 ms', round(d['ms_per_step'],4), 'idd', round(d['stage_ms_rank0']['msIddSigma'],4), 'checksum', d.get('dose_checksum'), '| c4 ms', round(d['c4']['ms_per_step'],3), 'idd', round(d['c4']['stage_ms_rank0']['msIddSigma'],3), d['c4'].get('dose_checksum'))
P
done
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_all.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_all.log
