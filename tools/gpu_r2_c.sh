#!/bin/bash
# A/B of the ks2 micro-optimisations (packed erf table / early G loads / work-item prefetch): variants built into gpurun_variants/.
mkdir -p gpurun_out
for f in gpurun_variants/librtdose_*.so; do
  n=$(basename $f .so)
  timeout 120 python bench.py --lib $f --steps 10 --no-e2e --no-cpu-baseline --no-c4 --no-cold > gpurun_out/ab_$n.json 2> gpurun_out/ab_$n.err || tail -3 gpurun_out/ab_$n.err
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/ab_librtdose_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); s=d['stage_ms_rank0']
        print(f.split('librtdose_')[1][:-5], 'plan %.3f ms'%d['ms_per_step'], 'superp %.3f'%s['msSuperp'], 'idd %.3f'%s['msIddSigma'], 'trace %.3f'%s['msTrace'], 'serial total %.3f'%s['msTotal'], 'checksum', d['dose_checksum'])
    except Exception as e: print(f,'failed',e)
P
timeout 600 python -m pytest tests -m gpu -q -k "mini or large_radius or slab or variants" > gpurun_out/pytest_subset.log 2>&1; tail -4 gpurun_out/pytest_subset.log
