#!/bin/bash
# A/B of Ks2Params::bodyStep (RTD_KS2_BODY_STEP = 1 | 2 | 4) on C4 and C3, then the GPU parity suite under step 2.
mkdir -p gpurun_out
for s in 1 2 4; do RTD_KS2_BODY_STEP=$s timeout 25 python bench.py --workload c4 --steps 5 --no-cpu-baseline --no-e2e > gpurun_out/bs_c4_$s.json 2> gpurun_out/bs_c4_$s.err; done
for s in 1 2; do RTD_KS2_BODY_STEP=$s timeout 25 python bench.py --steps 10 --no-cpu-baseline --no-e2e > gpurun_out/bs_c3_$s.json 2> gpurun_out/bs_c3_$s.err; done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/bs_c*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['ms_per_step'],3), round(d['stage_ms_rank0']['msSuperp'],3), d['dose_checksum'], d['dose_max'])
    except Exception as e: print(f, 'failed', e)
P
RTD_KS2_BODY_STEP=2 timeout 60 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_bs2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_bs2.log; tail -3 gpurun_out/pytest_bs2.log
