#!/bin/bash
# Multi-GPU bench line (torchrun, one rank per GPU). usage: gpu_multi.sh N [workload] [reduce: push|nccl] [extra bench args]
N=${1:-2}; W=${2:-c3}; R=${3:-push}; shift 3 2>/dev/null
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps ${STEPS:-10} --warmup 3 --workload $W --reduce $R "$@" 2> gpurun_out/bench_${W}_n${N}_$R.err | grep "^{" > gpurun_out/bench_${W}_n${N}_$R.json
echo "rc=${PIPESTATUS[0]}"
python - <<P
import json
try:
    d=json.loads(open('gpurun_out/bench_${W}_n${N}_$R.json').read().strip().splitlines()[-1])
    print('N=$N $W $R: plan %.3f ms, e2e %s ms, err_vs_1gpu %s, per-rank stage totals %s' % (d['ms_per_step'], d.get('e2e',{}).get('ms_per_plan'), d.get('dose_max_rel_err_vs_1gpu'), d['stage_total_ms_per_rank']))
    print('  steps', d['ms_each_step_rank0'][:10], 'launches/step', d['gpu_launches_per_step'])
    if 'c4' in d: print('  c4: plan %.3f ms e2e %s err %s ranks %s' % (d['c4']['ms_per_step'], d['c4'].get('e2e',{}).get('ms_per_plan'), d['c4'].get('dose_max_rel_err_vs_1gpu'), d['c4']['stage_total_ms_per_rank']))
except Exception as e: print('parse failed', e)
P
grep -v "OMP_NUM_THREADS\|^\*\*\*\*" gpurun_out/bench_${W}_n${N}_$R.err | tail -6
