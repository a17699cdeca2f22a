#!/bin/bash
# Multi-GPU bench line (torchrun, one rank per GPU). usage: gpu_multi.sh N [workload] [reduce: peer|nccl]
N=${1:-2}; W=${2:-c3}; R=${3:-peer}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps ${STEPS:-10} --warmup 3 --workload $W --reduce $R 2> gpurun_out/bench_${W}_n${N}_$R.err | grep "^{" > gpurun_out/bench_${W}_n${N}_$R.json
echo "rc=${PIPESTATUS[0]}"; cut -c1-300 gpurun_out/bench_${W}_n${N}_$R.json; grep -v "OMP_NUM_THREADS\|^\*\*\*\*" gpurun_out/bench_${W}_n${N}_$R.err | tail -8
