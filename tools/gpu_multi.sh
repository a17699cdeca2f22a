#!/bin/bash
# Multi-GPU bench line (torchrun, one rank per GPU, NCCL reduce of the partial dose grids). usage: gpu_multi.sh N [workload]
N=${1:-2}; W=${2:-c3}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --workload $W > gpurun_out/bench_${W}_n$N.json 2> gpurun_out/bench_${W}_n$N.err
echo "rc=$?"; cut -c1-2500 gpurun_out/bench_${W}_n$N.json; tail -5 gpurun_out/bench_${W}_n$N.err
