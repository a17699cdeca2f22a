#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "tma_brick" > gpurun_out/pytest_brick.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_brick.log
tail -25 gpurun_out/pytest_brick.log
timeout 600 python tools/trace_ab.py c3 c1a c5 > gpurun_out/trace_ab.jsonl 2> gpurun_out/trace_ab.err; echo "ab rc=$?"
cat gpurun_out/trace_ab.jsonl | cut -c1-400; tail -5 gpurun_out/trace_ab.err
