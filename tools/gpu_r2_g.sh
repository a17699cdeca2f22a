#!/bin/bash
# tracer variant B: parity tests, then ncu --set full of both tracer kernels on C3 (one plan, tracerBrick = 1 then 0)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "tma_brick" > gpurun_out/pytest_brick.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_brick.log
tail -5 gpurun_out/pytest_brick.log
cat > /tmp/one.py <<'P'
import sys, numpy as np
sys.path.insert(0, ".")
from raytracedicom_b200 import cases, engine
case = cases.make_case("c3"); nx, ny, nz = case.dose_dims
with engine.DoseEngine(0) as eng:
    eng.set_luts(case.luts); eng.set_ct(case.ct); eng.set_dose_grid(nx, ny, nz)
    for brick in (1, 0):
        eng.opts.tracerBrick = brick; eng.opts.noGraph = 1; eng.opts.serialBeams = 1
        eng.dose(case.beams, np.zeros((nz, ny, nx), np.float32))
P
timeout 900 ncu --set full --import-source on --clock-control none -k regex:trace_b -o gpurun_out/r02_trace_ab python /tmp/one.py > gpurun_out/ncu_trace.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu_trace.log
