"""Runs the mini case (all kernels of the hot path, both tracer variants, the no-alloc tail) once, for compute-sanitizer:
compute-sanitizer --tool memcheck python tools/sanitize_mini.py"""
import sys
import numpy as np
sys.path.insert(0, ".")
from raytracedicom_b200 import cases, engine

case = cases.make_case("mini")
case.ct = np.rint(case.ct).astype(np.float32)
nx, ny, nz = case.dose_dims
with engine.DoseEngine(0) as eng:
    eng.set_luts(case.luts); eng.set_ct(case.ct); eng.set_dose_grid(nx, ny, nz)
    for brick in (0, 1):
        eng.opts.tracerBrick = brick
        dose = np.zeros((nz, ny, nx), np.float32)
        t = eng.dose(case.beams, dose)
        print("brick", brick, "updates", t.updates, "dose max", float(dose.max()), "brick beams", t.tracerBrickBeams)
    dose2 = np.zeros((nz, ny, nx), np.float32)
    eng.opts.tracerBrick = -1
    t = eng.plan_dose(case.ct, case.beams, dose2)
    print("one-call", float(dose2.max()), float(np.abs(dose2 - dose).max()))
print("done")
