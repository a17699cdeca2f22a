"""Per-beam stage END times of the one-call host path (rtd_plan_dose, pinned CT and dose), from the library's own events
(RTD_TIMELINE=1): where does the end-to-end time of a plan go? usage: RTD_TIMELINE=1 python tools/e2e_timeline.py [workload]"""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from raytracedicom_b200 import cases, engine

case = cases.make_case(sys.argv[1] if len(sys.argv) > 1 else "c3")
nx, ny, nz = case.dose_dims
cnz, cny, cnx = case.ct.shape
ct = torch.from_numpy(case.ct).pin_memory()
dose = torch.zeros((nz, ny, nx), dtype=torch.float32).pin_memory()
packed, keep = engine.pack_beams(case.beams)
with engine.DoseEngine(0) as eng:
    eng.set_luts(case.luts); eng.set_dose_grid(nx, ny, nz)
    for i in range(6):
        if i == 5: print("---- last call", file=sys.stderr, flush=True)
        t0 = time.perf_counter()
        tm = eng.plan_dose_ptr(ct.data_ptr(), cnx, cny, cnz, packed, len(case.beams), dose.data_ptr())
        wall = (time.perf_counter() - t0) * 1e3
    print(f"wall {wall:.3f} ms, msTotal {tm.msTotal:.3f}, msUpload {tm.msUpload:.3f}, h2d {tm.h2dBytes}", file=sys.stderr)
