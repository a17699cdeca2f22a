"""Where does the cold drop-in call go? Times rtd_create / rtd_set_luts / rtd_set_dose_grid / first rtd_plan_dose / second rtd_plan_dose /
rtd_destroy separately, twice in one process (the second round shows what was one-time module loading)."""
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
torch.cuda.init(); torch.cuda.synchronize()
for rnd in range(2):
    t = [time.perf_counter()]
    eng = engine.DoseEngine(0); t.append(time.perf_counter())
    eng.set_luts(case.luts); t.append(time.perf_counter())
    eng.set_dose_grid(nx, ny, nz); t.append(time.perf_counter())
    eng.plan_dose_ptr(ct.data_ptr(), cnx, cny, cnz, packed, len(case.beams), dose.data_ptr()); t.append(time.perf_counter())
    eng.plan_dose_ptr(ct.data_ptr(), cnx, cny, cnz, packed, len(case.beams), dose.data_ptr()); t.append(time.perf_counter())
    eng.close(); t.append(time.perf_counter())
    names = ["create", "set_luts", "set_dose_grid", "plan #1", "plan #2", "destroy"]
    print("round", rnd, {n: round((b - a) * 1e3, 3) for n, a, b in zip(names, t, t[1:])})
