"""A/B of the two tracer variants on one B200 (tools/gpu_r2_f.sh): stage times of a plan with rtd_opts.tracerBrick = 0 (tex3D) and
1 (TMA brick), kernel by kernel (noGraph, serialBeams), median of N calls. Prints one JSON line per (workload, variant)."""
import json
import sys

import numpy as np

sys.path.insert(0, ".")
from raytracedicom_b200 import cases, engine

N = 15
for name in sys.argv[1:] or ["c3"]:
    case = cases.make_case(name)
    nx, ny, nz = case.dose_dims
    with engine.DoseEngine(0) as eng:
        eng.set_luts(case.luts); eng.set_ct(case.ct); eng.set_dose_grid(nx, ny, nz)
        for brick in (0, 1, 0, 1):
            eng.opts.tracerBrick = brick; eng.opts.noGraph = 1; eng.opts.serialBeams = 1
            rows = []
            for i in range(N):
                dose = np.zeros((nz, ny, nx), np.float32)
                rows.append(eng.dose(case.beams, dose).as_dict())
            med = {k: float(np.median([r[k] for r in rows[3:]])) for k in ("msTrace", "msSetup", "msIddSigma", "msSuperp")}
            print(json.dumps({"workload": name, "tracerBrick": brick, "median_ms": med, "tracerBrickBeams": rows[-1]["tracerBrickBeams"],
                              "tracerOutsideSamples": rows[-1]["tracerOutsideSamples"], "checksum": float(dose.sum(dtype=np.float64))}), flush=True)
