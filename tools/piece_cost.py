"""Stage times of ONE beam of C3 restricted to layer ranges of different length (what a rank of a multi-GPU plan runs): where does the
per-piece floor come from? Prints the stage times for shards [0, n) and the sum over a cover of the beam by shards of n layers."""
import json, sys
import numpy as np
sys.path.insert(0, ".")
from raytracedicom_b200 import cases, engine

case = cases.make_case(sys.argv[1] if len(sys.argv) > 1 else "c3")
beam = case.beams[0]
L = beam.n_layers
nx, ny, nz = case.dose_dims
KEYS = ("msTrace", "msSetup", "msConv", "msIddSigma", "msRadius", "msSuperp", "msTransfer", "msTotal")
with engine.DoseEngine(0) as eng:
    eng.set_luts(case.luts); eng.set_ct(case.ct); eng.set_dose_grid(nx, ny, nz)
    eng.opts.noGraph = 1; eng.opts.serialBeams = 1
    for n in (L, (L + 1) // 2, (L + 3) // 4, 2, 1):
        tot = {k: 0.0 for k in KEYS}; pieces = 0
        for lo in range(0, L, n):
            sh = beam.shard(lo, min(lo + n, L))
            rows = []
            for _ in range(6):
                rows.append(eng.dose([sh], np.zeros((nz, ny, nx), np.float32)).as_dict())
            for k in KEYS: tot[k] += float(np.median([r[k] for r in rows[2:]]))
            pieces += 1
        print(json.dumps({"layers_per_piece": n, "pieces": pieces, "sum_ms": {k: round(v, 4) for k, v in tot.items()}}), flush=True)
