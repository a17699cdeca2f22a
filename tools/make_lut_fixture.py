#!/usr/bin/env python
"""Convert the reference's physics LUT text files (/root/reference/LUTs/*.txt, read by
energy_reader.cpp:11-101) into raytracedicom_b200/data/luts.npz so that tests and benchmarks can run on the
GPU box, where /root/reference does not exist. Run in the build container:  python tools/make_lut_fixture.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raytracedicom_b200.energy_reader import energy_reader

src = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/LUTs/"
e = energy_reader(src, water_cube_test=False)
w = energy_reader(src, water_cube_test=True)
out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "raytracedicom_b200", "data", "luts.npz")
np.savez_compressed(out, nEnergySamples=e.nEnergySamples, nEnergies=e.nEnergies, energiesPerU=e.energiesPerU,
                    peakDepths=e.peakDepths, scaleFacts=e.scaleFacts,
                    ciddMatrix=e.ciddMatrix.reshape(e.nEnergies, e.nEnergySamples),
                    densityVector=e.densityVector, densityScaleFact=np.float32(e.densityScaleFact),
                    spVector=e.spVector, spScaleFact=np.float32(e.spScaleFact),
                    rRlVector=e.rRlVector, rRlVectorWater=w.rRlVector, rRlScaleFact=np.float32(e.rRlScaleFact))
print("wrote", out, os.path.getsize(out), "bytes; energies", e.nEnergies, "samples", e.nEnergySamples,
      "diff water entries:", int(np.sum(e.rRlVector != w.rRlVector)))

# the LUTs of the NUCLEAR_CORR builds (energy_reader.cpp:103-162): nuclear-halo weight and variance per model
from raytracedicom_b200.energy_reader import NUCLEAR_MODELS
nuc = {}
for name, model in NUCLEAR_MODELS.items():
    n = energy_reader(src, nuclear_corr=model)
    nuc[f"weight_{model}"] = n.nucWeightMatrix.reshape(e.nEnergies, e.nEnergySamples).astype(np.float32)
    nuc[f"sqsigma_{model}"] = n.nucSqSigmaMatrix.reshape(e.nEnergies, e.nEnergySamples).astype(np.float32)
nout = os.path.join(os.path.dirname(out), "luts_nuclear.npz")
np.savez_compressed(nout, **nuc)
print("wrote", nout, os.path.getsize(nout), "bytes")
