"""LUT loading for the dose engine.

Mirrors the reference's ``energyReader`` (/root/reference/src/energy_reader.cpp:11-163) and its
``EnergyStruct`` (/root/reference/src/energy_struct.h:13-38): whitespace separated text files

* ``proton_cumul_ddd_data.txt``   "nSamples nEnergies", energies[nE], peakDepths[nE], scaleFacts[nE], cidd[nE*nS]
* ``density_Schneider2000_adj.txt`` / ``HU_to_SP_H&N_adj.txt`` / ``radiation_length.txt``  "n scale", values[n]

The physics tables themselves are beam-model *data*; the repo ships them converted to a binary
``data/luts.npz`` (made by ``tools/make_lut_fixture.py`` from the reference's LUTs directory) so the GPU
box, which has no /root/reference, can run the same physics.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field

import numpy as np

_DATA_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")


@dataclass
class EnergyStruct:
    """Same fields, same meaning as the reference struct (energy_struct.h:14-31)."""

    nEnergySamples: int = 0
    nEnergies: int = 0
    energiesPerU: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float32))
    peakDepths: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float32))
    scaleFacts: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float32))
    ciddMatrix: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float32))  # [nEnergies*nEnergySamples]
    nDensitySamples: int = 0
    densityScaleFact: float = 0.0
    densityVector: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float32))
    nSpSamples: int = 0
    spScaleFact: float = 0.0
    spVector: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float32))
    nRRlSamples: int = 0
    rRlScaleFact: float = 0.0
    rRlVector: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float32))
    # NUCLEAR_CORR builds only (energy_struct.h:33-36): nuclear-halo weight and variance, same shape as ciddMatrix
    nucWeightMatrix: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float32))
    nucSqSigmaMatrix: np.ndarray = field(default_factory=lambda: np.zeros(0, np.float32))
    nuclearCorr: int = 0          # 0 = none, else the model the two matrices belong to (NUCLEAR_MODELS)

    def validate(self) -> None:
        if self.nEnergies < 2 or self.nEnergySamples < 2:
            raise ValueError("EnergyStruct: need at least 2 energies and 2 depth samples")
        if self.ciddMatrix.size != self.nEnergies * self.nEnergySamples:
            raise ValueError("EnergyStruct: ciddMatrix size mismatch")
        for name, n in (("energiesPerU", self.nEnergies), ("peakDepths", self.nEnergies),
                        ("scaleFacts", self.nEnergies), ("densityVector", self.nDensitySamples),
                        ("spVector", self.nSpSamples), ("rRlVector", self.nRRlSamples)):
            if getattr(self, name).size != n:
                raise ValueError(f"EnergyStruct: {name} has {getattr(self, name).size} entries, expected {n}")
        if self.nuclearCorr and (self.nucWeightMatrix.size != self.ciddMatrix.size or self.nucSqSigmaMatrix.size != self.ciddMatrix.size):
            raise ValueError("EnergyStruct: nuclear matrices must have the shape of ciddMatrix")
        if np.any(np.diff(self.energiesPerU) <= 0):
            raise ValueError("EnergyStruct: energiesPerU must be strictly increasing (vector_find.h:130)")


# rtd_opts.nuclearCorr / orc_opts.nuclearCorr: 0 = off (the reference's default build), 1.. = the reference's
# NUCLEAR_CORR models SOUKUP / FLUKA / GAUSS_FIT (CMakeLists.txt:57-62; the reference numbers them 0 / 1 / 2)
NUCLEAR_MODELS = {"soukup": 1, "fluka": 2, "gauss_fit": 3}
NUCLEAR_FILES = {1: "nuclear_weights_and_sigmas_Soukup.txt", 2: "nuclear_weights_and_sigmas_Fluka.txt", 3: "nuclear_weights_and_sigmas_fit.txt"}


def _tokens(path: str) -> np.ndarray:
    try:
        with open(path, "r") as f:
            txt = f.read()
    except OSError as e:  # energy_reader.cpp:19-25: "Failed to open <path>"
        raise RuntimeError(f"Failed to open {path}") from e
    return np.array(txt.split(), dtype=np.float64)


def _read_1d(path: str):
    t = _tokens(path)
    n = int(t[0])
    scale = float(np.float32(t[1]))
    if t.size < 2 + n:
        raise RuntimeError(f"{path}: expected {n} samples, found {t.size - 2}")
    return n, scale, t[2:2 + n].astype(np.float32)


def energy_reader(data_path: str, water_cube_test: bool = False, nuclear_corr: int = 0) -> EnergyStruct:
    """Text LUT loader; ``data_path`` must end with a separator exactly like the reference's
    PHYS_DATA_DIRECTORY (energy_reader.cpp:19 concatenates ``dataPath + file``).

    ``water_cube_test`` selects ``radiation_length_inc_water.txt`` as the WATER_CUBE_TEST build does
    (energy_reader.cpp:77-84)."""
    e = EnergyStruct()
    t = _tokens(data_path + "proton_cumul_ddd_data.txt")
    e.nEnergySamples, e.nEnergies = int(t[0]), int(t[1])
    nE, nS = e.nEnergies, e.nEnergySamples
    need = 2 + 3 * nE + nE * nS
    if t.size < need:
        raise RuntimeError("proton_cumul_ddd_data.txt: truncated")
    p = 2
    e.energiesPerU = t[p:p + nE].astype(np.float32); p += nE
    e.peakDepths = t[p:p + nE].astype(np.float32); p += nE
    e.scaleFacts = t[p:p + nE].astype(np.float32); p += nE
    e.ciddMatrix = t[p:p + nE * nS].astype(np.float32)
    e.nDensitySamples, e.densityScaleFact, e.densityVector = _read_1d(data_path + "density_Schneider2000_adj.txt")
    e.nSpSamples, e.spScaleFact, e.spVector = _read_1d(data_path + "HU_to_SP_H&N_adj.txt")
    rl = "radiation_length_inc_water.txt" if water_cube_test else "radiation_length.txt"
    e.nRRlSamples, e.rRlScaleFact, e.rRlVector = _read_1d(data_path + rl)
    if nuclear_corr:
        # energy_reader.cpp:103-162: same header as the IDD file (checked to 0.01), then the two matrices
        name = NUCLEAR_FILES[nuclear_corr]
        t = _tokens(data_path + name)
        if int(t[0]) != nS or int(t[1]) != nE:
            raise RuntimeError(f"Number of samples or energies in {name} different from proton_cumul_ddd_data.txt")
        p = 2
        for what, ref in (("Energies", e.energiesPerU), ("Peak depths", e.peakDepths), ("Scale facts", e.scaleFacts)):
            if np.any(np.abs(t[p:p + nE].astype(np.float32) - ref) > 0.01):
                raise RuntimeError(f"{what} in {name} different from proton_cumul_ddd_data.txt")
            p += nE
        if t.size < p + 2 * nE * nS:
            raise RuntimeError(f"{name}: truncated")
        e.nucWeightMatrix = t[p:p + nE * nS].astype(np.float32); p += nE * nS
        e.nucSqSigmaMatrix = t[p:p + nE * nS].astype(np.float32)
        e.nuclearCorr = nuclear_corr
    e.validate()
    return e


def load_packaged_luts(water_cube_test: bool = False, nuclear_corr: int = 0) -> EnergyStruct:
    """Load ``data/luts.npz`` (binary conversion of the reference's LUT text files); ``nuclear_corr`` (NUCLEAR_MODELS) adds the two
    matrices of that model from ``data/luts_nuclear.npz`` (nuclear_weights_and_sigmas_*.txt, energy_reader.cpp:103-162)."""
    path = os.path.join(_DATA_DIR, "luts.npz")
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path} missing: run tools/make_lut_fixture.py where /root/reference exists")
    z = np.load(path)
    e = EnergyStruct()
    e.nEnergySamples, e.nEnergies = int(z["nEnergySamples"]), int(z["nEnergies"])
    e.energiesPerU = z["energiesPerU"].astype(np.float32)
    e.peakDepths = z["peakDepths"].astype(np.float32)
    e.scaleFacts = z["scaleFacts"].astype(np.float32)
    e.ciddMatrix = z["ciddMatrix"].astype(np.float32).ravel()
    e.densityVector = z["densityVector"].astype(np.float32); e.nDensitySamples = e.densityVector.size
    e.densityScaleFact = float(z["densityScaleFact"])
    e.spVector = z["spVector"].astype(np.float32); e.nSpSamples = e.spVector.size
    e.spScaleFact = float(z["spScaleFact"])
    key = "rRlVectorWater" if water_cube_test else "rRlVector"
    e.rRlVector = z[key].astype(np.float32); e.nRRlSamples = e.rRlVector.size
    e.rRlScaleFact = float(z["rRlScaleFact"])
    if nuclear_corr:
        npath = os.path.join(_DATA_DIR, "luts_nuclear.npz")
        if not os.path.exists(npath):
            raise FileNotFoundError(f"{npath} missing: run tools/make_lut_fixture.py where /root/reference exists")
        zn = np.load(npath)
        e.nucWeightMatrix = zn[f"weight_{int(nuclear_corr)}"].astype(np.float32).ravel()
        e.nucSqSigmaMatrix = zn[f"sqsigma_{int(nuclear_corr)}"].astype(np.float32).ravel()
        e.nuclearCorr = int(nuclear_corr)
    e.validate()
    return e


def write_text_luts(data_path: str, e: EnergyStruct = None, water_cube_test: bool = False) -> None:
    """Write an EnergyStruct in the reference's text LUT format (the files ``energy_reader`` / the C++ ``energyReader``
    read; ``data_path`` ends with a separator). float32 values are printed with 9 significant digits, so reading them
    back reproduces every bit. Lets the DICOM input path of the host driver run where /root/reference/LUTs is absent."""
    e = e or load_packaged_luts(water_cube_test)
    os.makedirs(data_path, exist_ok=True)

    def fmt(a):
        return "\n".join("%.9g" % float(v) for v in np.asarray(a, np.float32).ravel())

    with open(data_path + "proton_cumul_ddd_data.txt", "w") as f:
        f.write(f"{e.nEnergySamples} {e.nEnergies}\n")
        for a in (e.energiesPerU, e.peakDepths, e.scaleFacts, e.ciddMatrix):
            f.write(fmt(a) + "\n")
    rl = "radiation_length_inc_water.txt" if water_cube_test else "radiation_length.txt"
    for name, n, s, v in (("density_Schneider2000_adj.txt", e.nDensitySamples, e.densityScaleFact, e.densityVector),
                          ("HU_to_SP_H&N_adj.txt", e.nSpSamples, e.spScaleFact, e.spVector), (rl, e.nRRlSamples, e.rRlScaleFact, e.rRlVector)):
        with open(data_path + name, "w") as f:
            f.write(f"{n} {'%.9g' % float(s)}\n" + fmt(v) + "\n")
