"""ctypes binding of oracle/librtd_oracle.so (CPU restatement of the reference dose path, oracle/rtd_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
legs as the checker.  Nothing under raytracedicom_b200/ imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librtd_oracle.so")
SRC_PATH = os.path.join(_HERE, "rtd_oracle.c")
BUILD_CMD = ["gcc", "-O2", "-std=c99", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared", SRC_PATH, "-o", LIB_PATH, "-lm"]

_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int)


class OrcLuts(C.Structure):
    _fields_ = [("nEnergySamples", C.c_int), ("nEnergies", C.c_int), ("energiesPerU", _fp), ("peakDepths", _fp),
                ("scaleFacts", _fp), ("ciddMatrix", _fp),
                ("nDensitySamples", C.c_int), ("densityScaleFact", C.c_float), ("densityVector", _fp),
                ("nSpSamples", C.c_int), ("spScaleFact", C.c_float), ("spVector", _fp),
                ("nRRlSamples", C.c_int), ("rRlScaleFact", C.c_float), ("rRlVector", _fp),
                ("nucWeightMatrix", _fp), ("nucSqSigmaMatrix", _fp)]


class OrcBeam(C.Structure):
    _fields_ = [("weights", _fp), ("sx", C.c_int), ("sy", C.c_int), ("nLayers", C.c_int), ("energies", _fp), ("sigmas", _fp),
                ("raySpacing", C.c_float * 2), ("steps", C.c_uint), ("sourceDist", C.c_float * 2),
                ("spotDelta", C.c_float * 3), ("spotOffset", C.c_float * 3),
                ("imM", C.c_float * 9), ("imV", C.c_float * 3), ("doseM", C.c_float * 9), ("doseV", C.c_float * 3),
                ("layerBegin", C.c_int), ("layerEnd", C.c_int)]


class OrcOpts(C.Structure):
    _fields_ = [("bpDepthCutoff", C.c_float), ("convSigmaCutoff", C.c_float), ("ksSigmaCutoff", C.c_float),
                ("rayWeightCutoff", C.c_float), ("minTilesInBatch", C.c_int), ("texFracMode", C.c_int),
                ("noNozzle", C.c_int), ("doseToMedium", C.c_int), ("nuclearCorr", C.c_int)]


class OrcResult(C.Structure):
    _fields_ = [("lSteps", C.c_int), ("bSteps", C.c_int), ("rSteps", C.c_int), ("tSteps", C.c_int),
                ("Rx", C.c_int), ("Ry", C.c_int), ("S", C.c_int), ("L", C.c_int),
                ("bfi", C.c_int), ("bfo", C.c_int), ("Zg", C.c_int), ("fcp", C.c_int),
                ("bboxMin", C.c_int * 3), ("bboxMax", C.c_int * 3), ("entryZ", C.c_float),
                ("updates", C.c_longlong), ("ksTaps", C.c_longlong),
                ("density", _fp), ("wepl", _fp), ("firstInside", _ip), ("firstOutside", _ip), ("weplMin", _fp),
                ("rayWeights", _fp), ("idd", _fp), ("rsig", _fp), ("firstPassive", _ip), ("layerF", _fp), ("layerI", _ip),
                ("hist", _ip), ("batchRad", _ip), ("bev", _fp), ("keepLayer", C.c_int), ("radiusError", C.c_int)]


_lib = None


def build(force: bool = False) -> str:
    """Compile the C restatement (gcc only; no reference sources involved)."""
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(SRC_PATH):
        subprocess.run(BUILD_CMD, check=True)
    return LIB_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        l = C.CDLL(LIB_PATH)
        l.orc_default_opts.argtypes = [C.POINTER(OrcOpts)]
        l.orc_free_result.argtypes = [C.POINTER(OrcResult)]
        l.orc_beam_dose.argtypes = [C.POINTER(OrcLuts), _fp, C.c_int, C.c_int, C.c_int, C.POINTER(OrcBeam), C.c_int, C.c_int, C.c_int,
                                    _fp, C.POINTER(OrcOpts), C.c_int, C.c_int, C.POINTER(OrcResult)]
        l.orc_beam_dose.restype = C.c_int
        l.orc_plan_dose.argtypes = [C.POINTER(OrcLuts), _fp, C.c_int, C.c_int, C.c_int, C.POINTER(OrcBeam), C.c_int, C.c_int, C.c_int, C.c_int,
                                    _fp, C.POINTER(OrcOpts), C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]
        l.orc_plan_dose.restype = C.c_int
        l.orc_find_decimal_ordered.argtypes = [_fp, C.c_int, C.c_float]; l.orc_find_decimal_ordered.restype = C.c_float
        l.orc_vector_interpolate.argtypes = [_fp, C.c_int, C.c_float]; l.orc_vector_interpolate.restype = C.c_float
        _lib = l
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, np.float32)


def _p(a):
    return a.ctypes.data_as(_fp)


def pack_luts(e):
    keep = [_f32(e.energiesPerU), _f32(e.peakDepths), _f32(e.scaleFacts), _f32(e.ciddMatrix), _f32(e.densityVector), _f32(e.spVector), _f32(e.rRlVector)]
    l = OrcLuts(e.nEnergySamples, e.nEnergies, _p(keep[0]), _p(keep[1]), _p(keep[2]), _p(keep[3]),
                e.nDensitySamples, e.densityScaleFact, _p(keep[4]), e.nSpSamples, e.spScaleFact, _p(keep[5]),
                e.nRRlSamples, e.rRlScaleFact, _p(keep[6]), None, None)
    if getattr(e, "nuclearCorr", 0):
        keep += [_f32(e.nucWeightMatrix), _f32(e.nucSqSigmaMatrix)]
        l.nucWeightMatrix, l.nucSqSigmaMatrix = _p(keep[7]), _p(keep[8])
    return l, keep


def pack_beam(b):
    w, e, s = _f32(b.weights), _f32(b.energies), _f32(b.sigmas)
    ob = OrcBeam()
    ob.weights, ob.energies, ob.sigmas = _p(w), _p(e), _p(s)
    ob.nLayers, ob.sy, ob.sx = w.shape
    ob.raySpacing = (C.c_float * 2)(*b.ray_spacing)
    ob.steps = int(b.steps)
    ob.sourceDist = (C.c_float * 2)(*b.source_dist)
    ob.spotDelta = (C.c_float * 3)(*b.spot_delta)
    ob.spotOffset = (C.c_float * 3)(*b.spot_offset)
    ob.imM = (C.c_float * 9)(*_f32(b.gantry_to_im_m).ravel())
    ob.imV = (C.c_float * 3)(*_f32(b.gantry_to_im_v).ravel())
    ob.doseM = (C.c_float * 9)(*_f32(b.gantry_to_dose_m).ravel())
    ob.doseV = (C.c_float * 3)(*_f32(b.gantry_to_dose_v).ravel())
    ob.layerBegin, ob.layerEnd = getattr(b, "layer_begin", 0), getattr(b, "layer_end", 0)
    return ob, [w, e, s]


def default_opts(**kw) -> OrcOpts:
    o = OrcOpts()
    lib().orc_default_opts(C.byref(o))
    for k, v in kw.items():
        setattr(o, k, v)
    return o


def _arr(ptr, shape, dtype):
    if not ptr:
        return None
    n = int(np.prod(shape))
    return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype, copy=True).reshape(shape)


def run_beam(case, beam_no: int = 0, dose: np.ndarray = None, keep_layer: int = -1, stop_after: int = 0, opts: OrcOpts = None) -> dict:
    """Run one beam of ``case`` through the CPU restatement; returns every intermediate as numpy arrays.
    ``dose`` ([nz][ny][nx] float32) is accumulated in place when given."""
    l = lib()
    luts, k1 = pack_luts(case.luts)
    ob, k2 = pack_beam(case.beams[beam_no])
    ct = _f32(case.ct)
    nz, ny, nx = ct.shape
    dnx, dny, dnz = case.dose_dims
    if dose is None:
        dose = np.zeros((dnz, dny, dnx), np.float32)
    assert dose.dtype == np.float32 and dose.flags["C_CONTIGUOUS"] and dose.shape == (dnz, dny, dnx)
    o = opts or default_opts()
    r = OrcResult()
    rc = l.orc_beam_dose(C.byref(luts), _p(ct), nx, ny, nz, C.byref(ob), dnx, dny, dnz, _p(dose), C.byref(o), keep_layer, stop_after, C.byref(r))
    Rx, Ry, S, L = r.Rx, r.Ry, r.S, r.L
    out = dict(rc=rc, lSteps=r.lSteps, bSteps=r.bSteps, rSteps=r.rSteps, tSteps=r.tSteps, Rx=Rx, Ry=Ry, S=S, L=L,
               beamFirstInside=r.bfi, beamFirstOutside=r.bfo, beamFirstGuaranteedPassive=r.Zg, beamFirstCalculatedPassive=r.fcp,
               bboxMin=list(r.bboxMin), bboxMax=list(r.bboxMax), entryZ=r.entryZ, updates=r.updates, ksTaps=r.ksTaps, dose=dose)
    out["density"] = _arr(r.density, (S, Ry, Rx), np.float32)
    out["wepl"] = _arr(r.wepl, (S, Ry, Rx), np.float32)
    out["first_inside"] = _arr(r.firstInside, (Ry, Rx), np.int32)
    out["first_outside"] = _arr(r.firstOutside, (Ry, Rx), np.int32)
    out["wepl_min"] = _arr(r.weplMin, (S,), np.float32)
    out["ray_weights"] = _arr(r.rayWeights, (L, Ry, Rx), np.float32)
    out["idd"] = _arr(r.idd, (S, Ry, Rx), np.float32)
    out["rsig"] = _arr(r.rsig, (S, Ry, Rx), np.float32)
    out["first_passive"] = _arr(r.firstPassive, (L, Ry, Rx), np.int32)
    out["layerF"] = _arr(r.layerF, (L, 5), np.float32)
    out["layerI"] = _arr(r.layerI, (L, 2), np.int32)
    out["hist"] = _arr(r.hist, (L, 34), np.int32)
    out["batchRad"] = _arr(r.batchRad, (L, 34), np.int32)
    out["bev_dose"] = _arr(r.bev, (r.Zg, Ry + 64, Rx + 64), np.float32) if r.Zg > 0 else None
    l.orc_free_result(C.byref(r))
    del k1, k2
    return out


def plan_dose(case, dose: np.ndarray = None, opts: OrcOpts = None, beams=None):
    """All beams of the case (cudaWrapperProtons' beam loop); returns (dose, updates, ksTaps)."""
    l = lib()
    luts, k1 = pack_luts(case.luts)
    bl = case.beams if beams is None else beams
    packed = [pack_beam(b) for b in bl]
    arr = (OrcBeam * len(bl))(*[p[0] for p in packed])
    ct = _f32(case.ct)
    nz, ny, nx = ct.shape
    dnx, dny, dnz = case.dose_dims
    if dose is None:
        dose = np.zeros((dnz, dny, dnx), np.float32)
    o = opts or default_opts()
    u, t = C.c_longlong(0), C.c_longlong(0)
    rc = l.orc_plan_dose(C.byref(luts), _p(ct), nx, ny, nz, arr, len(bl), dnx, dny, dnz, _p(dose), C.byref(o), C.byref(u), C.byref(t))
    if rc:
        raise RuntimeError(f"oracle: status {rc} (3 = kernel superposition radius > 32)")
    del k1, packed
    return dose, u.value, t.value
