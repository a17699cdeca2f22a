"""Test / baseline infrastructure only (never on the product path).

"Baseline B" of BASELINE.md 2B / SURVEY.md 8d: the reference's OWN CPU stage code, /root/reference/src/cpu_convolution_1d.cpp
(xConvCpuSparse :91, yConvCpuSparse :173), compiled unmodified into oracle/_ref/libref_cpuconv.so (oracle/ref_harness/Makefile,
C wrapper oracle/ref_harness/cpuconv_wrap.cpp) and run on stage K3a of a workload: spot weights -> ray weights
(gpu_convolution_2d.cu:16-71 on the GPU side). The library is built in the development container and travels to the GPU box
with the snapshot; nothing here reads /root/reference at run time.

Problem set-up: per layer sigma_eff = entrySigma / (raySpacing * (1 - entryZ / sourceDist)),
rSigmaEff = 1 / (sqrt(2) sigma_eff), rad = ceil(3 sigma_eff + 1/2) - 1 (the taps |i| <= rad are exactly those with
|distance| < 3 sigma_eff + 1/2 that the GPU kernel visits); spot s sits on ray s * (spotDelta / raySpacing) + (spotOffset /
raySpacing - lSteps). Ray grid, entry plane and entry sigmas come from the CPU oracle stopped after the pre-convolution.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import time

import numpy as np

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libref_cpuconv.so")
_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int)
_lib = None


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        l = C.CDLL(LIB_PATH)
        l.ref_cpuconv_layers.argtypes = [_fp, C.c_int, C.c_int, C.c_int, _fp, _fp, _ip, _ip, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _fp, C.c_int]
        l.ref_cpuconv_layers.restype = C.c_int
        l.ref_cpuconv_max_threads.restype = C.c_int
        _lib = l
    return _lib


def max_threads() -> int:
    return int(lib().ref_cpuconv_max_threads())


def _int_ratio(a: float, b: float, what: str) -> int:
    q = a / b
    if abs(q - round(q)) > 1e-4:
        raise ValueError(f"{what}: {a} / {b} is not an integer (the reference's sparse CPU convolution needs integer pitch / offset)")
    return int(round(q))


def k3a_problem(case, beam_no: int, cutoff: float = 3.0) -> dict:
    """Stage K3a of one beam as the arguments of xConvCpuSparse / yConvCpuSparse, plus the oracle's result for it."""
    from . import pyoracle
    b = case.beams[beam_no]
    o = pyoracle.run_beam(case, beam_no, stop_after=2)
    L, Sy, Sx = b.weights.shape
    rsx, rsy = float(b.ray_spacing[0]), float(b.ray_spacing[1])
    px = rsx * (1.0 - o["entryZ"] / float(b.source_dist[0]))
    py = rsy * (1.0 - o["entryZ"] / float(b.source_dist[1]))
    sig = o["layerF"][:, 3:5].astype(np.float64)
    sex, sey = sig[:, 0] / px, sig[:, 1] / py
    prob = dict(
        weights=np.ascontiguousarray(b.weights, np.float32), L=L, Sx=Sx, Sy=Sy, Rx=o["Rx"], Ry=o["Ry"],
        rSigX=(1.0 / (math.sqrt(2.0) * sex)).astype(np.float32), rSigY=(1.0 / (math.sqrt(2.0) * sey)).astype(np.float32),
        radX=(np.ceil(cutoff * sex + 0.5) - 1).astype(np.int32), radY=(np.ceil(cutoff * sey + 0.5) - 1).astype(np.int32),
        dX=_int_ratio(float(b.spot_delta[0]), rsx, "spot pitch x"), dY=_int_ratio(float(b.spot_delta[1]), rsy, "spot pitch y"),
        offX=_int_ratio(float(b.spot_offset[0]), rsx, "spot offset x") - o["lSteps"],
        offY=_int_ratio(float(b.spot_offset[1]), rsy, "spot offset y") - o["bSteps"],
        oracle_ray_weights=o["ray_weights"])
    return prob


def run(prob: dict, threads: int = 1, repeats: int = 1):
    """-> (ray weights [L][Ry][Rx] from the reference's CPU code, best wall seconds of ``repeats`` runs)."""
    out = np.empty((prob["L"], prob["Ry"], prob["Rx"]), np.float32)
    p = lambda a, t: a.ctypes.data_as(t)
    best = float("inf")
    for _ in range(max(1, repeats)):
        t0 = time.perf_counter()
        rc = lib().ref_cpuconv_layers(p(prob["weights"], _fp), prob["L"], prob["Sx"], prob["Sy"], p(prob["rSigX"], _fp), p(prob["rSigY"], _fp),
                                      p(prob["radX"], _ip), p(prob["radY"], _ip), prob["Rx"], prob["Ry"], prob["offX"], prob["offY"],
                                      prob["dX"], prob["dY"], p(out, _fp), int(threads))
        best = min(best, time.perf_counter() - t0)
        if rc != 0:
            raise RuntimeError(f"ref_cpuconv_layers: taps of layer {rc - 1} leave the ray grid (rad > offset); the reference code would exit(1)")
    return out, best
