"""ctypes binding of librtdose.so (include/rtdose.h) and the Python mirror of the reference's host interface.

``cuda_wrapper_protons(im_vol, dose_vol, beams, idd_data)`` has the argument meaning of the reference's
``cudaWrapperProtons`` (/root/reference/src/kernel_wrapper.cuh:161): CT volume, dose volume accumulated in
place, list of BeamSettings-shaped beams, EnergyStruct. There is NO CPU fallback: importing works without a
GPU (so the symbol table can be checked), creating a context without one raises ``RtdError``.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence

import numpy as np

from .cases import Beam
from .energy_reader import EnergyStruct

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librtdose.so")

HIST_BINS = 34
STATUS = {0: "RTD_OK", 1: "RTD_ERR_INVALID", 2: "RTD_ERR_CUDA", 3: "RTD_ERR_RADIUS", 4: "RTD_ERR_STATE", 5: "RTD_ERR_NO_DEVICE"}
FETCH = dict(density=0, wepl=1, first_inside=2, first_outside=3, wepl_min=4, ray_weights=5, idd=6, rsigma=7,
             first_passive=8, bev_dose=9, tile_radius=10)

_fp = C.POINTER(C.c_float)


class RtdLuts(C.Structure):
    _fields_ = [("nEnergySamples", C.c_int), ("nEnergies", C.c_int), ("energiesPerU", _fp), ("peakDepths", _fp),
                ("scaleFacts", _fp), ("ciddMatrix", _fp),
                ("nDensitySamples", C.c_int), ("densityScaleFact", C.c_float), ("densityVector", _fp),
                ("nSpSamples", C.c_int), ("spScaleFact", C.c_float), ("spVector", _fp),
                ("nRRlSamples", C.c_int), ("rRlScaleFact", C.c_float), ("rRlVector", _fp),
                ("nucWeightMatrix", _fp), ("nucSqSigmaMatrix", _fp)]


class RtdBeam(C.Structure):
    _fields_ = [("weights", _fp), ("sx", C.c_int), ("sy", C.c_int), ("nLayers", C.c_int), ("energies", _fp), ("sigmas", _fp),
                ("raySpacing", C.c_float * 2), ("steps", C.c_uint), ("sourceDist", C.c_float * 2),
                ("spotDelta", C.c_float * 3), ("spotOffset", C.c_float * 3),
                ("gantryToImIdxM", C.c_float * 9), ("gantryToImIdxV", C.c_float * 3),
                ("gantryToDoseIdxM", C.c_float * 9), ("gantryToDoseIdxV", C.c_float * 3),
                ("layerBegin", C.c_int), ("layerEnd", C.c_int)]


class RtdOpts(C.Structure):
    _fields_ = [("bpDepthCutoff", C.c_float), ("convSigmaCutoff", C.c_float), ("ksSigmaCutoff", C.c_float),
                ("rayWeightCutoff", C.c_float), ("minTilesInBatch", C.c_int), ("ks2BodyStep", C.c_int),
                ("serialBeams", C.c_int), ("noZeroCopyDose", C.c_int), ("noNozzle", C.c_int), ("doseToMedium", C.c_int),
                ("atomicDose", C.c_int), ("ctPull", C.c_int), ("nuclearCorr", C.c_int), ("noGraph", C.c_int), ("fullIntermediates", C.c_int), ("tracerBrick", C.c_int)]


class RtdTiming(C.Structure):
    _fields_ = [("msTotal", C.c_float), ("msUpload", C.c_float), ("msTrace", C.c_float), ("msSetup", C.c_float),
                ("msConv", C.c_float), ("msIddSigma", C.c_float), ("msRadius", C.c_float), ("msSuperp", C.c_float),
                ("msTransfer", C.c_float), ("msDownload", C.c_float), ("updates", C.c_int64), ("ksTaps", C.c_int64),
                ("kernelLaunches", C.c_int), ("hostSyncs", C.c_int), ("h2dBytes", C.c_int64), ("d2hBytes", C.c_int64),
                ("doseVoxelsModified", C.c_int64), ("ksLaunchesWithWork", C.c_int), ("tracerBrickBeams", C.c_int),
                ("tracerOutsideSamples", C.c_int), ("reserved0", C.c_int)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class RtdBeamReport(C.Structure):
    _fields_ = [("lSteps", C.c_int), ("bSteps", C.c_int), ("rSteps", C.c_int), ("tSteps", C.c_int),
                ("rayDimX", C.c_int), ("rayDimY", C.c_int), ("steps", C.c_int), ("nLayers", C.c_int),
                ("beamFirstInside", C.c_int), ("beamFirstOutside", C.c_int), ("beamFirstGuaranteedPassive", C.c_int),
                ("beamFirstCalculatedPassive", C.c_int), ("bboxMin", C.c_int * 3), ("bboxMax", C.c_int * 3),
                ("entryZ", C.c_float), ("updates", C.c_int64)]

    def as_dict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_}
        d["bboxMin"] = list(self.bboxMin); d["bboxMax"] = list(self.bboxMax)
        return d


class RtdLayerReport(C.Structure):
    _fields_ = [("energyIdx", C.c_float), ("energyScaleFact", C.c_float), ("peakDepth", C.c_float),
                ("entrySigmaX", C.c_float), ("entrySigmaY", C.c_float), ("afterLastStep", C.c_int),
                ("layerFirstPassive", C.c_int), ("radiusHist", C.c_int * HIST_BINS), ("batchRadius", C.c_int * HIST_BINS)]

    def as_dict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_}
        d["radiusHist"] = list(self.radiusHist); d["batchRadius"] = list(self.batchRadius)
        return d


class RtdWorkItem(C.Structure):
    _fields_ = [("rank", C.c_int), ("beam", C.c_int), ("layerBegin", C.c_int), ("layerEnd", C.c_int)]


class RtdError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"{STATUS.get(status, status)}: {msg}")
        self.status = status


EXPORTS = ["rtd_default_opts", "rtd_create", "rtd_create_on_stream", "rtd_destroy", "rtd_last_error", "rtd_version", "rtd_set_luts", "rtd_set_ct",
           "rtd_set_ct_device", "rtd_set_dose_grid", "rtd_dose", "rtd_dose_device", "rtd_plan_dose", "rtd_plan_dose_device", "rtd_add_device_to_host", "rtd_add_device_to_host_noalloc", "rtd_partition", "rtd_dose_slab", "rtd_group_create", "rtd_group_destroy", "rtd_group_ct_create", "rtd_group_ct_open", "rtd_group_dose", "rtd_group_barrier", "rtd_plan_dose_group", "rtd_shared_dose_create", "rtd_shared_dose_open", "rtd_shared_dose_close", "rtd_push_dose_to_peer",
           "rtd_get_beam_report",
           "rtd_get_layer_report", "rtd_debug_fetch", "rtd_debug_ct_boxes", "rtd_stream"]

_lib = None


def load_library() -> C.CDLL:
    """Load librtdose.so; fails loudly if the CUDA extension has not been built (no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          f"or `make -C raytracedicom_b200/csrc`. The dose path has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    lib.rtd_default_opts.argtypes = [C.POINTER(RtdOpts)]; lib.rtd_default_opts.restype = None
    lib.rtd_create.argtypes = [C.POINTER(vp), C.c_int]; lib.rtd_create.restype = C.c_int
    lib.rtd_create_on_stream.argtypes = [C.POINTER(vp), C.c_int, vp]; lib.rtd_create_on_stream.restype = C.c_int
    lib.rtd_destroy.argtypes = [vp]; lib.rtd_destroy.restype = None
    lib.rtd_last_error.argtypes = [vp]; lib.rtd_last_error.restype = C.c_char_p
    lib.rtd_version.argtypes = []; lib.rtd_version.restype = C.c_char_p
    lib.rtd_set_luts.argtypes = [vp, C.POINTER(RtdLuts)]; lib.rtd_set_luts.restype = C.c_int
    lib.rtd_set_ct.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int]; lib.rtd_set_ct.restype = C.c_int
    lib.rtd_set_ct_device.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int]; lib.rtd_set_ct_device.restype = C.c_int
    lib.rtd_set_dose_grid.argtypes = [vp, C.c_int, C.c_int, C.c_int]; lib.rtd_set_dose_grid.restype = C.c_int
    lib.rtd_dose.argtypes = [vp, C.POINTER(RtdBeam), C.c_int, vp, C.POINTER(RtdOpts), C.POINTER(RtdTiming)]; lib.rtd_dose.restype = C.c_int
    lib.rtd_dose_device.argtypes = lib.rtd_dose.argtypes; lib.rtd_dose_device.restype = C.c_int
    lib.rtd_plan_dose.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.POINTER(RtdBeam), C.c_int, vp, C.POINTER(RtdOpts), C.POINTER(RtdTiming)]
    lib.rtd_plan_dose.restype = C.c_int
    lib.rtd_plan_dose_device.argtypes = lib.rtd_plan_dose.argtypes; lib.rtd_plan_dose_device.restype = C.c_int
    lib.rtd_add_device_to_host.argtypes = [vp, vp, vp, C.c_size_t, C.POINTER(C.c_int64)]; lib.rtd_add_device_to_host.restype = C.c_int
    lib.rtd_group_ct_create.argtypes = [vp, C.c_size_t, C.c_char_p]; lib.rtd_group_ct_create.restype = C.c_int
    lib.rtd_group_ct_open.argtypes = [vp, C.c_char_p]; lib.rtd_group_ct_open.restype = C.c_int
    lib.rtd_add_device_to_host_noalloc.argtypes = [vp, vp, vp, C.c_size_t, C.POINTER(C.c_int64)]; lib.rtd_add_device_to_host_noalloc.restype = C.c_int
    lib.rtd_shared_dose_create.argtypes = [vp, C.c_size_t, C.POINTER(vp), C.c_char_p]; lib.rtd_shared_dose_create.restype = C.c_int
    lib.rtd_shared_dose_open.argtypes = [vp, C.c_char_p, C.POINTER(vp)]; lib.rtd_shared_dose_open.restype = C.c_int
    lib.rtd_shared_dose_close.argtypes = [vp, vp, C.c_int]; lib.rtd_shared_dose_close.restype = C.c_int
    lib.rtd_push_dose_to_peer.argtypes = [vp, vp, vp, C.c_size_t]; lib.rtd_push_dose_to_peer.restype = C.c_int
    lib.rtd_partition.argtypes = [C.POINTER(RtdBeam), C.c_int, C.POINTER(RtdLuts), C.c_int, C.c_int, C.POINTER(RtdWorkItem), C.c_int]; lib.rtd_partition.restype = C.c_int
    lib.rtd_dose_slab.argtypes = [C.POINTER(RtdBeam), C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.POINTER(C.c_int), C.POINTER(C.c_int)]; lib.rtd_dose_slab.restype = C.c_int
    lib.rtd_group_create.argtypes = [vp, C.c_int, C.c_int, C.c_size_t, C.c_char_p]; lib.rtd_group_create.restype = C.c_int
    lib.rtd_group_destroy.argtypes = [vp]; lib.rtd_group_destroy.restype = C.c_int
    lib.rtd_group_dose.argtypes = [vp]; lib.rtd_group_dose.restype = vp
    lib.rtd_group_barrier.argtypes = [vp]; lib.rtd_group_barrier.restype = C.c_int
    lib.rtd_plan_dose_group.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.POINTER(RtdBeam), C.c_int, vp, C.POINTER(RtdOpts), C.POINTER(RtdTiming)]
    lib.rtd_plan_dose_group.restype = C.c_int
    lib.rtd_get_beam_report.argtypes = [vp, C.c_int, C.POINTER(RtdBeamReport)]; lib.rtd_get_beam_report.restype = C.c_int
    lib.rtd_get_layer_report.argtypes = [vp, C.c_int, C.c_int, C.POINTER(RtdLayerReport)]; lib.rtd_get_layer_report.restype = C.c_int
    lib.rtd_debug_fetch.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, C.c_size_t]; lib.rtd_debug_fetch.restype = C.c_int
    lib.rtd_debug_ct_boxes.argtypes = [C.POINTER(RtdBeam), C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int]; lib.rtd_debug_ct_boxes.restype = C.c_int
    lib.rtd_stream.argtypes = [vp]; lib.rtd_stream.restype = vp
    _lib = lib
    return lib


def _f32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(_fp)


def pack_beams(beams: Sequence[Beam]):
    """Beam list -> (ctypes array of rtd_beam, keep-alive list)."""
    arr = (RtdBeam * max(len(beams), 1))()
    keep = []
    for i, b in enumerate(beams):
        w, e, s = _f32(b.weights), _f32(b.energies), _f32(b.sigmas)
        keep += [w, e, s]
        rb = arr[i]
        rb.weights, rb.energies, rb.sigmas = _ptr(w), _ptr(e), _ptr(s)
        rb.nLayers, rb.sy, rb.sx = w.shape
        rb.raySpacing = (C.c_float * 2)(*b.ray_spacing)
        rb.steps = int(b.steps)
        rb.sourceDist = (C.c_float * 2)(*b.source_dist)
        rb.spotDelta = (C.c_float * 3)(*b.spot_delta)
        rb.spotOffset = (C.c_float * 3)(*b.spot_offset)
        rb.gantryToImIdxM = (C.c_float * 9)(*_f32(b.gantry_to_im_m).ravel())
        rb.gantryToImIdxV = (C.c_float * 3)(*_f32(b.gantry_to_im_v).ravel())
        rb.gantryToDoseIdxM = (C.c_float * 9)(*_f32(b.gantry_to_dose_m).ravel())
        rb.gantryToDoseIdxV = (C.c_float * 3)(*_f32(b.gantry_to_dose_v).ravel())
        rb.layerBegin, rb.layerEnd = getattr(b, "layer_begin", 0), getattr(b, "layer_end", 0)
    return arr, keep


def pack_luts(e: EnergyStruct):
    """EnergyStruct -> (rtd_luts, keep-alive list)."""
    arrs = [_f32(e.energiesPerU), _f32(e.peakDepths), _f32(e.scaleFacts), _f32(e.ciddMatrix), _f32(e.densityVector),
            _f32(e.spVector), _f32(e.rRlVector)]
    l = RtdLuts(e.nEnergySamples, e.nEnergies, _ptr(arrs[0]), _ptr(arrs[1]), _ptr(arrs[2]), _ptr(arrs[3]),
                e.nDensitySamples, e.densityScaleFact, _ptr(arrs[4]), e.nSpSamples, e.spScaleFact, _ptr(arrs[5]),
                e.nRRlSamples, e.rRlScaleFact, _ptr(arrs[6]), None, None)
    if getattr(e, "nuclearCorr", 0):
        arrs += [_f32(e.nucWeightMatrix), _f32(e.nucSqSigmaMatrix)]
        l.nucWeightMatrix, l.nucSqSigmaMatrix = _ptr(arrs[7]), _ptr(arrs[8])
    return l, arrs


def partition(beams: Sequence[Beam], luts: EnergyStruct, world_size: int, chunks_per_beam: int = 0):
    """(rank, beam, first layer, one-past-last layer) work items of the plan on ``world_size`` ranks (rtd_partition: the
    partitioner rtd_plan_dose_group itself uses; host only, no device needed)."""
    lib = load_library()
    arr, keep = pack_beams(beams)
    l, keep2 = pack_luts(luts)
    n = lib.rtd_partition(arr, len(beams), C.byref(l), int(world_size), int(chunks_per_beam), None, 0)
    if n < 0:
        raise ValueError("rtd_partition: bad plan / LUTs / world size")
    items = (RtdWorkItem * max(n, 1))()
    lib.rtd_partition(arr, len(beams), C.byref(l), int(world_size), int(chunks_per_beam), items, n)
    del keep, keep2
    return [(it.rank, it.beam, it.layerBegin, it.layerEnd) for it in items[:n]]


def dose_slab(beams: Sequence[Beam], dose_dims_xyz, sigma_cutoff: float = 3.0):
    """Inclusive z-slab (z0, z1) of the dose grid the plan can reach, (0, -1) if none (rtd_dose_slab; host only)."""
    lib = load_library()
    arr, keep = pack_beams(beams)
    z0, z1 = C.c_int(0), C.c_int(-1)
    nx, ny, nz = (int(v) for v in dose_dims_xyz)
    r = lib.rtd_dose_slab(arr, len(beams), nx, ny, nz, float(sigma_cutoff), C.byref(z0), C.byref(z1))
    if r < 0:
        raise ValueError("rtd_dose_slab: bad plan / dims")
    del keep
    return (int(z0.value), int(z1.value)) if r == 1 else (0, -1)


def ct_boxes(beam: Beam, ct_dims_xyz) -> np.ndarray:
    """The CT sub-volumes rtd_plan_dose uploads for ``beam`` into a CT of (nx, ny, nz) voxels: int array [n][6] of inclusive
    boxes (x0, y0, z0, x1, y1, z1). Host geometry only - works without a GPU (rtd_debug_ct_boxes)."""
    lib = load_library()
    arr, keep = pack_beams([beam])
    nx, ny, nz = (int(v) for v in ct_dims_xyz)
    n = lib.rtd_debug_ct_boxes(arr, nx, ny, nz, None, 0)
    if n < 0:
        raise ValueError("rtd_debug_ct_boxes: bad beam or CT dims")
    out = np.zeros((max(n, 1), 6), np.int32)
    lib.rtd_debug_ct_boxes(arr, nx, ny, nz, out.ctypes.data_as(C.POINTER(C.c_int)), n)
    del keep
    return out[:n]


class DoseEngine:
    """One context per GPU. Mirrors the life cycle inside cudaWrapperProtons: LUT upload, CT upload, per-plan dose."""

    def __init__(self, device: int = 0, stream: Optional[int] = None):
        """``stream``: a cudaStream_t (integer handle) the caller owns, e.g. ``torch.cuda.Stream().cuda_stream``; the context then
        issues its work on it (rtd_create_on_stream) and never destroys it."""
        self.lib = load_library()
        self._h = C.c_void_p()
        if stream is None:
            st = self.lib.rtd_create(C.byref(self._h), int(device))
        else:
            st = self.lib.rtd_create_on_stream(C.byref(self._h), int(device), C.c_void_p(int(stream)))
        if st != 0:
            self._h = C.c_void_p()
            raise RtdError(st, "rtd_create failed (a CUDA device is required; there is no CPU fallback)")
        self.device = device
        self.opts = RtdOpts()
        self.lib.rtd_default_opts(C.byref(self.opts))
        self.timing = RtdTiming()
        self._keep = []
        self.dose_dims = None

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self.lib.rtd_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, st: int):
        if st != 0:
            raise RtdError(st, self.lib.rtd_last_error(self._h).decode())

    def set_luts(self, e: EnergyStruct):
        e.validate()
        l, arrs = pack_luts(e)
        self._check(self.lib.rtd_set_luts(self._h, C.byref(l)))

    def set_ct(self, ct: np.ndarray):
        """ct: float32 [nz][ny][nx] HU+1000 (host)."""
        ct = _f32(ct)
        nz, ny, nx = ct.shape
        self._check(self.lib.rtd_set_ct(self._h, ct.ctypes.data, nx, ny, nz))

    def set_ct_ptr(self, ptr: int, nx: int, ny: int, nz: int, device: bool):
        fn = self.lib.rtd_set_ct_device if device else self.lib.rtd_set_ct
        self._check(fn(self._h, C.c_void_p(ptr), nx, ny, nz))

    def set_dose_grid(self, nx: int, ny: int, nz: int):
        self.dose_dims = (nx, ny, nz)
        self._check(self.lib.rtd_set_dose_grid(self._h, nx, ny, nz))

    def dose(self, beams: Sequence[Beam], dose: np.ndarray) -> RtdTiming:
        """Accumulate the plan's dose onto the HOST array ``dose`` ([nz][ny][nx] float32, modified in place)."""
        if dose.dtype != np.float32 or not dose.flags["C_CONTIGUOUS"]:
            raise ValueError("dose must be a C-contiguous float32 array")
        arr, keep = pack_beams(beams)
        self._check(self.lib.rtd_dose(self._h, arr, len(beams), dose.ctypes.data, C.byref(self.opts), C.byref(self.timing)))
        return self.timing

    def dose_ptr(self, packed, n_beams: int, ptr: int, device: bool) -> RtdTiming:
        fn = self.lib.rtd_dose_device if device else self.lib.rtd_dose
        self._check(fn(self._h, packed, n_beams, C.c_void_p(ptr), C.byref(self.opts), C.byref(self.timing)))
        return self.timing

    def plan_dose(self, ct: np.ndarray, beams: Sequence[Beam], dose: np.ndarray) -> RtdTiming:
        """One-call equivalent of cudaWrapperProtons: HOST CT ([nz][ny][nx] float32 HU+1000) and HOST dose (accumulated
        in place); only the sub-volumes the plan touches are copied (rtd_plan_dose). Needs set_luts + set_dose_grid."""
        if dose.dtype != np.float32 or not dose.flags["C_CONTIGUOUS"]:
            raise ValueError("dose must be a C-contiguous float32 array")
        ct = _f32(ct)
        nz, ny, nx = ct.shape
        arr, keep = pack_beams(beams)
        self._check(self.lib.rtd_plan_dose(self._h, ct.ctypes.data, nx, ny, nz, arr, len(beams), dose.ctypes.data,
                                           C.byref(self.opts), C.byref(self.timing)))
        return self.timing

    def plan_dose_ptr(self, ct_ptr: int, nx: int, ny: int, nz: int, packed, n_beams: int, dose_ptr: int, device: bool = False) -> RtdTiming:
        fn = self.lib.rtd_plan_dose_device if device else self.lib.rtd_plan_dose
        self._check(fn(self._h, C.c_void_p(ct_ptr), nx, ny, nz, packed, n_beams, C.c_void_p(dose_ptr),
                                           C.byref(self.opts), C.byref(self.timing)))
        return self.timing

    def add_device_to_host(self, d_src_ptr: int, host_dst_ptr: int, n: int) -> int:
        """host[i] += device[i] where device[i] != 0 (multi-GPU tail, rtd_add_device_to_host); returns the voxels updated."""
        m = C.c_int64(0)
        self._check(self.lib.rtd_add_device_to_host(self._h, C.c_void_p(d_src_ptr), C.c_void_p(host_dst_ptr), n, C.byref(m)))
        return int(m.value)

    def add_device_to_host_noalloc(self, d_src_ptr: int, dst_ptr: int, n: int) -> int:
        """rtd_add_device_to_host_noalloc: the same tail without any allocation (fixed scratch inside the library, n <= 2^28, dst device or
        page-locked host memory); returns the voxels updated."""
        m = C.c_int64(0)
        self._check(self.lib.rtd_add_device_to_host_noalloc(self._h, C.c_void_p(d_src_ptr), C.c_void_p(dst_ptr), n, C.byref(m)))
        return int(m.value)

    # ---- one plan on the GPUs of a node (rtd_group_*) ----
    def group_create(self, rank: int, world: int, n_dose_floats: int, handle: Optional[bytes] = None) -> bytes:
        """Rank 0 creates the plan's dose grid and returns its 64-byte IPC handle; the other ranks pass that handle in."""
        h = C.create_string_buffer(handle if handle is not None else b"\0" * 64, 64)
        self._check(self.lib.rtd_group_create(self._h, int(rank), int(world), int(n_dose_floats), h))
        return h.raw

    def group_ct_create(self, n_bytes: int) -> bytes:
        """CT relay of the group: this rank's staging buffer (same size on every rank); returns its 64-byte IPC handle."""
        buf = C.create_string_buffer(64)
        self._check(self.lib.rtd_group_ct_create(self._h, int(n_bytes), buf))
        return buf.raw

    def group_ct_open(self, handles: Sequence[bytes]):
        """Maps the staging buffers of all ranks (handles in rank order): rtd_plan_dose_group then uploads every CT box once per node."""
        blob = b"".join(handles)
        self._check(self.lib.rtd_group_ct_open(self._h, blob))

    def group_destroy(self):
        self._check(self.lib.rtd_group_destroy(self._h))

    def group_dose_ptr(self) -> int:
        return int(self.lib.rtd_group_dose(self._h) or 0)

    def group_barrier(self):
        self._check(self.lib.rtd_group_barrier(self._h))

    def plan_dose_group(self, ct_ptr: int, nx: int, ny: int, nz: int, packed, n_beams: int, host_dose_ptr: int = 0) -> RtdTiming:
        """This rank's share of the WHOLE plan ``packed`` (rtd_plan_dose_group). ct_ptr = 0: the CT resident from set_ct."""
        self._check(self.lib.rtd_plan_dose_group(self._h, C.c_void_p(ct_ptr) if ct_ptr else None, nx, ny, nz, packed, n_beams,
                                                 C.c_void_p(host_dose_ptr) if host_dose_ptr else None, C.byref(self.opts), C.byref(self.timing)))
        return self.timing

    def shared_dose_create(self, n_floats: int):
        """(device pointer, 64-byte IPC handle) of a zeroed dose grid other ranks of the node can map (rtd_shared_dose_create)."""
        p = C.c_void_p(); h = C.create_string_buffer(64)
        self._check(self.lib.rtd_shared_dose_create(self._h, n_floats, C.byref(p), h))
        return int(p.value), h.raw

    def shared_dose_open(self, handle: bytes) -> int:
        p = C.c_void_p()
        self._check(self.lib.rtd_shared_dose_open(self._h, handle, C.byref(p)))
        return int(p.value)

    def shared_dose_close(self, ptr: int, owner: bool):
        self._check(self.lib.rtd_shared_dose_close(self._h, C.c_void_p(ptr), 1 if owner else 0))

    def push_dose_to_peer(self, d_src_ptr: int, d_peer_dst_ptr: int, n: int):
        """peer[i] += device[i] for the float4 groups holding dose (rtd_push_dose_to_peer); asynchronous on the engine stream."""
        self._check(self.lib.rtd_push_dose_to_peer(self._h, C.c_void_p(d_src_ptr), C.c_void_p(d_peer_dst_ptr), n))

    def beam_report(self, beam_no: int) -> dict:
        r = RtdBeamReport()
        self._check(self.lib.rtd_get_beam_report(self._h, beam_no, C.byref(r)))
        return r.as_dict()

    def layer_report(self, beam_no: int, layer_no: int) -> dict:
        r = RtdLayerReport()
        self._check(self.lib.rtd_get_layer_report(self._h, beam_no, layer_no, C.byref(r)))
        return r.as_dict()

    def fetch(self, beam_no: int, what: str, layer: int = 0) -> np.ndarray:
        rep = self.beam_report(beam_no)
        rx, ry, s, L = rep["rayDimX"], rep["rayDimY"], rep["steps"], rep["nLayers"]
        nz = rep["beamFirstGuaranteedPassive"] - rep["beamFirstInside"]
        shapes = dict(density=((s, ry, rx), np.float32), wepl=((s, ry, rx), np.float32), first_inside=((ry, rx), np.int32),
                      first_outside=((ry, rx), np.int32), wepl_min=((s,), np.float32), ray_weights=((L, ry, rx), np.float32),
                      idd=((nz, ry, rx), np.float32), rsigma=((nz, ry, rx), np.float32), first_passive=((ry, rx), np.int32),
                      bev_dose=((rep["beamFirstGuaranteedPassive"], ry + 64, rx + 64), np.float32),
                      tile_radius=((nz, ry // 8, rx // 32), np.uint8))
        shape, dt = shapes[what]
        out = np.empty(shape, dt)
        self._check(self.lib.rtd_debug_fetch(self._h, beam_no, FETCH[what], layer, out.ctypes.data, out.nbytes))
        return out

    @property
    def stream(self) -> int:
        return int(self.lib.rtd_stream(self._h) or 0)


def cuda_wrapper_protons(im_vol: np.ndarray, dose_vol: np.ndarray, beams: List[Beam], idd_data: EnergyStruct,
                         out_stream=None, device: int = 0, engine: Optional[DoseEngine] = None) -> RtdTiming:
    """Python mirror of ``cudaWrapperProtons(imVol, doseVol, beams, iddData, outStream)`` (kernel_wrapper.cuh:161).

    ``dose_vol`` is accumulated in place. Unlike the reference nothing is deleted and the device is not reset."""
    own = engine is None
    eng = engine or DoseEngine(device)
    try:
        eng.set_luts(idd_data)
        nz, ny, nx = dose_vol.shape
        eng.set_dose_grid(nx, ny, nz)
        t = eng.plan_dose(im_vol, beams, dose_vol)
        if out_stream is not None:
            out_stream.write(f"    Total global execution time (excluding GPU initialisation): {t.msTotal} ms.\n\n")
        return t
    finally:
        if own:
            eng.close()
