"""RT Ion Plan + CT geometry -> ``BeamSettings``: the step the reference's DICOM path never performs.
``main()`` reads the plan through ``rti`` and prints its spots, but constructs the BeamSettings from all-zero weights,
energies and sigmas (/root/reference/src/main.cu:171-197) and rejects more than one beam (main.cu:117-120).
This module completes that mapping with the reference's own conventions wherever it has one:

* spot-to-layer assignment: ``rti::beam_module_ion`` (rti_beam_module_ion.hpp:86-104) - control points in pairs, the
  odd-indexed ones skipped, spots of a kept control point in file order -> done by ``dicom_lite.read_ion_plan``;
* spot sigma = FWHM / 2.355 (rbe_1p1.hpp:90);
* weights as one regular grid ``[layer][sy][sx]`` (beam_settings.h:21): spots are rasterised onto the lattice spanned by
  the distinct x / y positions of the whole beam; positions off the lattice are an error, not an approximation;
* ray spacing (1, 1) mm and step length 1 mm as in main.cu:192, 57-64; the number of tracer steps and the start plane
  come from the CT's extent along the beam axis instead of the hard-coded 512 / 128;
* coordinate chain = rti::coordinate_transform (rti_coordinate_transform.hpp:48-58: iec2dicom * patient_support * gantry
  [* collimator]) with IEC 61217 angles in radians; ScanSpotPositionMap is defined in the IEC GANTRY system, so the
  beam-limiting-device angle does not rotate the spots. The engine's gantry frame is IEC gantry: x, y lateral at the
  isocentre plane, z towards the source; the beam travels along -z (fill_idd_and_sigma_params.cu:22).
The C++ host driver implements the same mapping (raytracedicom_b200/host/plan_reader.cpp); tests compare the two.
"""
from __future__ import annotations

from typing import List, Sequence

import numpy as np

from . import cases

FWHM_TO_SIGMA = 2.355     # rbe_1p1.hpp:90
LATTICE_TOL = 1.0e-3      # mm
MAX_POSITION = 1.0e4      # mm: anything farther from the axis than 10 m is a corrupt file
MAX_LATTICE_NODES = 4096


def gantry_to_patient(gantry_deg: float, couch_deg: float) -> np.ndarray:
    """3x3 rotation taking IEC gantry coordinates to DICOM patient coordinates (HFS): p = T Rz(-couch) Ry(gantry) g,
    T = IEC fixed -> DICOM patient (x, y, z)_d = (x, -z, y)_f."""
    g, s = np.deg2rad(gantry_deg), np.deg2rad(-couch_deg)
    ry = np.array([[np.cos(g), 0.0, np.sin(g)], [0.0, 1.0, 0.0], [-np.sin(g), 0.0, np.cos(g)]])
    rz = np.array([[np.cos(s), -np.sin(s), 0.0], [np.sin(s), np.cos(s), 0.0], [0.0, 0.0, 1.0]])
    t = np.array([[1.0, 0.0, 0.0], [0.0, 0.0, -1.0], [0.0, 1.0, 0.0]])
    return t @ rz @ ry


def _lattice(values: np.ndarray, default_pitch: float, min_count: int):
    """Distinct positions -> (origin, pitch, count) of the regular lattice they lie on; padded with zero-weight nodes to
    ``min_count`` (a one-spot plan still gets a ray grid of at least 32 x 32 rays)."""
    values = np.asarray(values, np.float64)
    if values.size == 0:
        raise ValueError("beam has no scan spots")
    if not np.all(np.isfinite(values)) or np.max(np.abs(values)) > MAX_POSITION:
        raise ValueError("scan spot position is not a finite distance in mm")
    u = np.unique(np.round(values / LATTICE_TOL) * LATTICE_TOL)
    if u.size == 1:
        pitch = float(default_pitch)
    else:
        pitch = float(np.min(np.diff(u)))
        k = (u - u[0]) / pitch
        if np.max(np.abs(k - np.round(k))) * pitch > 10 * LATTICE_TOL:
            raise ValueError("scan spot positions do not lie on a regular lattice")
    nodes = round((u[-1] - u[0]) / pitch) + 1
    if not nodes <= MAX_LATTICE_NODES:
        raise ValueError("scan spot lattice has too many nodes")
    n = int(nodes)
    x0 = float(u[0])
    side = 0
    while n < min_count:          # alternately one node before, one after: a single spot ends up in the centre
        if side == 0:
            x0 -= pitch
        n += 1; side ^= 1
    return x0, pitch, n


def beam_settings_from_plan(beam: dict, ct_shape: Sequence[int], im_idx_to_world: np.ndarray, origin: np.ndarray,
                            ray_spacing=(1.0, 1.0), step_mm: float = 1.0, min_grid: int = 3, default_pitch: float = 5.0) -> cases.Beam:
    """One beam dict of ``dicom_lite.read_ion_plan`` + CT geometry of ``dicom_lite.read_ct_series`` -> cases.Beam
    (field-for-field BeamSettings). ct_shape = (nz, ny, nx)."""
    layers = beam["layers"]
    if not layers:
        raise ValueError("beam has no energy layers")
    allx = np.concatenate([l["x"] for l in layers]); ally = np.concatenate([l["y"] for l in layers])
    x0, dx, sx = _lattice(allx, default_pitch, min_grid)
    y0, dy, sy = _lattice(ally, default_pitch, min_grid)
    L = len(layers)
    w = np.zeros((L, sy, sx), np.float32)
    for li, l in enumerate(layers):
        ix = np.rint((l["x"].astype(np.float64) - x0) / dx).astype(int)
        iy = np.rint((l["y"].astype(np.float64) - y0) / dy).astype(int)
        if ix.size and (ix.min() < 0 or ix.max() >= sx or iy.min() < 0 or iy.max() >= sy or not np.all(np.isfinite(l["w"]))):
            raise ValueError("scan spot outside its lattice or non-finite weight in beam " + str(beam.get("name", "")))
        np.add.at(w[li], (iy, ix), l["w"].astype(np.float32))      # repaintings of one position add up
    energies = np.array([l["energy"] for l in layers], np.float32)
    sigmas = np.array([[l["fwhm"][0] / FWHM_TO_SIGMA, l["fwhm"][1] / FWHM_TO_SIGMA] for l in layers], np.float32)
    rot = gantry_to_patient(beam["gantry_angle"], beam.get("couch_angle", 0.0))
    a_inv = np.linalg.inv(np.asarray(im_idx_to_world, np.float64))
    m = (a_inv @ rot).astype(np.float32)
    v = (a_inv @ (np.asarray(beam["isocenter"], np.float64) - np.asarray(origin, np.float64))).astype(np.float32)
    nz, ny, nx = ct_shape
    zstart, steps = cases.steps_and_start(m, v, (nx, ny, nz), dz=step_mm)
    return cases.Beam(w, energies, sigmas, tuple(ray_spacing), steps, (float(beam["sad"][0]), float(beam["sad"][1])),
                      (dx, dy, -step_mm), (x0, y0, zstart), m, v)


def case_from_dicom(ct_dir: str, rtplan: str, beam_names: Sequence[str] = (), luts=None, name: str = "dicom") -> cases.Case:
    """The reference's command line (--ct_dir, --rtplan, --beams; config.cpp:13-51) -> a runnable case, any number of beams
    (the reference stops at one, main.cu:117-120, although cudaWrapperProtons loops over beams, kernel_wrapper.cu:601)."""
    from . import dicom_lite
    from .energy_reader import load_packaged_luts
    ct, m, o = dicom_lite.read_ct_series(ct_dir)
    plan = dicom_lite.read_ion_plan(rtplan)
    by_name = {b["name"]: b for b in plan}
    chosen: List[dict] = []
    for n in (beam_names or [b["name"] for b in plan]):
        if n not in by_name:
            raise ValueError(f"beam {n} not found in {rtplan} (has: {sorted(by_name)})")
        chosen.append(by_name[n])
    beams = [beam_settings_from_plan(b, ct.shape, m, o) for b in chosen]
    sp = (float(np.linalg.norm(m[:, 0])), float(np.linalg.norm(m[:, 1])), float(np.linalg.norm(m[:, 2])))
    return cases.Case(name, ct, beams, luts or load_packaged_luts(), sp)
