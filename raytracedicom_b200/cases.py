"""Synthetic phantoms + plans for the configurations BASELINE.json names, expressed at
the reference's real boundary: raw CT array + ``BeamSettings``-shaped beams + ``EnergyStruct`` LUTs
(/root/reference/src/main.cu:39-99, 192-206).  The same arrays are fed to the new engine (C ABI), to the
CPU oracle and - through an RTDCASE1 file - to the rebuilt reference (oracle/ref_harness).

Geometry conventions (identical to the reference): CT is float32 HU+1000, x fastest,
array shape [nz][ny][nx]; voxel centres sit at integer image indices; beams travel along -z gantry;
``gantryToImIdx`` maps gantry mm -> image index; dose grid == CT grid.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import List

import numpy as np

from .energy_reader import EnergyStruct, load_packaged_luts

SEED = 20261104


class Lcg:
    """x = x*1664525 + 1013904223 (mod 2^32); u = (x>>8)/2^24  (SURVEY.md App. C)."""

    def __init__(self, seed: int = SEED):
        self.x = seed & 0xFFFFFFFF

    def uniform(self, n: int) -> np.ndarray:
        out = np.empty(n, np.float32)
        x = self.x
        for i in range(n):
            x = (x * 1664525 + 1013904223) & 0xFFFFFFFF
            out[i] = np.float32((x >> 8) / 16777216.0)
        self.x = x
        return out


@dataclass
class Beam:
    """Field-for-field mirror of the reference's BeamSettings (beam_settings.h:101-109)."""

    weights: np.ndarray           # [L][Sy][Sx] float32
    energies: np.ndarray          # [L]
    sigmas: np.ndarray            # [L][2] (x, y) spot sigma at iso in air, mm
    ray_spacing: tuple = (1.0, 1.0)
    steps: int = 512
    source_dist: tuple = (float("inf"), float("inf"))
    spot_delta: tuple = (1.0, 1.0, -1.0)      # spotIdxToGantry delta
    spot_offset: tuple = (0.0, 0.0, 0.0)      # spotIdxToGantry offset
    gantry_to_im_m: np.ndarray = field(default_factory=lambda: np.eye(3, dtype=np.float32))
    gantry_to_im_v: np.ndarray = field(default_factory=lambda: np.zeros(3, np.float32))
    gantry_to_dose_m: np.ndarray = None
    gantry_to_dose_v: np.ndarray = None
    layer_begin: int = 0          # layer shard [layer_begin, layer_end) (0,0 = all); rtd_beam.layerBegin/End
    layer_end: int = 0

    def __post_init__(self):
        self.weights = np.ascontiguousarray(self.weights, np.float32)
        self.energies = np.ascontiguousarray(self.energies, np.float32)
        self.sigmas = np.ascontiguousarray(self.sigmas, np.float32).reshape(-1, 2)
        self.gantry_to_im_m = np.ascontiguousarray(self.gantry_to_im_m, np.float32).reshape(3, 3)
        self.gantry_to_im_v = np.ascontiguousarray(self.gantry_to_im_v, np.float32).reshape(3)
        if self.gantry_to_dose_m is None:
            self.gantry_to_dose_m = self.gantry_to_im_m.copy()
            self.gantry_to_dose_v = self.gantry_to_im_v.copy()
        self.gantry_to_dose_m = np.ascontiguousarray(self.gantry_to_dose_m, np.float32).reshape(3, 3)
        self.gantry_to_dose_v = np.ascontiguousarray(self.gantry_to_dose_v, np.float32).reshape(3)
        L = self.weights.shape[0]
        if self.weights.ndim != 3 or self.energies.shape != (L,) or self.sigmas.shape != (L, 2):
            raise ValueError("Beam: weights must be [L][Sy][Sx], energies [L], sigmas [L][2]")

    @property
    def n_layers(self) -> int:
        return self.weights.shape[0]

    def shard(self, lo: int, hi: int) -> "Beam":
        """Same beam (same ray grid, tracer and cut-offs) restricted to energy layers [lo, hi): the multi-GPU work
        item of SURVEY.md 8e. The doses of disjoint shards add up to the dose of the whole beam."""
        return Beam(self.weights, self.energies, self.sigmas, self.ray_spacing, self.steps, self.source_dist, self.spot_delta,
                    self.spot_offset, self.gantry_to_im_m, self.gantry_to_im_v, self.gantry_to_dose_m, self.gantry_to_dose_v, lo, hi)

    def layer_slice(self, lo: int, hi: int) -> "Beam":
        """Sub-beam holding energy layers [lo, hi) - dose is linear in layers."""
        return Beam(self.weights[lo:hi], self.energies[lo:hi], self.sigmas[lo:hi], self.ray_spacing, self.steps,
                    self.source_dist, self.spot_delta, self.spot_offset, self.gantry_to_im_m, self.gantry_to_im_v,
                    self.gantry_to_dose_m, self.gantry_to_dose_v)


@dataclass
class Case:
    name: str
    ct: np.ndarray                # [nz][ny][nx] float32 HU+1000
    beams: List[Beam]
    luts: EnergyStruct
    spacing: tuple = (1.0, 1.0, 1.0)
    dose_dims: tuple = None       # (nx, ny, nz); defaults to the CT dims

    def __post_init__(self):
        self.ct = np.ascontiguousarray(self.ct, np.float32)
        if self.dose_dims is None:
            self.dose_dims = (self.ct.shape[2], self.ct.shape[1], self.ct.shape[0])

    @property
    def ct_dims(self):
        return (self.ct.shape[2], self.ct.shape[1], self.ct.shape[0])

    def write(self, path: str) -> None:
        """RTDCASE1 file (layout documented in oracle/ref_harness/case_io.h)."""
        l = self.luts
        with open(path, "wb") as f:
            f.write(b"RTDCASE1")
            f.write(struct.pack("<3i", *self.ct_dims)); f.write(self.ct.tobytes())
            f.write(struct.pack("<3i", *self.dose_dims))
            f.write(struct.pack("<2i", l.nEnergySamples, l.nEnergies))
            for a in (l.energiesPerU, l.peakDepths, l.scaleFacts, l.ciddMatrix):
                f.write(np.ascontiguousarray(a, np.float32).tobytes())
            for n, s, v in ((l.nDensitySamples, l.densityScaleFact, l.densityVector),
                            (l.nSpSamples, l.spScaleFact, l.spVector), (l.nRRlSamples, l.rRlScaleFact, l.rRlVector)):
                f.write(struct.pack("<if", n, s)); f.write(np.ascontiguousarray(v, np.float32).tobytes())
            f.write(struct.pack("<i", len(self.beams)))
            for b in self.beams:
                L, Sy, Sx = b.weights.shape
                f.write(struct.pack("<4i", Sx, Sy, L, b.steps))
                f.write(b.weights.tobytes()); f.write(b.energies.tobytes()); f.write(b.sigmas.tobytes())
                f.write(struct.pack("<2f", *b.ray_spacing)); f.write(struct.pack("<2f", *b.source_dist))
                f.write(struct.pack("<3f", *b.spot_delta)); f.write(struct.pack("<3f", *b.spot_offset))
                f.write(b.gantry_to_im_m.tobytes()); f.write(b.gantry_to_im_v.tobytes())
                f.write(b.gantry_to_dose_m.tobytes()); f.write(b.gantry_to_dose_v.tobytes())
            if getattr(l, "nuclearCorr", 0):      # optional trailer: the LUTs of a NUCLEAR_CORR build (energy_struct.h:33-36)
                f.write(b"RTDNUCL1"); f.write(struct.pack("<i", int(l.nuclearCorr)))
                f.write(np.ascontiguousarray(l.nucWeightMatrix, np.float32).tobytes())
                f.write(np.ascontiguousarray(l.nucSqSigmaMatrix, np.float32).tobytes())


# ----------------------------------------------------------------------------------------------------
# helpers
def find_decimal_ordered(ordered: np.ndarray, value: float) -> np.float32:
    """vector_find.h:130-145 in float32."""
    v = np.float32(value)
    if v >= ordered[-1]:
        return np.float32(len(ordered) - 1)
    if v < ordered[0]:
        return np.float32(0.0)
    i = int(np.searchsorted(ordered, v, side="right") - 1)
    return np.float32(i) + (v - ordered[i]) / np.float32(ordered[i + 1] - ordered[i])


def vector_interpolate(lst: np.ndarray, idx) -> np.float32:
    """vector_interpolate.h:19-30 in float32."""
    idx = np.float32(idx)
    if idx <= 0:
        return np.float32(lst[0])
    if idx >= len(lst) - 1:
        return np.float32(lst[-1])
    i = int(np.floor(idx))
    dec = np.float32(idx - np.float32(i))
    return np.float32(lst[i] + np.float32(lst[i + 1] - lst[i]) * dec)


def peak_depth(luts: EnergyStruct, energy: float) -> float:
    return float(vector_interpolate(luts.peakDepths, find_decimal_ordered(luts.energiesPerU, energy)))


def gantry_frame(angle_deg: float):
    """Rows: gantry x, y, z axes in world coordinates. The beam travels along -z gantry; the gantry rotates
    in the axial (world x-y) plane; gantry y is the world z (slice) axis."""
    t = np.deg2rad(angle_deg)
    xg = np.array([np.cos(t), np.sin(t), 0.0])
    yg = np.array([0.0, 0.0, 1.0])
    zg = np.cross(xg, yg)
    return xg, yg, zg


def gantry_to_im_idx(angle_deg: float, iso_world, origin_world, spacing):
    """idx = (iso + G^T g - origin) / spacing  ->  (M, v) float32."""
    xg, yg, zg = gantry_frame(angle_deg)
    g2w = np.stack([xg, yg, zg], axis=1)             # columns = axes
    sp = np.asarray(spacing, np.float64)
    m = g2w / sp[:, None]
    v = (np.asarray(iso_world, np.float64) - np.asarray(origin_world, np.float64)) / sp
    return m.astype(np.float32), v.astype(np.float32)


def steps_and_start(m: np.ndarray, v: np.ndarray, ct_dims, dz: float = 1.0, margin: float = 2.0):
    """Tracer start plane (gantry z) and number of steps so the rays span the CT box along the beam axis
    (the reference hard-codes 512 steps from z=128/104, main.cu:62-64,192)."""
    minv = np.linalg.inv(m.astype(np.float64))
    zs = []
    for k in (0, 1):
        for j in (0, 1):
            for i in (0, 1):
                idx = np.array([i * ct_dims[0] - 0.5, j * ct_dims[1] - 0.5, k * ct_dims[2] - 0.5])
                zs.append((minv @ (idx - v.astype(np.float64)))[2])
    zmax = float(np.ceil(max(zs) + margin))
    zmin = float(np.floor(min(zs) - margin))
    return zmax, int(np.ceil((zmax - zmin) / dz))


def ellipsoid_mask(shape, centre, radii, spacing):
    nz, ny, nx = shape
    z = (np.arange(nz, dtype=np.float32) - centre[2]) * spacing[2] / radii[2]
    y = (np.arange(ny, dtype=np.float32) - centre[1]) * spacing[1] / radii[1]
    x = (np.arange(nx, dtype=np.float32) - centre[0]) * spacing[0] / radii[0]
    return (z[:, None, None] ** 2 + y[None, :, None] ** 2 + x[None, None, :] ** 2) <= 1.0


def lut_energies(luts: EnergyStruct, e_lo: float, e_hi: float, n: int) -> np.ndarray:
    return np.linspace(e_lo, e_hi, n).astype(np.float32)


def empirical_sigma(luts: EnergyStruct, energies: np.ndarray) -> np.ndarray:
    """sigma = 2.3 + 290/(peak+15) (main.cu:93-94), float32 arithmetic."""
    s = np.array([np.float32(2.3) + np.float32(290.0) / (np.float32(peak_depth(luts, e)) + np.float32(15.0)) for e in energies], np.float32)
    return np.stack([s, s], axis=1)


# ----------------------------------------------------------------------------------------------------
# configurations
def case_c1a(luts: EnergyStruct = None) -> Case:
    """WATER_CUBE_TEST (main.cu:39-43, 61-62, 74-99): 256^3 water, one parallel beam, 33x33x20 spots.
    The reference draws weights with libc rand(); here the documented LCG supplies 90+10u."""
    luts = luts or load_packaged_luts(water_cube_test=True)
    ct = np.full((256, 256, 256), 1000.0, np.float32)
    L = 20
    energies = (np.float32(118.12) + np.arange(L, dtype=np.float32) * np.float32((172.51 - 118.12) / (L - 1))).astype(np.float32)
    w = (90.0 + 10.0 * Lcg().uniform(33 * 33 * L)).astype(np.float32).reshape(L, 33, 33)
    b = Beam(w, energies, empirical_sigma(luts, energies), (1.0, 1.0), 512, (np.inf, np.inf), (3.0, 3.0, -1.0),
             (-48.0, -48.0, 128.0), np.eye(3, dtype=np.float32), np.array([128.0, 128.0, 106.0], np.float32))
    return Case("c1a_water_cube_test", ct, [b], luts)


def case_mini(luts: EnergyStruct = None) -> Case:
    """Small heterogeneous, divergent, slightly oblique case used for golden fixtures: 64x64x80 CT,
    32x32 rays (1024 = the smallest ray grid the reference's reductions handle, kernel_wrapper.cuh:180),
    4 layers, bone/lung/air inserts so the stale-dose (H9) and multi-radius paths are exercised."""
    luts = luts or load_packaged_luts()
    nx, ny, nz = 64, 64, 80
    ct = np.full((nz, ny, nx), 1000.0, np.float32)
    ct[:, :, :4] = 0.0; ct[:, :, -4:] = 0.0; ct[:, :4, :] = 0.0; ct[:, -4:, :] = 0.0; ct[-6:, :, :] = 0.0
    ct[50:58, :, :32] = 1900.0           # half-width bone slab
    ct[30:44, 20:44, 20:44] = 260.0      # lung block
    ct[60:66, 28:36, 28:36] = 0.0        # air pocket
    rng = Lcg(SEED + 7)
    ct += (rng.uniform(ct.size).reshape(ct.shape) - 0.5) * 20.0 * (ct > 100)
    energies = np.array([70.31, 78.0, 86.05, 93.9], np.float32)
    w = (60.0 + 40.0 * rng.uniform(4 * 5 * 5)).astype(np.float32).reshape(4, 5, 5)
    w[1, 0, :] = 0.0
    sig = np.array([[3.1, 3.3], [3.0, 3.2], [2.9, 3.0], [2.8, 2.9]], np.float32)
    t = np.deg2rad(4.0)
    m = np.array([[np.cos(t), 0.0, np.sin(t)], [0.0, 1.0, 0.0], [-np.sin(t), 0.0, np.cos(t)]], np.float32)
    v = np.array([31.5, 31.5, 40.0], np.float32)
    b = Beam(w, energies, sig, (1.0, 1.0), 96, (2000.0, 1800.0), (4.0, 4.0, -1.0), (-8.0, -8.0, 44.0), m, v)
    return Case("mini", ct, [b], luts)


def case_c2(luts: EnergyStruct = None) -> Case:
    """Heterogeneous slab phantom, single divergent beam, full energy-layer stack."""
    luts = luts or load_packaged_luts()
    n = 256
    ct = np.full((n, n, n), 1000.0, np.float32)
    ct[:, :, :6] = 0; ct[:, :, -6:] = 0; ct[:, :6, :] = 0; ct[:, -6:, :] = 0; ct[-6:, :, :] = 0; ct[:6, :, :] = 0
    # beam enters at high z and travels to low z; slabs perpendicular to z
    ct[200:220, 6:-6, 6:-6] = 2000.0                 # bone, 20 mm
    ct[140:180, 6:-6, 6:-6] = 250.0                  # lung, 40 mm
    ct[100:120, 6:-6, 6:128] = 1800.0                # half-width bone
    yy, zz = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    cyl = ((yy - 128) ** 2 + (zz - 232) ** 2) <= 100  # 20 mm air cylinder along x
    ct[cyl.T[:, :, None].repeat(n, 2) & (np.arange(n)[None, None, :] >= 6) & (np.arange(n)[None, None, :] < n - 6)] = 0.0
    idx = np.arange(0, luts.nEnergies, 2)
    energies = luts.energiesPerU[idx]
    energies = energies[(energies >= 70.0) & (energies <= 200.0)].astype(np.float32)
    L = energies.size
    rng = Lcg()
    w = (50.0 + 50.0 * rng.uniform(L * 21 * 21)).astype(np.float32).reshape(L, 21, 21)
    b = Beam(w, energies, empirical_sigma(luts, energies), (1.0, 1.0), 300, (2000.0, 1800.0), (4.0, 4.0, -1.0),
             (-40.0, -40.0, 130.0), np.eye(3, dtype=np.float32), np.array([127.5, 127.5, 127.5], np.float32))
    return Case("c2_slab", ct, [b], luts)


def _head_phantom(nx, ny, nz, spacing, seed):
    shape = (nz, ny, nx)
    c = (nx / 2 - 0.5, ny / 2 - 0.5, nz / 2 - 0.5)
    ct = np.zeros(shape, np.float32)
    outer = ellipsoid_mask(shape, c, (95.0, 115.0, 0.42 * nz * spacing[2]), spacing)
    inner = ellipsoid_mask(shape, c, (89.0, 109.0, 0.42 * nz * spacing[2] - 6.0), spacing)
    ct[outer] = 1900.0
    ct[inner] = 1040.0
    sinus = ellipsoid_mask(shape, (c[0], c[1] - 60.0 / spacing[1], c[2]), (18.0, 14.0, 25.0), spacing)
    ct[sinus & inner] = 0.0
    rng = np.random.default_rng(seed)
    noise = rng.integers(-15, 16, size=shape).astype(np.float32)
    ct += noise * (ct > 500)
    return ct


def _spot_grid_beam(luts, ct_dims, spacing, angle, n_side, pitch, energies, mask_ellipse, rng, w_lo, w_hi,
                    source_dist=(2000.0, 1800.0), iso_idx=None):
    nx, ny, nz = ct_dims
    iso_idx = iso_idx if iso_idx is not None else np.array([nx / 2 - 0.5, ny / 2 - 0.5, nz / 2 - 0.5])
    m, v = gantry_to_im_idx(angle, iso_idx * np.asarray(spacing), (0.0, 0.0, 0.0), spacing)
    zstart, steps = steps_and_start(m, v, ct_dims)
    L = len(energies)
    w = (w_lo + (w_hi - w_lo) * rng.uniform(L * n_side * n_side)).astype(np.float32).reshape(L, n_side, n_side)
    if mask_ellipse:
        g = (np.arange(n_side) - (n_side - 1) / 2) / ((n_side - 1) / 2 + 0.5)
        w *= ((g[:, None] ** 2 + g[None, :] ** 2) <= 1.0)[None].astype(np.float32)
    off = -(n_side - 1) / 2 * pitch
    return Beam(w, np.asarray(energies, np.float32), empirical_sigma(luts, energies), (1.0, 1.0), steps, source_dist,
                (pitch, pitch, -1.0), (off, off, zstart), m, v)


def case_c3(luts: EnergyStruct = None) -> Case:
    """512x512x200 head-and-neck-size CT @1x1x2 mm, 3 fields (0, 70, 290 deg), 15 layers x 15x15 spots."""
    luts = luts or load_packaged_luts()
    dims, sp = (512, 512, 200), (1.0, 1.0, 2.0)
    ct = _head_phantom(*dims, sp, SEED)
    rng = Lcg()
    energies = lut_energies(luts, 100.0, 160.0, 15)
    beams = [_spot_grid_beam(luts, dims, sp, a, 15, 5.0, energies, True, rng, 50.0, 100.0) for a in (0.0, 70.0, 290.0)]
    return Case("c3_head_3field", ct, beams, luts, sp)


def case_c4(luts: EnergyStruct = None) -> Case:
    """512x512x300 pelvis-size CT @1x1x1.5 mm, 5 beams, 20 layers x 25x25 spots @5 mm."""
    luts = luts or load_packaged_luts()
    dims, sp = (512, 512, 300), (1.0, 1.0, 1.5)
    nx, ny, nz = dims
    shape = (nz, ny, nx)
    c = (nx / 2 - 0.5, ny / 2 - 0.5, nz / 2 - 0.5)
    ct = np.zeros(shape, np.float32)
    ct[ellipsoid_mask(shape, c, (190.0, 130.0, 0.48 * nz * sp[2]), sp)] = 1030.0
    ct[ellipsoid_mask(shape, c, (150.0, 100.0, 0.30 * nz * sp[2]), sp) & ~ellipsoid_mask(shape, c, (135.0, 88.0, 0.30 * nz * sp[2]), sp)] = 1700.0
    for sx in (-1, 1):
        ct[ellipsoid_mask(shape, (c[0] + sx * 95.0, c[1], c[2]), (26.0, 26.0, 26.0), sp)] = 1850.0
    ct[ellipsoid_mask(shape, (c[0] + 20, c[1] - 30, c[2] + 10), (22.0, 15.0, 30.0), sp)] = 60.0
    rng = Lcg()
    energies = lut_energies(luts, 150.0, 220.0, 20)
    beams = [_spot_grid_beam(luts, dims, sp, a, 25, 5.0, energies, True, rng, 40.0, 100.0) for a in (0.0, 72.0, 144.0, 216.0, 288.0)]
    return Case("c4_pelvis_5beam", ct, beams, luts, sp)


def case_c5(luts: EnergyStruct = None) -> Case:
    """High-resolution 1 mm CT (512x512x400), 9 beams every 40 deg, 25 layers x 41x41 spots @2.5 mm."""
    luts = luts or load_packaged_luts()
    dims, sp = (512, 512, 400), (1.0, 1.0, 1.0)
    ct = _head_phantom(*dims, sp, SEED + 5)
    rng = Lcg()
    energies = lut_energies(luts, 90.0, 170.0, 25)
    beams = [_spot_grid_beam(luts, dims, sp, 40.0 * i, 41, 2.5, energies, True, rng, 30.0, 60.0) for i in range(9)]
    return Case("c5_hires_9beam", ct, beams, luts, sp)


def case_c1b(luts: EnergyStruct = None) -> Case:
    """README water cube (generate_water_cube.py:96-105, 328-363): 512^3 @1 mm, 6-px air rim, one beam G000,
    29 layers 90..230 MeV, one spot (0,0) of 1e9 per layer presented as the centre of a 3x3 grid @5 mm
    (keeps the ray grid at the reference-safe 1024 rays); sigma = 6 mm FWHM / 2.355."""
    luts = luts or load_packaged_luts()
    n = 512
    ct = np.zeros((n, n, n), np.float32)
    ct[6:-6, 6:-6, 6:-6] = 1000.0
    energies = np.arange(90.0, 231.0, 5.0, dtype=np.float32)
    L = energies.size
    w = np.zeros((L, 3, 3), np.float32); w[:, 1, 1] = 1.0e9
    sig = np.full((L, 2), np.float32(6.0 / 2.355), np.float32)
    m, v = gantry_to_im_idx(0.0, (255.5, 255.5, 255.5), (0.0, 0.0, 0.0), (1.0, 1.0, 1.0))
    zstart, steps = steps_and_start(m, v, (n, n, n))
    b = Beam(w, energies, sig, (1.0, 1.0), steps, (2000.0, 1800.0), (5.0, 5.0, -1.0), (-5.0, -5.0, zstart), m, v)
    return Case("c1b_readme_water_cube", ct, [b], luts)


CASES = {"c1a": case_c1a, "c1b": case_c1b, "mini": case_mini, "c2": case_c2, "c3": case_c3, "c4": case_c4, "c5": case_c5}


def make_case(name: str) -> Case:
    return CASES[name]()
