"""CPU property test of the CT sub-volume logic of rtd_plan_dose (csrc/rtdose_api.cu: ctBoxesOfBeam, depth segments and
disjoint axis slabs): for random beam geometries every texel the tracer can fetch lies inside one of the boxes. The tracer's
sample positions are restated here exactly as the kernel forms them (density_and_sp_tracer_params.cu:15-17: start + 1/2,
then repeated fp32 addition of the increment, kernel_wrapper.cu:183); a linear fetch at p reads texels floor(p - 1/2) and
the one after it on every axis. The GPU test with a NaN-poisoned array covers three fixed cases; this one covers the
geometry space (gantry / couch angles, divergence, offsets, anisotropic voxels) without a device."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from raytracedicom_b200 import cases, engine  # noqa: E402
from raytracedicom_b200.plan import gantry_to_patient  # noqa: E402


def ray_grid(b):
    L, sy, sx = b.weights.shape
    res = np.array(b.ray_spacing, np.float32)
    smax = np.max(np.asarray(b.sigmas, np.float32), axis=0)
    off = np.array(b.spot_offset[:2], np.float32); d = np.array(b.spot_delta[:2], np.float32)
    n = np.array([sx, sy], np.float32)
    half = np.float32(0.5)
    lo = np.ceil((off - (np.float32(3.0) * smax + half * res)) / res).astype(int)            # kernel_wrapper.cu:650-653
    hi = np.floor(((n - 1) * d + off + (np.float32(3.0) * smax + half * res)) / res).astype(int)
    dim = hi - lo + 1
    dim[0] = (dim[0] + 31) // 32 * 32; dim[1] = (dim[1] + 7) // 8 * 8
    return lo, dim, res


def sample_texels(b, stride=1):
    """floor(p - 1/2) of every tracer sample, [n][3] int (x, y, z), positions accumulated in fp32 like the kernel."""
    lo, dim, res = ray_grid(b)
    f = np.float32
    m = np.asarray(b.gantry_to_im_m, f); v = np.asarray(b.gantry_to_im_v, f)
    sd = np.array(b.source_dist, f)
    cz = f(b.spot_offset[2]); dz = f(b.spot_delta[2])
    ii, jj = np.meshgrid(np.arange(0, dim[0], stride), np.arange(0, dim[1], stride), indexing="ij")
    # include the last ray of each axis (the frustum's faces)
    ii = np.concatenate([ii.ravel(), np.full(dim[1], dim[0] - 1), np.arange(dim[0])]); jj = np.concatenate([jj.ravel(), np.arange(dim[1]), np.full(dim[0], dim[1] - 1)])
    gx = (lo[0] + ii).astype(f) * res[0]; gy = (lo[1] + jj).astype(f) * res[1]
    s1 = f(1) - cz / sd[0]; s2 = f(1) - cz / sd[1]
    start = (np.outer(gx * s1, m[:, 0]) + np.outer(gy * s2, m[:, 1]) + cz * m[:, 2] + v).astype(f) + f(0.5)
    inc = ((m[:, 2][None, :] - np.outer(gx / sd[0], m[:, 0]) - np.outer(gy / sd[1], m[:, 1])) * dz).astype(f)
    out = []
    p = start.copy()
    for _ in range(int(b.steps)):
        out.append(np.floor(p - f(0.5)).astype(np.int64))
        p = (p + inc).astype(f)
    return np.concatenate(out)


def covered(tex, boxes, dims):
    nx, ny, nz = dims
    need = np.concatenate([tex + np.array(o) for o in ((0, 0, 0), (1, 0, 0), (0, 1, 0), (1, 1, 0), (0, 0, 1), (1, 0, 1), (0, 1, 1), (1, 1, 1))])
    inside = ((need >= 0) & (need < np.array([nx, ny, nz]))).all(1)
    need = np.unique(need[inside], axis=0)
    ok = np.zeros(len(need), bool)
    for bx in boxes:
        ok |= ((need >= bx[:3]) & (need <= bx[3:])).all(1)
    return need, ok


def make_beam(rng, dims, spacing):
    nx, ny, nz = dims
    L = 3
    sxy = int(rng.integers(3, 9))
    pitch = float(rng.choice([3.0, 5.0, 8.0]))
    w = np.ones((L, sxy, sxy), np.float32)
    sig = np.full((L, 2), float(rng.uniform(2.0, 6.0)), np.float32)
    gantry, couch = float(rng.uniform(0, 360)), float(rng.choice([0.0, 0.0, 15.0, 340.0]))
    rot = gantry_to_patient(gantry, couch)
    a = np.diag(spacing).astype(np.float64)
    iso = np.array([nx * spacing[0] * rng.uniform(0.35, 0.65), ny * spacing[1] * rng.uniform(0.35, 0.65), nz * spacing[2] * rng.uniform(0.35, 0.65)])
    m = (np.linalg.inv(a) @ rot).astype(np.float32)
    v = (np.linalg.inv(a) @ iso).astype(np.float32)
    sad = float(rng.choice([np.inf, 2000.0, 800.0])); sad2 = sad if rng.random() < 0.5 else float(rng.choice([np.inf, 2500.0, 900.0]))
    z0, steps = cases.steps_and_start(m, v, dims, dz=1.0)
    off = -(sxy - 1) * pitch / 2 + float(rng.uniform(-20, 20))
    return cases.Beam(w, np.full(L, 120.0, np.float32), sig, (1.0, 1.0), steps, (sad, sad2), (pitch, pitch, -1.0), (off, off * 0.5, z0), m, v)


@pytest.mark.parametrize("seed", range(12))
def test_boxes_cover_every_texel_the_tracer_can_fetch(seed):
    rng = np.random.default_rng(100 + seed)
    dims = (int(rng.choice([96, 128, 160])), int(rng.choice([96, 128])), int(rng.choice([40, 64, 100])))
    spacing = (float(rng.choice([1.0, 1.5, 2.0])),) * 2 + (float(rng.choice([1.0, 2.0, 3.0])),)
    b = make_beam(rng, dims, spacing)
    boxes = engine.ct_boxes(b, dims)
    assert len(boxes) >= 1 and (boxes[:, :3] >= 0).all() and (boxes[:, 3:] < np.array(dims)).all() and (boxes[:, :3] <= boxes[:, 3:]).all()
    tex = sample_texels(b, stride=3)
    need, ok = covered(tex, boxes, dims)
    assert len(need) > 0
    assert ok.all(), f"{(~ok).sum()} of {len(need)} texels outside the boxes, e.g. {need[~ok][:3]}"
    # and the boxes are a real sub-volume unless the beam fills the CT
    vox = int(np.prod(boxes[:, 3:] - boxes[:, :3] + 1, axis=1).sum())
    assert vox <= 1.6 * np.prod(dims)


def test_c3_boxes_are_well_below_the_whole_ct():
    c = cases.make_case("c3")
    nz, ny, nx = c.ct.shape
    tot = 0
    for b in c.beams:
        bx = engine.ct_boxes(b, (nx, ny, nz))
        tot += int(np.prod(bx[:, 3:] - bx[:, :3] + 1, axis=1).sum())
    assert tot * 4 < 0.32 * c.ct.nbytes * 1.0        # 60 MB of 210 MB (depth segments alone: 80 MB)
