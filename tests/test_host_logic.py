"""Host-side pieces that need no GPU: LUT loader, case generators, LCG, work partition."""
import os

import numpy as np
import pytest

from raytracedicom_b200 import cases, sharding
from raytracedicom_b200.energy_reader import load_packaged_luts


def test_packaged_luts_shapes():
    """Shapes and anchor values of the reference's LUT files (energy_reader.cpp:11-101)."""
    e = load_packaged_luts()
    assert (e.nEnergies, e.nEnergySamples) == (147, 1024)
    assert e.ciddMatrix.size == 147 * 1024
    assert e.nDensitySamples == e.nSpSamples == e.nRRlSamples == 3072
    assert e.densityVector[1000] == pytest.approx(1.0) and e.spVector[1000] == pytest.approx(1.0)
    assert np.all(np.diff(e.energiesPerU) > 0)
    assert e.energiesPerU[0] == pytest.approx(62.39, abs=0.01) and e.energiesPerU[-1] == pytest.approx(226.64, abs=0.01)
    w = load_packaged_luts(water_cube_test=True)
    assert int(np.sum(w.rRlVector != e.rRlVector)) == 1


@pytest.mark.skipif(not os.path.isdir("/root/reference/LUTs"), reason="reference text LUTs only exist in the build container")
def test_energy_reader_matches_packaged_fixture():
    from raytracedicom_b200.energy_reader import energy_reader
    e = energy_reader("/root/reference/LUTs/")
    p = load_packaged_luts()
    for k in ("energiesPerU", "peakDepths", "scaleFacts", "densityVector", "spVector", "rRlVector"):
        assert np.array_equal(getattr(e, k), getattr(p, k)), k
    assert np.array_equal(np.ravel(e.ciddMatrix), np.ravel(p.ciddMatrix))


def test_lcg_is_the_documented_generator():
    g = cases.Lcg(20261104)
    x = 20261104
    want = []
    for _ in range(4):
        x = (x * 1664525 + 1013904223) % 2 ** 32
        want.append(np.float32((x >> 8) / 2 ** 24))
    assert np.array_equal(g.uniform(4), np.array(want, np.float32))


def test_find_and_interpolate_match_oracle(oracle):
    import ctypes as C
    e = load_packaged_luts()
    l = oracle.lib()
    en = np.ascontiguousarray(e.energiesPerU, np.float32)
    pk = np.ascontiguousarray(e.peakDepths, np.float32)
    fp = C.POINTER(C.c_float)
    for v in (50.0, 62.39, 100.0, 118.12, 172.51, 226.0, 230.0):
        a = cases.find_decimal_ordered(en, v)
        b = l.orc_find_decimal_ordered(en.ctypes.data_as(fp), len(en), C.c_float(v))
        assert np.float32(a) == np.float32(b)
        assert np.float32(cases.vector_interpolate(pk, a)) == np.float32(l.orc_vector_interpolate(pk.ctypes.data_as(fp), len(pk), C.c_float(b)))


def test_c1a_matches_water_cube_test_constants():
    """main.cu:39-43, 61-62, 74-99."""
    c = cases.make_case("c1a")
    b = c.beams[0]
    assert c.ct.shape == (256, 256, 256) and float(c.ct.min()) == 1000.0
    assert b.weights.shape == (20, 33, 33) and b.steps == 512
    assert b.energies[0] == pytest.approx(118.12) and b.energies[-1] == pytest.approx(172.51, abs=1e-3)
    assert tuple(b.spot_delta) == (3.0, 3.0, -1.0) and tuple(b.spot_offset) == (-48.0, -48.0, 128.0)
    assert 90.0 <= b.weights.min() and b.weights.max() <= 100.0


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_partition_covers_every_layer_once(world):
    luts = load_packaged_luts()
    rng = cases.Lcg()
    en = cases.lut_energies(luts, 100.0, 160.0, 15)
    beams = [cases.Beam(rng.uniform(15 * 9).reshape(15, 3, 3), en, cases.empirical_sigma(luts, en)) for _ in range(3)]
    parts = sharding.partition(beams, luts, world)
    assert len(parts) == world
    seen = np.zeros((3, 15), int)
    for items in parts:
        for b, lo, hi in items:
            assert 0 <= lo < hi <= 15
            seen[b, lo:hi] += 1
    assert np.all(seen == 1)
    if world > 1:
        cost = [sum(sharding.layer_cost(beams[b], luts)[lo:hi].sum() for b, lo, hi in items) for items in parts]
        assert max(cost) <= 1.6 * (sum(cost) / world)        # LPT keeps the ranks balanced


def test_shard_keeps_whole_beam_geometry():
    c = cases.make_case("mini")
    s = c.beams[0].shard(1, 3)
    assert s.weights is c.beams[0].weights or np.array_equal(s.weights, c.beams[0].weights)
    assert (s.layer_begin, s.layer_end) == (1, 3) and s.n_layers == c.beams[0].n_layers


@pytest.mark.skipif(not os.path.isdir("/root/reference/LUTs"), reason="needs the reference's LUT text files (build container only)")
def test_energy_reader_loads_the_nuclear_luts_like_the_reference():
    """energy_reader.cpp:103-162 (NUCLEAR_CORR builds): header checked against the IDD file, then two matrices of the
    IDD matrix's shape. The dose engine does not consume them: the reference's nuclear path is not functional as shipped
    (kernel_wrapper.cu:925 passes nucMemStep = 0)."""
    from raytracedicom_b200 import energy_reader as er
    base = er.energy_reader("/root/reference/LUTs/")
    for name, model in er.NUCLEAR_MODELS.items():
        e = er.energy_reader("/root/reference/LUTs/", nuclear_corr=model)
        assert e.nuclearCorr == model and e.nucWeightMatrix.shape == base.ciddMatrix.shape == e.nucSqSigmaMatrix.shape
        assert -0.05 < e.nucWeightMatrix.min() and e.nucWeightMatrix.max() < 1.0 and np.isfinite(e.nucSqSigmaMatrix).all()
        assert np.array_equal(e.ciddMatrix, base.ciddMatrix)


def test_partition_properties_on_random_plans():
    """Random plans (1-5 beams, 1-24 layers each) on 1-16 ranks, both partition rules: every (beam, layer) exactly once,
    items ordered and non-empty; the contiguous rule creates at most world - 1 extra beam pieces and its makespan stays
    within one largest item (layer + piece cost) of the average load."""
    luts = load_packaged_luts()
    rng = np.random.default_rng(7)
    for _ in range(40):
        beams = []
        for _b in range(int(rng.integers(1, 6))):
            lo = float(rng.uniform(70, 150))
            en = cases.lut_energies(luts, lo, lo + float(rng.uniform(1, 60)), int(rng.integers(1, 25)))
            beams.append(cases.Beam(rng.uniform(size=(len(en), 3, 3)).astype(np.float32), en, cases.empirical_sigma(luts, en)))
        lc = [sharding.layer_cost(b, luts) for b in beams]
        pc = [sharding.piece_cost(b) for b in beams]
        for world in (1, 2, 3, 5, 8, 16):
            for chunks in (0, 2):
                parts = sharding.partition(beams, luts, world, chunks)
                assert len(parts) == world
                seen = [np.zeros(b.n_layers, int) for b in beams]
                for items in parts:
                    assert items == sorted(items)
                    for b, l0, l1 in items:
                        assert 0 <= l0 < l1 <= beams[b].n_layers
                        seen[b][l0:l1] += 1
                assert all(np.all(s == 1) for s in seen)
                if chunks == 0 and world > 1:
                    assert sum(len(items) for items in parts) <= len(beams) + world - 1
                    loads = [sum(lc[b][l0:l1].sum() + pc[b] for b, l0, l1 in items) for items in parts]
                    average = (sum(c.sum() for c in lc) + sum(pc)) / world
                    assert max(loads) <= average + 1.01 * max(c.max() + p for c, p in zip(lc, pc))
