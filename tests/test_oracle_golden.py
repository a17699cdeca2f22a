"""The CPU oracle (oracle/rtd_oracle.c) against the reference's own outputs (tests/golden/golden_mini.npz: dumps of
the reference's kernels rebuilt for sm_100 and run on a B200) and against reference-free known answers
(SURVEY.md 8c).  Runs without a GPU."""
import numpy as np
import pytest

from conftest import rel_to_max
from raytracedicom_b200 import cases

# Float tolerances (relative to the array's max): the reference is compiled with -use_fast_math (approximate
# division, sqrt, exp2/log2), a CPU cannot reproduce those bit-for-bit; observed differences are ~1e-7..1e-4.
TOL_TRACE = 1e-6      # density, WEPL
TOL_RAYW = 5e-6       # ray weights (erff)
TOL_IDD = 5e-4        # dose per BEV voxel of one layer (division by small voxel masses amplifies ulps)
TOL_RSIG = 1e-5
TOL_DOSE = 2e-5       # BEV dose and final dose vs the reference (BASELINE.json bound is 1e-3)


@pytest.fixture(scope="module")
def orun(oracle, mini_case, golden_mini):
    return oracle.run_beam(mini_case, 0, keep_layer=int(golden_mini["golden_layer"]))


def test_index_outputs_bit_exact(orun, golden_mini):
    g = golden_mini
    got = [orun[k] for k in ("lSteps", "bSteps", "rSteps", "tSteps", "Rx", "Ry", "beamFirstInside", "beamFirstOutside",
                             "beamFirstGuaranteedPassive", "beamFirstCalculatedPassive", "updates")]
    assert got == list(g["meta_ints"])
    assert orun["bboxMin"] + orun["bboxMax"] == list(g["bbox"])
    assert np.array_equal(orun["first_inside"], g["first_inside"])
    assert np.array_equal(orun["first_outside"], g["first_outside"])
    gl = int(g["golden_layer"])
    assert np.array_equal(orun["first_passive"][gl], g["first_passive_l"])
    for l in range(orun["L"]):
        assert np.array_equal(orun["hist"][l], g[f"hist_{l}"]), f"radius histogram of layer {l}"
        assert list(orun["layerI"][l]) == list(g[f"layer_ints_{l}"]), f"afterLast / layerFirstPassive of layer {l}"


def test_float_stages_within_tolerance(orun, golden_mini):
    g = golden_mini
    assert rel_to_max(orun["density"], g["density"]) <= (TOL_TRACE, 0)
    assert rel_to_max(orun["wepl"], g["wepl"]) <= (TOL_TRACE, 0)
    assert rel_to_max(orun["ray_weights"], g["ray_weights"]) <= (TOL_RAYW, 0)
    bfi, n = orun["beamFirstInside"], g["idd_l"].shape[0]
    e, nf = rel_to_max(orun["idd"][bfi:bfi + n], g["idd_l"]); assert e <= TOL_IDD and nf == 0
    e, nf = rel_to_max(orun["rsig"][bfi:bfi + n], g["rsig_l"]); assert e <= TOL_RSIG and nf == 0   # +inf pattern identical
    assert rel_to_max(orun["bev_dose"], g["bev_dose"]) <= (TOL_DOSE, 0)
    assert rel_to_max(orun["dose"], g["dose_beam0"]) <= (TOL_DOSE, 0)


def test_golden_discriminates_texture_model(oracle, mini_case, golden_mini):
    """A plain separable fp32 trilinear filter (what a naive port would use) must FAIL the pin: shows that the
    fixture is sensitive to the texture-unit arithmetic the reference's numerics depend on."""
    o = oracle.run_beam(mini_case, 0, opts=oracle.default_opts(texFracMode=2))
    e, _ = rel_to_max(o["dose"], golden_mini["dose_beam0"])
    assert e > 20 * TOL_DOSE


def _water_case(source_dist=(np.inf, np.inf), n=64, L=2):
    luts = cases.load_packaged_luts(water_cube_test=True)
    ct = np.full((n, n, n), 1000.0, np.float32)
    energies = np.array([70.31, 86.05][:L], np.float32)
    w = np.full((L, 3, 3), 100.0, np.float32)
    b = cases.Beam(w, energies, cases.empirical_sigma(luts, energies), (1.0, 1.0), 80, source_dist, (4.0, 4.0, -1.0),
                   (-4.0, -4.0, 36.0), np.eye(3, dtype=np.float32), np.array([31.5, 31.5, 31.5], np.float32))
    return cases.Case("water", ct, [b], luts)


def test_known_answer_water_wepl(oracle):
    """Homogeneous water, parallel beam, 1 mm steps: SP[1000] = rho[1000] = 1, so wepl[k] = (number of steps taken
    inside the cube up to k) and density = 1 inside."""
    c = _water_case()
    o = oracle.run_beam(c, 0, stop_after=1)
    cx, cy = o["Rx"] // 2, o["Ry"] // 2
    fi = o["first_inside"][cy, cx]; fo = o["first_outside"][cy, cx]
    assert fo - fi >= 60
    col = o["wepl"][:, cy, cx]
    inside = np.arange(fi + 1, fo - 1)
    assert np.allclose(np.diff(col[inside]), 1.0, atol=1e-5)
    assert np.allclose(o["density"][inside, cy, cx], 1.0, atol=1e-6)


def test_preconvolution_conserves_weight(oracle):
    """Interior spots: each spot spreads its weight over the ray pixels with erf-difference weights that sum to 1
    per axis, so sum(ray weights) = sum(spot weights) up to the 3-sigma truncation erf(3/sqrt2)^2."""
    c = _water_case()
    o = oracle.run_beam(c, 0, stop_after=2)
    b = c.beams[0]
    for l in range(b.n_layers):
        total = o["ray_weights"][l].sum(dtype=np.float64)
        assert abs(total / b.weights[l].sum() - 1.0) < 0.012


def test_superposition_conserves_idd(oracle):
    """Kernel superposition only redistributes dose laterally: per depth slice, sum(BEV dose) = sum over layers of
    sum(idd) up to the truncation of the Gaussian tails.  One layer -> compare directly."""
    c = _water_case(L=1)
    o = oracle.run_beam(c, 0, keep_layer=0, stop_after=3)
    bfi, fcp = o["beamFirstInside"], o["beamFirstCalculatedPassive"]
    for k in range(bfi + 2, fcp - 1, 7):
        s_in = o["idd"][k].sum(dtype=np.float64); s_out = o["bev_dose"][k].sum(dtype=np.float64)
        assert s_in > 0 and abs(s_out / s_in - 1.0) < 0.01


def test_bragg_peak_depth(oracle):
    """Depth of the dose maximum on the central axis within 2 steps of the LUT peak depth."""
    c = _water_case(L=1)
    o = oracle.run_beam(c, 0)
    peak = o["layerF"][0][2]
    col = o["dose"][:, 31, 31] + o["dose"][:, 32, 32]
    z_peak = int(np.argmax(col))
    # beam enters at the high-z face (gantry z = +32 -> index 63.5) and travels towards -z
    depth = 63.5 - z_peak
    assert abs(depth - peak) <= 2.0


def test_shards_add_up(oracle, mini_case):
    """Dose is linear in energy layers: layer shards of a beam sum to the whole beam (basis of the multi-GPU split)."""
    whole = oracle.run_beam(mini_case, 0)["dose"]
    acc = np.zeros_like(whole)
    b = mini_case.beams[0]
    for lo, hi in ((0, 1), (1, 3), (3, 4)):
        c2 = cases.Case("s", mini_case.ct, [b.shard(lo, hi)], mini_case.luts)
        oracle.run_beam(c2, 0, dose=acc)
    assert rel_to_max(acc, whole)[0] < 1e-6


def test_radius_over_32_is_an_error(oracle):
    """0.08 mm rays: 3 sigma exceeds 32 ray pitches near the peak -> the reference throws (kernel_wrapper.cu:965)."""
    c = _water_case(L=1)
    b = c.beams[0]
    b2 = cases.Beam(b.weights * 1.0e4, b.energies, b.sigmas, (0.08, 0.08), b.steps, b.source_dist, b.spot_delta, b.spot_offset,
                    b.gantry_to_im_m, b.gantry_to_im_v)
    o = oracle.run_beam(cases.Case("r", c.ct, [b2], c.luts), 0)
    assert o["rc"] == 3
