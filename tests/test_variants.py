"""The reference's build variants NOZZLE=OFF and DOSE_TO_WATER=OFF (CMakeLists.txt:40-51) in the CPU oracle: known answers
that need no reference run (the GPU engine is checked against the oracle and the rebuilt variants in test_gpu_parity.py)."""
import numpy as np

from raytracedicom_b200 import cases


def _water_case():
    luts = cases.load_packaged_luts()
    ct = np.full((48, 64, 64), 1000.0, np.float32)
    w = np.full((2, 3, 3), 50.0, np.float32)
    b = cases.Beam(w, np.array([70.0, 75.0], np.float32), np.full((2, 2), 3.0, np.float32), (1.0, 1.0), 60, (2000.0, 1800.0),
                   (4.0, 4.0, -1.0), (-4.0, -4.0, 30.0), np.eye(3, dtype=np.float32), np.array([32.0, 32.0, 24.0], np.float32))
    return cases.Case("water", ct, [b], luts)


def test_no_nozzle_leaves_the_spot_sigma_unchanged_at_entry(oracle):
    c = _water_case()
    on = oracle.run_beam(c, 0)
    off = oracle.run_beam(c, 0, opts=oracle.default_opts(noNozzle=1))
    # sigmaSqAirCoefs = (0, 0) (fill_idd_and_sigma_params.cu:80-81) -> entrySigma = sqrt(sigma^2) (kernel_wrapper.cu:838-841)
    assert np.array_equal(off["layerF"][:, 3:5], np.full((2, 2), 3.0, np.float32))
    assert np.all(on["layerF"][:, 3:5] != 3.0)
    assert off["updates"] == on["updates"]
    assert np.abs(off["dose"] - on["dose"]).max() > 1e-3 * on["dose"].max()


def test_dose_to_medium_equals_dose_to_water_in_water(oracle):
    """mass = density * stepVol (kernel_wrapper.cu:317) against (wepl - weplOld) * stepVol (cu:315): identical where the
    relative stopping power and the density are both 1 and the ray is parallel to the axis; rays off axis are longer by
    stepLen = sqrt(1 + dx^2 + dy^2) (density_and_sp_tracer_params.cu:32-36), which only dose-to-water sees."""
    c = _water_case()
    wat = oracle.run_beam(c, 0)
    med = oracle.run_beam(c, 0, opts=oracle.default_opts(doseToMedium=1))
    assert med["updates"] == wat["updates"]
    ratio = med["dose"].sum() / wat["dose"].sum()
    assert 1.0 <= ratio < 1.0 + 1e-3
    # in lung (density 0.26, stopping power ~0.26 as well) the two differ by construction only through rho/SP
    c.ct[:, :, :] = 250.0
    wat = oracle.run_beam(c, 0); med = oracle.run_beam(c, 0, opts=oracle.default_opts(doseToMedium=1))
    sp = float(c.luts.spVector[250]); rho = float(c.luts.densityVector[250])
    if wat["dose"].max() > 0:
        assert np.isclose(med["dose"].sum() / wat["dose"].sum(), sp / rho, rtol=2e-2)


def test_nuclear_corr_switch_in_the_oracle(oracle):
    """NUCLEAR_CORR as shipped (kernel_wrapper.cu:228-245, 300-302, 320-341, 842-847, 925): the integer outputs of a beam that starts
    outside the patient are those of the default build where the scattering constants do not move a cut-off (the dose cut-offs
    depend on WEPL only), the dose is the primary dose scaled by 1 - nucWeight (< 1) with a changed lateral spread; the three
    models differ; the option without its LUT, and a beam whose first step is inside the patient, are refused."""
    from raytracedicom_b200.energy_reader import NUCLEAR_MODELS, load_packaged_luts
    c = cases.make_case("mini")
    base = oracle.run_beam(c, 0)
    assert oracle.run_beam(c, 0, opts=oracle.default_opts(nuclearCorr=2))["rc"] == 1      # no LUT
    sums = {}
    for name, m in NUCLEAR_MODELS.items():
        luts = load_packaged_luts(nuclear_corr=m)
        assert luts.nucWeightMatrix.shape == luts.ciddMatrix.shape and 0.0 < float(luts.nucWeightMatrix.max()) < 0.3
        o = oracle.run_beam(cases.Case("n", c.ct, c.beams, luts), 0, opts=oracle.default_opts(nuclearCorr=m))
        assert o["rc"] == 0 and o["updates"] == base["updates"] and o["beamFirstCalculatedPassive"] == base["beamFirstCalculatedPassive"]
        if m != 3:      # GAUSS_FIT's 0.97 entry sigma changes the ray weights, hence which rays pass RAY_WEIGHT_CUTOFF
            assert np.array_equal(o["first_passive"], base["first_passive"])
        sums[name] = float(o["dose"].sum(dtype=np.float64))
        assert 0.75 * base["dose"].sum(dtype=np.float64) < sums[name] < base["dose"].sum(dtype=np.float64)
    assert len({round(v / sums["soukup"], 4) for v in sums.values()}) == 3
    # a beam that starts inside the patient: the reference's slice-0 halo residue is not restated
    w = _water_case()
    b = w.beams[0]
    inside = cases.Beam(b.weights, b.energies, b.sigmas, b.ray_spacing, b.steps, b.source_dist, b.spot_delta, (b.spot_offset[0], b.spot_offset[1], 20.0),
                        b.gantry_to_im_m, b.gantry_to_im_v)
    o = oracle.run_beam(cases.Case("inside", w.ct, [inside], load_packaged_luts(nuclear_corr=1)), 0, opts=oracle.default_opts(nuclearCorr=1))
    assert o["beamFirstInside"] == 0 and o["rc"] == 4
