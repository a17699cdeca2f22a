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
