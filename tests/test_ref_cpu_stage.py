"""The oracle's pre-convolution stage against the reference's OWN CPU code for it (/root/reference/src/cpu_convolution_1d.cpp,
compiled unmodified into oracle/_ref/libref_cpuconv.so by oracle/ref_harness/Makefile): a second, GPU-free pin of the oracle
(the first one is the rebuilt reference run on the GPU box, tests/golden). Also covers the guard around the reference code's
exit(1) / unchecked writes."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_cpu_stage as rcs  # noqa: E402
from raytracedicom_b200 import cases  # noqa: E402

pytestmark = pytest.mark.skipif(not rcs.available(), reason="oracle/_ref/libref_cpuconv.so not built (needs /root/reference at build time)")


@pytest.mark.parametrize("name", ["mini", "c1a"])
def test_oracle_preconvolution_matches_reference_cpu_code(name):
    case = cases.make_case(name)
    p = rcs.k3a_problem(case, 0)
    out, _ = rcs.run(p, threads=1)
    want = p["oracle_ray_weights"]
    assert out.shape == want.shape
    # same tap set (|i| <= ceil(3 sigma + 1/2) - 1  <=>  |distance| < 3 sigma + 1/2); differences: std::erf vs erff, x/y
    # summation order
    assert float(np.abs(out - want).max()) <= 2e-6 * float(want.max())
    # a Gaussian truncated at >= 3 sigma per axis (the last tap reaches up to one ray further) keeps between
    # erf(3 / sqrt 2)^2 and all of the spot weight
    kept = float(out.sum()) / float(case.beams[0].weights.sum())
    assert 0.9973 ** 2 - 1e-4 <= kept <= 1.0 + 1e-5
    out_mt, _ = rcs.run(p, threads=0)
    assert np.array_equal(out, out_mt)          # layers are independent: the OpenMP wrapper changes nothing


def test_guard_reports_taps_outside_the_ray_grid_instead_of_exiting():
    p = rcs.k3a_problem(cases.make_case("mini"), 0)
    p["radX"] = p["radX"].copy(); p["radX"][1] = p["offX"] + 1
    with pytest.raises(RuntimeError, match="layer 1"):
        rcs.run(p)
