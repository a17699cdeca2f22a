"""The C++ host driver (raytracedicom_b200/host: Config, energyReader, BeamSettings, main) over the C ABI, and the
C++ shim (include/rtdose_shim.hpp) against the reference's own value types."""
import os
import subprocess
import tempfile

import numpy as np
import pytest

from conftest import ROOT, rel_to_max
from raytracedicom_b200 import cases

EXE = os.path.join(ROOT, "raytracedicom_b200", "rtdose_main")


def _build():
    subprocess.run(["make", "-C", os.path.join(ROOT, "raytracedicom_b200", "host")], check=True, capture_output=True)


def test_driver_builds_and_has_reference_options():
    _build()
    r = subprocess.run([EXE, "--help"], capture_output=True, text=True)
    for opt in ("--gpu_id", "--ct_dir", "--rtplan", "--beams", "--output_directory", "--config_file"):   # src/config.cpp:13-51
        assert opt in r.stdout


def test_driver_fails_loudly_without_gpu_or_input(tmp_path):
    import torch
    _build()
    r = subprocess.run([EXE], capture_output=True, text=True)
    assert r.returncode == 1 and "no input" in r.stderr
    if not torch.cuda.is_available():
        c = cases.make_case("mini"); p = str(tmp_path / "mini.rtdcase"); c.write(p)
        r = subprocess.run([EXE, "--case", p, "--output_directory", str(tmp_path)], capture_output=True, text=True)
        assert r.returncode == 1 and "no CPU fallback" in r.stderr


@pytest.mark.skipif(not os.path.isdir("/root/reference/src"), reason="needs the reference's headers (build container only)")
def test_shim_compiles_against_reference_types(tmp_path):
    """rtdose::cudaWrapperProtons instantiated with the reference's HostPinnedImage3D / BeamSettings / EnergyStruct."""
    src = tmp_path / "shim_check.cu"
    src.write_text('#include <iostream>\n#include "beam_settings.h"\n#include "energy_struct.h"\n#include "rtdose_shim.hpp"\n'
                   'void call(HostPinnedImage3D<float>* im, HostPinnedImage3D<float>* dose, std::vector<BeamSettings> beams, EnergyStruct e) {\n'
                   '    rtdose::cudaWrapperProtons(im, dose, beams, e, std::cout);\n}\n')
    subprocess.run(["nvcc", "-std=c++14", "-c", str(src), "-o", str(tmp_path / "shim_check.o"), "-I/root/reference/src",
                    "-I" + os.path.join(ROOT, "oracle", "ref_harness"), "-I" + os.path.join(ROOT, "include"),
                    "-Wno-deprecated-declarations"], check=True, capture_output=True)


@pytest.mark.skipif(not os.path.isdir("/root/reference/LUTs"), reason="needs the reference's LUT text files (build container only)")
def test_driver_reads_the_nuclear_tables_like_the_reference(tmp_path):
    """--nuclear_corr MODEL (the reference's NUCLEAR_CORR build option at run time): the C++ energyReader loads the model's
    nuclear_weights_and_sigmas_*.txt with the header checks of energy_reader.cpp:103-162; the assembled plan (--dump_case) carries
    them in the case file's trailer, equal to what the Python reader produces."""
    import struct
    from raytracedicom_b200 import energy_reader as er
    _build()
    for name, model in er.NUCLEAR_MODELS.items():
        out = tmp_path / f"{name}.rtdcase"
        r = subprocess.run([EXE, "--water_cube_test", "--phys_data", "/root/reference/LUTs/", "--nuclear_corr", name, "--dump_case", str(out)],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        blob = out.read_bytes()
        e = er.energy_reader("/root/reference/LUTs/", water_cube_test=True, nuclear_corr=model)
        n = e.nucWeightMatrix.size
        tail = blob[-(8 + 4 + 8 * n):]
        assert tail[:8] == b"RTDNUCL1" and struct.unpack("<i", tail[8:12])[0] == model
        got = np.frombuffer(tail[12:], np.float32)
        assert np.array_equal(got[:n], e.nucWeightMatrix.ravel()) and np.array_equal(got[n:], e.nucSqSigmaMatrix.ravel())
    r = subprocess.run([EXE, "--water_cube_test", "--phys_data", "/root/reference/LUTs/", "--nuclear_corr", "nonsense", "--dump_case", str(tmp_path / "x")],
                       capture_output=True, text=True)
    assert r.returncode == 1 and "Unknown nuclear model" in r.stderr


@pytest.mark.gpu
def test_driver_matches_python_binding(tmp_path):
    from raytracedicom_b200.engine import DoseEngine
    _build()
    c = cases.make_case("mini"); p = str(tmp_path / "mini.rtdcase"); c.write(p)
    r = subprocess.run([EXE, "--case", p, "--output_directory", str(tmp_path), "--timing"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "Total global execution time" in r.stdout and "Max:" in r.stdout
    nx, ny, nz = c.dose_dims
    got = np.fromfile(tmp_path / "dose.dat", np.float32).reshape(nz, ny, nx)
    want = np.zeros_like(got)
    with DoseEngine(0) as eng:
        eng.set_luts(c.luts); eng.set_ct(c.ct); eng.set_dose_grid(nx, ny, nz); eng.dose(c.beams, want)
    assert rel_to_max(got, want)[0] <= 2e-6
