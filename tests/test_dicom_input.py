"""DICOM input of the dose path: CT series reader (role of itk_reader, src/dicom_reader.cpp:15-129),
RT Ion Plan reader (role of rti::beam_module_ion) and the plan -> BeamSettings mapping the reference's main() leaves out
(src/main.cu:171-197). The C++ host driver and the Python mirror must produce byte-identical plans."""
import os
import shutil
import subprocess

import numpy as np
import pytest

from conftest import ROOT
from raytracedicom_b200 import cases, dicom_lite as dl, energy_reader, plan

EXE = os.path.join(ROOT, "raytracedicom_b200", "rtdose_main")


@pytest.fixture(scope="module")
def exe():
    subprocess.run(["make", "-C", os.path.join(ROOT, "raytracedicom_b200", "host")], check=True, capture_output=True)
    return EXE


@pytest.fixture(scope="module")
def lut_dir(tmp_path_factory):
    d = str(tmp_path_factory.mktemp("luts")) + "/"
    energy_reader.write_text_luts(d)
    return d


def test_text_lut_writer_round_trips_bit_exact(lut_dir):
    a, b = energy_reader.load_packaged_luts(), energy_reader.energy_reader(lut_dir)
    for k in ("energiesPerU", "peakDepths", "scaleFacts", "ciddMatrix", "densityVector", "spVector", "rRlVector"):
        assert np.array_equal(np.asarray(getattr(a, k), np.float32).view(np.uint32), np.asarray(getattr(b, k), np.float32).view(np.uint32)), k


@pytest.mark.parametrize("explicit", [False, True])
def test_dataset_round_trip_with_nested_sequences(tmp_path, explicit):
    beams = [{"name": "B1", "gantry_angle": 30.0, "couch_angle": 10.0, "collimator_angle": 5.0, "isocenter": (1.5, -2.0, 3.25), "sad": (2100.0, 1900.0),
              "layers": [{"energy": 120.5, "fwhm": (7.0, 8.0), "x": np.array([-5, 0, 5], np.float32), "y": np.array([0, 0, 5], np.float32),
                          "w": np.array([1.0, 2.5, 4.0], np.float32)},
                         {"energy": 110.0, "fwhm": (7.5, 8.5), "x": np.array([0], np.float32), "y": np.array([-5], np.float32), "w": np.array([3.0], np.float32)}]}]
    p = str(tmp_path / "plan.dcm")
    dl.write_ion_plan(p, beams, explicit=explicit)
    ds = dl.read_file(p)
    assert ds["TransferSyntaxUID"] == (dl.EXPLICIT_LE if explicit else dl.IMPLICIT_LE)
    cps = ds["IonBeamSequence"][0]["IonControlPointSequence"]
    assert len(cps) == 4 and [int(c["ControlPointIndex"]) for c in cps] == [0, 1, 2, 3]
    got = dl.read_ion_plan(p)[0]
    assert got["name"] == "B1" and got["gantry_angle"] == 30.0 and got["couch_angle"] == 10.0 and got["sad"] == (2100.0, 1900.0)
    assert np.allclose(got["isocenter"], (1.5, -2.0, 3.25))
    # spot-to-layer assignment: odd control points (zero weights) skipped, spots in file order (rti_beam_module_ion.hpp:86-104)
    assert [l["energy"] for l in got["layers"]] == [120.5, 110.0]
    assert np.array_equal(got["layers"][0]["x"], [-5, 0, 5]) and np.array_equal(got["layers"][0]["w"], [1.0, 2.5, 4.0])
    assert got["layers"][1]["fwhm"] == (7.5, 8.5)


def test_reader_rejects_what_it_cannot_decode(tmp_path):
    p = str(tmp_path / "x.dcm")
    open(p, "wb").write(b"not dicom at all")
    with pytest.raises(ValueError):
        dl.read_file(p)
    ds = dl.Dataset(); ds["Modality"] = "CT"
    dl.write_file(p, ds, dl.CT_IMAGE_STORAGE, "1.2.3", transfer_syntax="1.2.840.10008.1.2.4.70")   # JPEG lossless
    with pytest.raises(ValueError, match="transfer syntax"):
        dl.read_file(p)
    with pytest.raises(ValueError, match="no DICOM Series"):
        dl.read_ct_series(str(tmp_path))


def test_water_cube_ct_series(tmp_path):
    """generate_water_cube.py:96-105,180: -1000 HU rim, 0 HU cube, slices written with DESCENDING z and read back in
    ascending order; HU + 1000 as float (dicom_reader.cpp:104); transform = direction * spacing, origin of the first slice."""
    d = str(tmp_path)
    dl.generate_water_cube(d, n=40, margin=6)
    ct, m, o = dl.read_ct_series(os.path.join(d, "ct"))
    assert ct.shape == (40, 40, 40) and ct.dtype == np.float32
    want = np.zeros((40, 40, 40), np.float32); want[6:-6, 6:-6, 6:-6] = 1000.0
    assert np.array_equal(ct, want)
    assert np.array_equal(m, np.eye(3, dtype=np.float32))
    assert np.allclose(o, (-19.5, -5.5, -19.5))
    # a marked voxel proves the slice order: file 1.dcm is the slice of HIGHEST z
    hu = np.full((3, 4, 5), -1000, np.int16); hu[0, 1, 2] = 77
    dl.write_ct_volume(os.path.join(d, "m"), hu, (0.0, 0.0, 10.0), (2.0, 3.0, -4.0))
    ct, m, o = dl.read_ct_series(os.path.join(d, "m"))
    assert ct[2, 1, 2] == 1077.0 and np.allclose(np.diag(m), (2.0, 3.0, 4.0)) and np.allclose(o, (0.0, 0.0, 2.0))


def _oblique_plan():
    rng = np.random.default_rng(7)
    beams = []
    for name, g, c in (("A30", 30.0, 0.0), ("B270", 270.0, 12.0)):
        layers = []
        for e in (150.0, 140.0, 130.0):
            xs, ys = np.meshgrid(np.arange(-12, 13, 4.0), np.arange(-8, 9, 4.0))
            keep = rng.random(xs.size) > 0.3
            x, y = xs.ravel()[keep].astype(np.float32), ys.ravel()[keep].astype(np.float32)
            x = np.concatenate([x, x[:2]]); y = np.concatenate([y, y[:2]])             # two repainted positions
            layers.append({"energy": e, "fwhm": (8.0, 9.0), "x": x, "y": y, "w": (rng.random(x.size) * 1e8 + 1e7).astype(np.float32)})
        beams.append({"name": name, "gantry_angle": g, "couch_angle": c, "isocenter": (3.0, 40.0, -2.0), "sad": (2000.0, 1800.0), "layers": layers})
    return beams


def test_plan_to_beam_settings_mapping(tmp_path):
    d = str(tmp_path)
    dl.generate_water_cube(d, n=64, margin=6)
    ct, m, o = dl.read_ct_series(os.path.join(d, "ct"))
    pl = _oblique_plan()
    b = plan.beam_settings_from_plan(pl[0], ct.shape, m, o)
    # regular grid [layer][sy][sx] over the distinct positions of the whole beam; repaintings add up; total weight conserved
    assert b.weights.shape == (3, 5, 7) and b.spot_delta == (4.0, 4.0, -1.0) and b.spot_offset[:2] == (-12.0, -8.0)
    for li, l in enumerate(pl[0]["layers"]):
        assert np.isclose(b.weights[li].sum(), l["w"].sum(), rtol=1e-6)
        j = 3
        assert b.weights[li][int((l["y"][j] + 8) / 4), int((l["x"][j] + 12) / 4)] >= l["w"][j]
    assert np.allclose(b.sigmas, [[8.0 / 2.355, 9.0 / 2.355]] * 3) and np.array_equal(b.energies, [150, 140, 130])
    assert b.source_dist == (2000.0, 1800.0) and b.ray_spacing == (1.0, 1.0)
    # geometry: the isocentre (gantry origin) lands on its image index, the beam axis (-z gantry) points along the IEC
    # direction of a 30 degree gantry: from the source at (sin30, -cos30, 0)*SAD (patient LPS, HFS) towards the isocentre
    iso_idx = (np.array(pl[0]["isocenter"]) - o) / 1.0
    assert np.allclose(b.gantry_to_im_v, iso_idx, atol=1e-5)
    axis = b.gantry_to_im_m @ np.array([0.0, 0.0, -1.0])
    assert np.allclose(axis, (-np.sin(np.deg2rad(30)), np.cos(np.deg2rad(30)), 0.0), atol=1e-6)
    assert np.allclose(b.gantry_to_im_m @ np.array([0.0, 1.0, 0.0]), (0.0, 0.0, 1.0), atol=1e-6)      # gantry y = patient superior
    # the traced range covers the CT box along the beam
    zs = [(np.linalg.inv(b.gantry_to_im_m) @ (np.array(c) - b.gantry_to_im_v))[2] for c in ((-0.5,) * 3, (63.5,) * 3, (-0.5, 63.5, -0.5), (63.5, -0.5, 63.5))]
    assert b.spot_offset[2] >= max(zs) and b.spot_offset[2] - b.steps <= min(zs)
    # a single spot becomes the centre of a 3 x 3 grid @ 5 mm with zero-weight neighbours
    one = plan.beam_settings_from_plan(dl.read_ion_plan(os.path.join(d, "rtplan.dcm"))[0], ct.shape, m, o)
    assert one.weights.shape == (29, 3, 3) and one.spot_offset[:2] == (-5.0, -5.0) and one.spot_delta[:2] == (5.0, 5.0)
    assert np.count_nonzero(one.weights) == 29 and np.all(one.weights[:, 1, 1] == 1.0e9)
    assert list(one.energies) == list(range(230, 85, -5))                     # layers in file order: descending energy
    # spots off the lattice are an error, not an approximation
    bad = _oblique_plan()[0]; bad["layers"][0]["x"][0] += 1.3
    with pytest.raises(ValueError, match="regular lattice"):
        plan.beam_settings_from_plan(bad, ct.shape, m, o)


def test_cpp_driver_reads_dicom_like_the_python_mirror(tmp_path, exe, lut_dir):
    """--ct_dir / --rtplan / --beams (src/config.cpp:13-51) through host/dicom_reader.cpp: the assembled plan (CT, LUTs, every
    BeamSettings field) is byte-identical to the Python mirror's, for the README water cube and a two-beam oblique plan."""
    d = str(tmp_path)
    dl.generate_water_cube(d, n=48, margin=6)
    dl.write_ion_plan(os.path.join(d, "two.dcm"), _oblique_plan(), explicit=True)
    for rtplan, names in (("rtplan.dcm", ["G000"]), ("two.dcm", ["B270", "A30"]), ("two.dcm", [])):
        out = os.path.join(d, "cpp.rtdcase")
        cmd = [exe, "--ct_dir", os.path.join(d, "ct"), "--rtplan", os.path.join(d, rtplan), "--phys_data", lut_dir, "--dump_case", out]
        if names:
            cmd += ["--beams", ",".join(names)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        c = plan.case_from_dicom(os.path.join(d, "ct"), os.path.join(d, rtplan), names)
        ref = os.path.join(d, "py.rtdcase"); c.write(ref)
        assert open(out, "rb").read() == open(ref, "rb").read()
        assert len(c.beams) == (len(names) or 2)
    r = subprocess.run([exe, "--ct_dir", os.path.join(d, "ct"), "--rtplan", os.path.join(d, "two.dcm"), "--beams", "nope", "--phys_data", lut_dir,
                        "--dump_case", out], capture_output=True, text=True)
    assert r.returncode == 1 and "not found" in r.stderr


def test_dicom_water_cube_bragg_peaks_with_the_oracle(tmp_path, oracle):
    """Reference-free known answer (check 6): the Bragg peak of every layer sits within one step of the LUT's
    peak depth below the water surface, which the generator puts at patient y = 0 (generate_water_cube.py:180)."""
    d = str(tmp_path)
    dl.generate_water_cube(d, n=96, margin=6, energies=[80])
    c = plan.case_from_dicom(os.path.join(d, "ct"), os.path.join(d, "rtplan.dcm"), ["G000"])
    r = oracle.run_beam(c, 0)
    prof = r["dose"][46:50, :, 46:50].sum(axis=(0, 2))
    depth = int(np.argmax(prof)) - 6 + 0.5            # voxel j spans depth [j-6, j-5] mm
    assert abs(depth - cases.peak_depth(c.luts, 80.0)) <= 1.0
    assert prof[:5].max() == 0.0                       # nothing in the air rim (the step just before entry is kept, kernel_wrapper.cu:361)


@pytest.mark.gpu
def test_driver_dose_from_dicom_matches_binding_and_oracle(tmp_path, exe, lut_dir, oracle):
    """BASELINE.json configs[0] (water-cube CT + single-beam G000 plan from generate_water_cube.py) at reduced size plus a
    second oblique beam, end to end through the C++ driver: DICOM in -> dose.dat, against the Python binding on the same
    plan (fp32 atomic order only) and the CPU oracle (BASELINE tolerance 1e-3 of max)."""
    from conftest import rel_to_max
    from raytracedicom_b200.engine import DoseEngine
    d = str(tmp_path)
    dl.generate_water_cube(d, n=128, margin=6, energies=[100, 120, 140])
    beams = dl.read_ion_plan(os.path.join(d, "rtplan.dcm"))
    second = dict(beams[0]); second["name"] = "G300"; second["gantry_angle"] = 300.0; second["isocenter"] = (0.0, 60.0, 0.0)
    first = dict(beams[0]); first["isocenter"] = (0.0, 60.0, 0.0)
    dl.write_ion_plan(os.path.join(d, "two.dcm"), [first, second])
    r = subprocess.run([exe, "--ct_dir", os.path.join(d, "ct"), "--rtplan", os.path.join(d, "two.dcm"), "--beams", "G000,G300", "--phys_data", lut_dir,
                        "--output_directory", d], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got = np.fromfile(os.path.join(d, "dose.dat"), np.float32).reshape(128, 128, 128)
    c = plan.case_from_dicom(os.path.join(d, "ct"), os.path.join(d, "two.dcm"), ["G000", "G300"])
    want = np.zeros_like(got)
    with DoseEngine(0) as eng:
        eng.set_luts(c.luts); eng.set_dose_grid(128, 128, 128)
        eng.plan_dose(c.ct, c.beams, want)
    assert want.max() > 0 and rel_to_max(got, want)[0] <= 2e-6
    ref = np.zeros_like(got)
    for b in range(2):
        oracle.run_beam(c, b, dose=ref)
    assert rel_to_max(got, ref)[0] <= 1e-3


def _mutate(buf, rng):
    b = bytearray(buf); k = int(rng.integers(0, 4))
    if k == 0:
        b = b[:int(rng.integers(0, len(b)))]                                   # truncated file
    elif k == 1:
        for _ in range(int(rng.integers(1, 6))):
            b[int(rng.integers(0, len(b)))] = int(rng.integers(0, 256))        # flipped bytes (tags, VRs, lengths, values)
    elif k == 2:
        p = int(rng.integers(132, min(len(b), 1500))); b[p:p + 4] = bytes(rng.integers(0, 256, 4, dtype=np.uint8))
    else:
        p = int(rng.integers(132, len(b))); b[p:p + 4] = b"\xff\xff\xff\xff"   # undefined length / NaN float
    return bytes(b)


def test_corrupt_dicom_is_an_error_never_a_crash(tmp_path, exe, lut_dir):
    """Truncated and byte-flipped CT slices / plans: the Python mirror raises ValueError (nothing else) or reads the file;
    the C++ driver exits with status 0 or 1 and a message - no signal, no hang (a NaN spot position used to index
    outside the weight grid in beamFromPlan)."""
    base = str(tmp_path / "base")
    dl.generate_water_cube(base, n=24, margin=4)
    plan_bytes = open(os.path.join(base, "rtplan.dcm"), "rb").read()
    names = sorted(os.listdir(os.path.join(base, "ct")))
    slice_bytes = open(os.path.join(base, "ct", names[0]), "rb").read()
    rng = np.random.default_rng(11)
    d = str(tmp_path / "t")
    rejected = 0
    for trial in range(90):
        shutil.rmtree(d, ignore_errors=True); shutil.copytree(base, d)
        if trial % 2 == 0:
            open(os.path.join(d, "rtplan.dcm"), "wb").write(_mutate(plan_bytes, rng))
        else:
            open(os.path.join(d, "ct", names[int(rng.integers(0, len(names)))]), "wb").write(_mutate(slice_bytes, rng))
        try:
            plan.case_from_dicom(os.path.join(d, "ct"), os.path.join(d, "rtplan.dcm"))
        except ValueError:
            rejected += 1
        if trial % 5 == 0:
            r = subprocess.run([exe, "--ct_dir", os.path.join(d, "ct"), "--rtplan", os.path.join(d, "rtplan.dcm"), "--phys_data", lut_dir,
                                "--dump_case", str(tmp_path / "o.rtdcase")], capture_output=True, timeout=60)
            assert r.returncode in (0, 1), (trial, r.returncode, r.stderr[-200:])
    assert 10 < rejected < 90
    # the case that used to crash: a NaN in the scan spot position map
    beams = dl.read_ion_plan(os.path.join(base, "rtplan.dcm"))
    beams[0]["layers"][0]["x"][0] = np.nan
    ct, m, o = dl.read_ct_series(os.path.join(base, "ct"))
    with pytest.raises(ValueError, match="finite"):
        plan.beam_settings_from_plan(beams[0], ct.shape, m, o)
    dl.write_ion_plan(os.path.join(base, "nan.dcm"), beams)
    r = subprocess.run([exe, "--ct_dir", os.path.join(base, "ct"), "--rtplan", os.path.join(base, "nan.dcm"), "--phys_data", lut_dir,
                        "--dump_case", str(tmp_path / "o.rtdcase")], capture_output=True, text=True, errors="replace", timeout=60)
    assert r.returncode == 1 and "finite" in (r.stdout + r.stderr)


def test_non_hfs_plans_and_hostile_nesting_are_rejected(tmp_path, exe, lut_dir):
    """beamFromPlan's gantry / couch chain assumes head-first supine: a plan whose PatientSetupSequence says anything else is
    refused by both readers instead of being dose-calculated mirrored. Files that nest sequences without end, or end right after
    a tag, are errors, not stack overflows or out-of-bounds reads."""
    base = str(tmp_path / "base")
    dl.generate_water_cube(base, n=24, margin=4)
    plan_bytes = open(os.path.join(base, "rtplan.dcm"), "rb").read()
    assert plan_bytes.count(b"HFS") >= 1
    for pos in (b"HFP ", b"FFS ", b"FFP "):
        p = os.path.join(base, "pos.dcm")
        open(p, "wb").write(plan_bytes.replace(b"HFS ", pos))
        with pytest.raises(ValueError, match="PatientPosition"):
            dl.read_ion_plan(p)
        r = subprocess.run([exe, "--ct_dir", os.path.join(base, "ct"), "--rtplan", p, "--phys_data", lut_dir, "--dump_case", str(tmp_path / "o.rtdcase")],
                           capture_output=True, text=True, timeout=60)
        assert r.returncode == 1 and "PatientPosition" in r.stderr
    # 5000 nested undefined-length sequences (explicit VR LE): (300A,0180) SQ -> item -> (300A,0180) SQ -> ...
    dl.write_ion_plan(os.path.join(base, "explicit.dcm"), dl.read_ion_plan(os.path.join(base, "rtplan.dcm")), explicit=True)
    ex_bytes = open(os.path.join(base, "explicit.dcm"), "rb").read()
    head = ex_bytes[:ex_bytes.index(b"DICM") + 4]
    meta = ex_bytes[len(head):]
    meta_len = int.from_bytes(meta[8:12], "little") if meta[4:6] == b"UL" else None
    assert meta_len is not None
    meta = meta[:12 + meta_len]
    nest = (bytes([0x0A, 0x30, 0x80, 0x01]) + b"SQ\x00\x00" + b"\xff\xff\xff\xff" + bytes([0xFE, 0xFF, 0x00, 0xE0]) + b"\xff\xff\xff\xff") * 5000
    deep = os.path.join(base, "deep.dcm")
    open(deep, "wb").write(head + meta + nest)
    with pytest.raises(ValueError):
        dl.read_file(deep)
    r = subprocess.run([exe, "--ct_dir", os.path.join(base, "ct"), "--rtplan", deep, "--phys_data", lut_dir, "--dump_case", str(tmp_path / "o.rtdcase")],
                       capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and "nested" in r.stderr
    # a file that ends right after a tag
    cut = os.path.join(base, "cut.dcm")
    open(cut, "wb").write(head + meta + bytes([0x0A, 0x30, 0x80, 0x01]) + b"S")
    r = subprocess.run([exe, "--ct_dir", os.path.join(base, "ct"), "--rtplan", cut, "--phys_data", lut_dir, "--dump_case", str(tmp_path / "o.rtdcase")],
                       capture_output=True, text=True, timeout=60)
    assert r.returncode == 1
