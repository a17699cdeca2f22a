"""The C-ABI shared library loads without a GPU, exports every symbol include/rtdose.h declares, its structs have
the layout the ctypes mirror assumes, and creating a context without a device fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess
import sys
import tempfile

import pytest

from conftest import ROOT
from raytracedicom_b200 import engine

HEADER = os.path.join(ROOT, "include", "rtdose.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rtd_[a-z_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = engine.load_library()
    names = declared_functions()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"librtdose.so does not export {n}"
    assert sorted(engine.EXPORTS) == names, "engine.EXPORTS and include/rtdose.h disagree"


def test_struct_layouts_match_header():
    """Compile a tiny C program against include/rtdose.h and compare sizeof/offsetof with the ctypes mirrors."""
    checks = {
        "rtd_luts": (engine.RtdLuts, ["nEnergySamples", "ciddMatrix", "densityScaleFact", "spVector", "rRlVector", "nucWeightMatrix", "nucSqSigmaMatrix"]),
        "rtd_beam": (engine.RtdBeam, ["weights", "nLayers", "sigmas", "steps", "spotOffset", "gantryToDoseIdxV", "layerBegin", "layerEnd"]),
        "rtd_opts": (engine.RtdOpts, ["bpDepthCutoff", "minTilesInBatch", "ks2BodyStep", "ctPull", "nuclearCorr", "noGraph", "fullIntermediates", "tracerBrick"]),
        "rtd_timing": (engine.RtdTiming, ["msTotal", "msDownload", "updates", "ksTaps", "hostSyncs", "doseVoxelsModified", "ksLaunchesWithWork", "tracerBrickBeams", "tracerOutsideSamples"]),
        "rtd_beam_report": (engine.RtdBeamReport, ["lSteps", "beamFirstCalculatedPassive", "bboxMax", "entryZ", "updates"]),
        "rtd_layer_report": (engine.RtdLayerReport, ["energyIdx", "layerFirstPassive", "radiusHist", "batchRadius"]),
    }
    lines = ["#include <stdio.h>", "#include <stddef.h>", f'#include "{HEADER}"', "int main(void){"]
    for s, (_, fields) in checks.items():
        lines.append(f'printf("{s} %zu\\n", sizeof({s}));')
        for f in fields:
            lines.append(f'printf("{s}.{f} %zu\\n", offsetof({s}, {f}));')
    lines.append("return 0;}")
    with tempfile.TemporaryDirectory() as d:
        cpath, exe = os.path.join(d, "t.c"), os.path.join(d, "t")
        open(cpath, "w").write("\n".join(lines))
        subprocess.run(["gcc", "-std=c99", cpath, "-o", exe], check=True)
        out = subprocess.run([exe], capture_output=True, text=True, check=True).stdout
    got = dict(l.split() for l in out.strip().splitlines())
    for s, (ct, fields) in checks.items():
        assert int(got[s]) == C.sizeof(ct), s
        for f in fields:
            assert int(got[f"{s}.{f}"]) == getattr(ct, f).offset, f"{s}.{f}"


def test_header_is_plain_c():
    """The boundary must be bindable from C / cgo / JNI: compiles as C99 with no C++ or torch types."""
    with tempfile.TemporaryDirectory() as d:
        cpath = os.path.join(d, "t.c")
        open(cpath, "w").write(f'#include "{HEADER}"\nint main(void){{ rtd_opts o; rtd_beam b; (void)o; (void)b; return (int)sizeof(rtd_luts) == 0; }}\n')
        subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-c", cpath, "-o", os.path.join(d, "t.o")], check=True)


def test_default_opts_and_version_without_gpu():
    lib = engine.load_library()
    o = engine.RtdOpts()
    lib.rtd_default_opts(C.byref(o))
    assert (o.bpDepthCutoff, o.convSigmaCutoff, o.ksSigmaCutoff, o.rayWeightCutoff, o.minTilesInBatch) == pytest.approx((1.05, 3.0, 3.0, 1.0, 16))
    assert b"sm_100a" in lib.rtd_version()


def test_no_cpu_fallback():
    """Without a CUDA device the product refuses to run (RTD_ERR_NO_DEVICE); with one this test only checks that
    invalid device numbers are rejected."""
    import torch
    lib = engine.load_library()
    h = C.c_void_p()
    if not torch.cuda.is_available():
        assert lib.rtd_create(C.byref(h), 0) == 5
        with pytest.raises(engine.RtdError):
            engine.DoseEngine(0)
    else:
        assert lib.rtd_create(C.byref(h), 9999) == 1


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under raytracedicom_b200/ may reference it."""
    pkg = os.path.join(ROOT, "raytracedicom_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h", ".cpp")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "pyoracle" not in txt and "rtd_oracle" not in txt and "librtd_oracle" not in txt, os.path.join(dp, f)
