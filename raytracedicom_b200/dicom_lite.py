"""Minimal DICOM Part-10 reader / writer for the two object types the dose path consumes: CT Image slices and
RT Ion Plans.  Stands in for pydicom (generator side) and for ITK/GDCM + the reference's ``rti`` dataset layer
(reader side), none of which exist in this image.

Scope: Implicit VR Little Endian (what generate_water_cube.py writes: /root/reference/extern/dicom-interface/rti/
test/dicom/generate_water_cube.py:118,212) and Explicit VR Little Endian data sets, nested sequences with defined or
undefined length, uncompressed pixel data.  The C++ host driver has its own reader (raytracedicom_b200/host/
dicom_reader.cpp); this module is the Python mirror used by the tests and by ``generate_water_cube`` below, which
reproduces the reference generator's CT + plan tag for tag where the dose path (or ``rti``) reads them.
"""
from __future__ import annotations

import os
import struct
from typing import Dict, List, Tuple, Union

import numpy as np

# (group, element) -> (VR, keyword) for every attribute read or written here.
TAGS: Dict[Tuple[int, int], Tuple[str, str]] = {
    (0x0002, 0x0000): ("UL", "FileMetaInformationGroupLength"), (0x0002, 0x0001): ("OB", "FileMetaInformationVersion"),
    (0x0002, 0x0002): ("UI", "MediaStorageSOPClassUID"), (0x0002, 0x0003): ("UI", "MediaStorageSOPInstanceUID"),
    (0x0002, 0x0010): ("UI", "TransferSyntaxUID"), (0x0002, 0x0012): ("UI", "ImplementationClassUID"),
    (0x0008, 0x0005): ("CS", "SpecificCharacterSet"), (0x0008, 0x0008): ("CS", "ImageType"),
    (0x0008, 0x0016): ("UI", "SOPClassUID"), (0x0008, 0x0018): ("UI", "SOPInstanceUID"),
    (0x0008, 0x0020): ("DA", "StudyDate"), (0x0008, 0x0021): ("DA", "SeriesDate"), (0x0008, 0x0030): ("TM", "StudyTime"),
    (0x0008, 0x0060): ("CS", "Modality"), (0x0008, 0x0070): ("LO", "Manufacturer"), (0x0008, 0x103E): ("LO", "SeriesDescription"),
    (0x0010, 0x0010): ("PN", "PatientName"), (0x0010, 0x0020): ("LO", "PatientID"),
    (0x0018, 0x0050): ("DS", "SliceThickness"), (0x0018, 0x5100): ("CS", "PatientPosition"),
    (0x0020, 0x000D): ("UI", "StudyInstanceUID"), (0x0020, 0x000E): ("UI", "SeriesInstanceUID"),
    (0x0020, 0x0011): ("IS", "SeriesNumber"), (0x0020, 0x0013): ("IS", "InstanceNumber"),
    (0x0020, 0x0032): ("DS", "ImagePositionPatient"), (0x0020, 0x0037): ("DS", "ImageOrientationPatient"),
    (0x0020, 0x0052): ("UI", "FrameOfReferenceUID"), (0x0020, 0x1041): ("DS", "SliceLocation"),
    (0x0028, 0x0002): ("US", "SamplesPerPixel"), (0x0028, 0x0004): ("CS", "PhotometricInterpretation"),
    (0x0028, 0x0010): ("US", "Rows"), (0x0028, 0x0011): ("US", "Columns"), (0x0028, 0x0030): ("DS", "PixelSpacing"),
    (0x0028, 0x0100): ("US", "BitsAllocated"), (0x0028, 0x0101): ("US", "BitsStored"), (0x0028, 0x0102): ("US", "HighBit"),
    (0x0028, 0x0103): ("US", "PixelRepresentation"), (0x0028, 0x1052): ("DS", "RescaleIntercept"),
    (0x0028, 0x1053): ("DS", "RescaleSlope"), (0x0028, 0x1054): ("LO", "RescaleType"),
    (0x7FE0, 0x0010): ("OW", "PixelData"),
    (0x300A, 0x0002): ("SH", "RTPlanLabel"), (0x300A, 0x0003): ("LO", "RTPlanName"), (0x300A, 0x000C): ("CS", "RTPlanGeometry"),
    (0x300A, 0x0070): ("SQ", "FractionGroupSequence"), (0x300A, 0x0071): ("IS", "FractionGroupNumber"),
    (0x300A, 0x0078): ("IS", "NumberOfFractionsPlanned"), (0x300A, 0x0080): ("IS", "NumberOfBeams"),
    (0x300C, 0x0004): ("SQ", "ReferencedBeamSequence"), (0x300C, 0x0006): ("IS", "ReferencedBeamNumber"),
    (0x300A, 0x0086): ("DS", "BeamMeterset"),
    (0x300A, 0x0180): ("SQ", "PatientSetupSequence"), (0x300A, 0x0182): ("IS", "PatientSetupNumber"),
    (0x300A, 0x03A2): ("SQ", "IonBeamSequence"), (0x300A, 0x00B2): ("SH", "TreatmentMachineName"),
    (0x300A, 0x00B3): ("CS", "PrimaryDosimeterUnit"), (0x300A, 0x00C0): ("IS", "BeamNumber"), (0x300A, 0x00C2): ("LO", "BeamName"),
    (0x300A, 0x00C4): ("CS", "BeamType"), (0x300A, 0x00C6): ("CS", "RadiationType"), (0x300A, 0x00CE): ("CS", "TreatmentDeliveryType"),
    (0x300A, 0x010E): ("DS", "FinalCumulativeMetersetWeight"), (0x300A, 0x0110): ("IS", "NumberOfControlPoints"),
    (0x300A, 0x0308): ("CS", "ScanMode"), (0x300A, 0x030A): ("FL", "VirtualSourceAxisDistances"),
    (0x300A, 0x0312): ("IS", "NumberOfRangeShifters"),
    (0x300A, 0x03A8): ("SQ", "IonControlPointSequence"), (0x300A, 0x0112): ("IS", "ControlPointIndex"),
    (0x300A, 0x0114): ("DS", "NominalBeamEnergy"), (0x300A, 0x0015): ("CS", "NominalBeamEnergyUnit"),
    (0x300A, 0x011E): ("DS", "GantryAngle"), (0x300A, 0x011F): ("CS", "GantryRotationDirection"),
    (0x300A, 0x0120): ("DS", "BeamLimitingDeviceAngle"), (0x300A, 0x0122): ("DS", "PatientSupportAngle"),
    (0x300A, 0x012C): ("DS", "IsocenterPosition"), (0x300A, 0x0134): ("DS", "CumulativeMetersetWeight"),
    (0x300A, 0x030D): ("FL", "SnoutPosition"), (0x300A, 0x0390): ("SH", "ScanSpotTuneID"),
    (0x300A, 0x0392): ("IS", "NumberOfScanSpotPositions"), (0x300A, 0x0394): ("FL", "ScanSpotPositionMap"),
    (0x300A, 0x0396): ("FL", "ScanSpotMetersetWeights"), (0x300A, 0x0398): ("FL", "ScanningSpotSize"),
    (0x300A, 0x039A): ("IS", "NumberOfPaintings"),
}
KEYWORD_TO_TAG = {kw: tag for tag, (vr, kw) in TAGS.items()}

IMPLICIT_LE = "1.2.840.10008.1.2"
EXPLICIT_LE = "1.2.840.10008.1.2.1"
CT_IMAGE_STORAGE = "1.2.840.10008.5.1.4.1.1.2"
RT_ION_PLAN_STORAGE = "1.2.840.10008.5.1.4.1.1.481.8"
_LONG_VRS = ("OB", "OW", "OF", "SQ", "UT", "UN", "OD", "OL", "UC", "UR")
_STRING_VRS = ("AE", "AS", "CS", "DA", "DS", "DT", "IS", "LO", "LT", "PN", "SH", "ST", "TM", "UI", "UT", "UC", "UR")
_BIN_FMT = {"US": "<H", "SS": "<h", "UL": "<I", "SL": "<i", "FL": "<f", "FD": "<d"}

Value = Union[str, bytes, int, float, list]


class MissingAttribute(KeyError, ValueError):
    """A required attribute is absent (the C++ reader's std::runtime_error "... missing", host/dicom_reader.cpp)."""

    def __str__(self):
        return str(self.args[0]) if self.args else ""


class Dataset(dict):
    """tag -> value. Values: str / list of str for string VRs (numbers for DS, IS), list of numbers for binary VRs,
    bytes for OB/OW/UN, list of Dataset for SQ. Access by keyword: ``ds["Rows"]``, ``ds.get_kw("Rows")``."""

    def __getitem__(self, key):
        try:
            return dict.__getitem__(self, KEYWORD_TO_TAG[key] if isinstance(key, str) else key)
        except KeyError:
            raise MissingAttribute(f"DICOM attribute {key} missing") from None

    def __setitem__(self, key, value):
        dict.__setitem__(self, KEYWORD_TO_TAG[key] if isinstance(key, str) else key, value)

    def __contains__(self, key):
        return dict.__contains__(self, KEYWORD_TO_TAG.get(key, (-1, -1)) if isinstance(key, str) else key)

    def get_kw(self, key, default=None):
        return self[key] if key in self else default

    def floats(self, key) -> List[float]:
        v = self[key]
        return [float(x) for x in (v if isinstance(v, list) else [v])]


# ---- encoding ---------------------------------------------------------------------------------------------
def _fmt_number(x) -> str:
    if isinstance(x, (int, np.integer)):
        return str(int(x))
    s = repr(float(x))
    if len(s) > 16:                       # DS is limited to 16 characters
        s = "%.10g" % float(x)
    return s


def _encode_value(vr: str, value: Value) -> bytes:
    if vr in ("OB", "OW", "UN", "OF"):
        b = bytes(value)
    elif vr in _BIN_FMT:
        vals = value if isinstance(value, (list, tuple, np.ndarray)) else [value]
        b = b"".join(struct.pack(_BIN_FMT[vr], v) for v in vals)
    elif vr in _STRING_VRS:
        vals = value if isinstance(value, (list, tuple)) else [value]
        b = "\\".join(v if isinstance(v, str) else _fmt_number(v) for v in vals).encode("latin-1")
    else:
        raise ValueError(f"unsupported VR {vr}")
    if len(b) & 1:
        b += b"\0" if vr in ("UI", "OB", "UN") else b" "
    return b


def _encode_dataset(ds: Dataset, explicit: bool) -> bytes:
    out = []
    for tag in sorted(ds.keys()):
        vr = TAGS[tag][0] if tag in TAGS else "UN"
        v = ds[tag]
        if vr == "SQ":
            items = b"".join(struct.pack("<HHI", 0xFFFE, 0xE000, len(body)) + body for body in (_encode_dataset(it, explicit) for it in v))
            payload = items
        else:
            payload = _encode_value(vr, v)
        if explicit:
            if vr in _LONG_VRS:
                out.append(struct.pack("<HH2sHI", tag[0], tag[1], vr.encode(), 0, len(payload)))
            else:
                out.append(struct.pack("<HH2sH", tag[0], tag[1], vr.encode(), len(payload)))
        else:
            out.append(struct.pack("<HHI", tag[0], tag[1], len(payload)))
        out.append(payload)
    return b"".join(out)


def write_file(path: str, ds: Dataset, sop_class: str, sop_instance: str, transfer_syntax: str = IMPLICIT_LE) -> None:
    """Part-10 file: 128-byte preamble, 'DICM', group 0002 in Explicit VR LE, data set in ``transfer_syntax``."""
    meta = Dataset()
    meta[(0x0002, 0x0001)] = b"\x00\x01"
    meta[(0x0002, 0x0002)] = sop_class
    meta[(0x0002, 0x0003)] = sop_instance
    meta[(0x0002, 0x0010)] = transfer_syntax
    meta[(0x0002, 0x0012)] = "1.2.826.0.1.3680043.10.1337.1"
    body = _encode_dataset(meta, True)
    glen = struct.pack("<HH2sHI", 0x0002, 0x0000, b"UL", 4, len(body))
    with open(path, "wb") as f:
        f.write(b"\0" * 128 + b"DICM" + glen + body + _encode_dataset(ds, transfer_syntax == EXPLICIT_LE))


# ---- decoding ---------------------------------------------------------------------------------------------
def _decode_value(vr: str, raw: bytes) -> Value:
    if vr in _BIN_FMT:
        n = struct.calcsize(_BIN_FMT[vr])
        vals = [struct.unpack_from(_BIN_FMT[vr], raw, i)[0] for i in range(0, len(raw) - n + 1, n)]
        return vals[0] if len(vals) == 1 else vals
    if vr in _STRING_VRS:
        s = raw.decode("latin-1").rstrip(" \0")
        parts = s.split("\\") if vr not in ("LT", "ST", "UT") else [s]
        if vr in ("DS", "IS"):
            conv = float if vr == "DS" else int
            parts = [conv(p) for p in parts if p.strip() != ""]
        return parts[0] if len(parts) == 1 else parts
    return raw


def _parse(buf: bytes, pos: int, end: int, explicit: bool, stop_at_item_end: bool = False) -> Tuple[Dataset, int]:
    ds = Dataset()
    while pos < end:
        g, e = struct.unpack_from("<HH", buf, pos)
        if (g, e) == (0xFFFE, 0xE00D) or (g, e) == (0xFFFE, 0xE0DD):      # item / sequence delimitation
            return ds, pos + 8
        if explicit:
            vr = buf[pos + 4:pos + 6].decode("latin-1")
            if vr in _LONG_VRS:
                (length,) = struct.unpack_from("<I", buf, pos + 8); pos += 12
            else:
                (length,) = struct.unpack_from("<H", buf, pos + 6); pos += 8
        else:
            (length,) = struct.unpack_from("<I", buf, pos + 4); pos += 8
            vr = TAGS[(g, e)][0] if (g, e) in TAGS else ("SQ" if length == 0xFFFFFFFF else "UN")
        if vr == "SQ" or (vr == "UN" and length == 0xFFFFFFFF):
            items = []
            seq_end = end if length == 0xFFFFFFFF else pos + length
            while pos < seq_end:
                ig, ie, ilen = struct.unpack_from("<HHI", buf, pos)
                pos += 8
                if (ig, ie) == (0xFFFE, 0xE0DD):
                    break
                if (ig, ie) != (0xFFFE, 0xE000):
                    raise ValueError(f"expected item tag at {pos - 8}, found ({ig:04x},{ie:04x})")
                if ilen == 0xFFFFFFFF:
                    item, pos = _parse(buf, pos, seq_end, explicit, True)
                else:
                    item, _ = _parse(buf, pos, pos + ilen, explicit)
                    pos += ilen
                items.append(item)
            ds[(g, e)] = items
            continue
        if length == 0xFFFFFFFF:
            raise ValueError(f"undefined length on non-sequence ({g:04x},{e:04x}): encapsulated pixel data is not supported")
        if pos + length > len(buf):
            raise ValueError(f"value of ({g:04x},{e:04x}) runs past the end of the file (truncated?)")
        ds[(g, e)] = _decode_value(vr, buf[pos:pos + length])
        pos += length
    return ds, pos


def read_file(path: str) -> Dataset:
    buf = open(path, "rb").read()
    if len(buf) < 132 or buf[128:132] != b"DICM":
        raise ValueError(f"{path}: not a DICOM Part-10 file")
    pos = 132
    try:
        g, e, vr, ln, glen = struct.unpack_from("<HH2sHI", buf, pos)
        if (g, e) != (0x0002, 0x0000):
            raise ValueError(f"{path}: file meta group length missing")
        if pos + 12 + glen > len(buf):
            raise ValueError(f"{path}: file meta group runs past the end of the file")
        meta, _ = _parse(buf, pos + 12, pos + 12 + glen, True)
        ts = meta[(0x0002, 0x0010)]
        if ts not in (IMPLICIT_LE, EXPLICIT_LE):
            raise ValueError(f"{path}: transfer syntax {ts} not supported (uncompressed little endian only)")
        ds, _ = _parse(buf, pos + 12 + glen, len(buf), ts == EXPLICIT_LE)
    except (struct.error, IndexError, UnicodeDecodeError, RecursionError, TypeError) as err:      # cut-off or corrupt element headers / values
        raise ValueError(f"{path}: truncated or corrupt DICOM data ({err})") from None
    ds.update(meta)
    return ds


# ---- CT series (the role of itk_reader, /root/reference/src/dicom_reader.cpp:15-129) -----------------------
def read_ct_series(directory: str):
    """All CT slices of the first series in ``directory`` -> (ct float32 [nz][ny][nx] = stored*slope+intercept+1000,
    im_idx_to_world 3x3 = direction*spacing, origin (3,)), like itk_reader: slices ordered along the slice normal
    by ImagePositionPatient (GDCMSeriesFileNames), pixel type short, HU offset +1000 (dicom_reader.cpp:24,104),
    transform = imDir*imSpacing with the origin of the first slice (dicom_reader.cpp:117-128)."""
    slices = []
    for name in sorted(os.listdir(directory)):
        p = os.path.join(directory, name)
        if not os.path.isfile(p):
            continue
        try:
            ds = read_file(p)
        except ValueError:
            continue
        if ds.get_kw("Modality") != "CT" or "PixelData" not in ds:
            continue
        slices.append(ds)
    if not slices:
        raise ValueError(f"The directory {directory} contains no DICOM Series")
    uid = sorted({s.get_kw("SeriesInstanceUID", "") for s in slices})[0]
    slices = [s for s in slices if s.get_kw("SeriesInstanceUID", "") == uid]
    iop = slices[0].floats("ImageOrientationPatient")
    row, col = np.array(iop[:3]), np.array(iop[3:])
    normal = np.cross(row, col)
    slices.sort(key=lambda s: float(np.dot(normal, s.floats("ImagePositionPatient"))))
    ny, nx = int(slices[0]["Rows"]), int(slices[0]["Columns"])
    ps = slices[0].floats("PixelSpacing")          # (row spacing = dy, column spacing = dx)
    vol = np.empty((len(slices), ny, nx), np.float32)
    for k, s in enumerate(slices):
        if int(s["Rows"]) != ny or int(s["Columns"]) != nx or int(s["BitsAllocated"]) != 16:
            raise ValueError("inconsistent slice geometry / only 16-bit pixels are supported")
        dt = np.int16 if int(s.get_kw("PixelRepresentation", 1)) == 1 else np.uint16
        px = np.frombuffer(s["PixelData"], dtype=dt, count=ny * nx).reshape(ny, nx).astype(np.float64)
        px = px * float(s.get_kw("RescaleSlope", 1.0)) + float(s.get_kw("RescaleIntercept", 0.0))
        vol[k] = (px.astype(np.int16).astype(np.float32) + np.float32(1000.0))
    p0 = np.array(slices[0].floats("ImagePositionPatient"))
    if len(slices) > 1:
        dz = float(np.dot(normal, np.array(slices[1].floats("ImagePositionPatient")) - p0))
    else:
        dz = float(slices[0].get_kw("SliceThickness", 1.0))
    direction = np.stack([row, col, normal], axis=1)                 # columns = index axes in patient coordinates
    m = direction @ np.diag([ps[1], ps[0], dz])
    return vol, m.astype(np.float32), p0.astype(np.float32)


# ---- RT Ion Plan (the role of rti::treatment_session / beam_module_ion) -------------------------------------
def read_ion_plan(path: str) -> List[dict]:
    """Beams of an RT Ion Plan: name, angles, isocentre, virtual SADs and the spot list per energy layer with the
    reference's rule (rti_beam_module_ion.hpp:86-104): control points come in pairs, the odd-indexed ones (all-zero
    weights) are skipped; each kept one yields energy, N spots, ScanningSpotSize (FWHM x, y), positions, weights."""
    ds = read_file(path)
    if ds.get_kw("Modality") != "RTPLAN" or "IonBeamSequence" not in ds:
        raise ValueError(f"{path}: not an RT Ion Plan")
    # plan.beam_from_plan's gantry / couch chain is the head-first-supine one: refuse anything else instead of computing a mirrored dose
    if "PatientSetupSequence" in ds:
        for ps in ds["PatientSetupSequence"]:
            pp = str(ps.get_kw("PatientPosition", "") or "").strip()
            if pp and pp != "HFS":
                raise ValueError(f"{path}: PatientPosition {pp} is not supported (head-first supine, HFS, only)")
    beams = []
    for be in ds["IonBeamSequence"]:
        cps = be["IonControlPointSequence"]
        cp0 = cps[0]
        layers = []
        for i, cp in enumerate(cps):
            if i % 2:
                continue
            n = int(cp["NumberOfScanSpotPositions"])
            xy = cp.floats("ScanSpotPositionMap")
            w = cp.floats("ScanSpotMetersetWeights")
            fw = cp.floats("ScanningSpotSize")
            if len(fw) != 2:
                raise ValueError(f"{path}: incomplete ion control point (ScanningSpotSize)")
            en = float(cp["NominalBeamEnergy"])
            if not (0.0 < en < 1.0e4 and 0.0 < fw[0] < 1.0e4 and 0.0 < fw[1] < 1.0e4 and 0 <= n <= 10 ** 7):
                raise ValueError(f"{path}: ion control point with an energy, spot size or spot count out of range")
            if len(xy) < 2 * n or len(w) < n:
                raise ValueError(f"{path}: scan spot map shorter than NumberOfScanSpotPositions")
            layers.append({"energy": float(cp["NominalBeamEnergy"]), "fwhm": (fw[0], fw[1]),
                           "x": np.array(xy[0:2 * n:2], np.float32), "y": np.array(xy[1:2 * n:2], np.float32),
                           "w": np.array(w[:n], np.float32)})
        sad = be.floats("VirtualSourceAxisDistances")
        beams.append({"name": str(be["BeamName"]).strip(), "gantry_angle": float(cp0["GantryAngle"]),
                      "collimator_angle": float(cp0.get_kw("BeamLimitingDeviceAngle", 0.0)),
                      "couch_angle": float(cp0.get_kw("PatientSupportAngle", 0.0)),
                      "isocenter": np.array(cp0.floats("IsocenterPosition"), np.float32),
                      "sad": (sad[0], sad[1]), "layers": layers})
    return beams


# ---- generators -------------------------------------------------------------------------------------------
def write_ct_volume(outdir: str, hu: np.ndarray, origin, spacing, uid_root: str = "1.2.826.0.1.3680043.10.1337") -> None:
    """hu: int16 [nz][ny][nx]; one file per slice, named like generate_water_cube.py (1.dcm ...), slice k at
    origin + k*spacing[2] along +z when spacing[2] > 0 (the generator writes descending z: pass a negative dz)."""
    os.makedirs(outdir, exist_ok=True)
    nz, ny, nx = hu.shape
    for k in range(nz):
        ds = Dataset()
        sop = f"{uid_root}.3.{k + 1}"
        ds["SpecificCharacterSet"] = "ISO_IR 100"; ds["ImageType"] = ["DERIVED", "SECONDARY", "AXIAL"]
        ds["SOPClassUID"] = CT_IMAGE_STORAGE; ds["SOPInstanceUID"] = sop
        ds["StudyDate"] = "20191011"; ds["SeriesDate"] = "20191011"; ds["StudyTime"] = "150000"
        ds["Modality"] = "CT"; ds["Manufacturer"] = "MGH Physics Research"; ds["SeriesDescription"] = "Water cube"
        ds["PatientName"] = "Cube^Water"; ds["PatientID"] = "WaterCube50"
        ds["SliceThickness"] = abs(float(spacing[2])); ds["PatientPosition"] = "HFS"
        ds["StudyInstanceUID"] = uid_root + ".1"; ds["SeriesInstanceUID"] = uid_root + ".2"
        ds["SeriesNumber"] = 1; ds["InstanceNumber"] = k + 1
        ds["ImagePositionPatient"] = [float(origin[0]), float(origin[1]), float(origin[2]) + k * float(spacing[2])]
        ds["ImageOrientationPatient"] = [1, 0, 0, 0, 1, 0]
        ds["FrameOfReferenceUID"] = uid_root + ".4"; ds["SliceLocation"] = float(origin[2]) + k * float(spacing[2])
        ds["SamplesPerPixel"] = 1; ds["PhotometricInterpretation"] = "MONOCHROME2"
        ds["Rows"] = ny; ds["Columns"] = nx; ds["PixelSpacing"] = [float(spacing[1]), float(spacing[0])]
        ds["BitsAllocated"] = 16; ds["BitsStored"] = 16; ds["HighBit"] = 15; ds["PixelRepresentation"] = 1
        ds["RescaleIntercept"] = 0; ds["RescaleSlope"] = 1; ds["RescaleType"] = "HU"
        ds["PixelData"] = np.ascontiguousarray(hu[k], "<i2").tobytes()
        write_file(os.path.join(outdir, f"{k + 1}.dcm"), ds, CT_IMAGE_STORAGE, sop)


def write_ion_plan(path: str, beams: List[dict], uid_root: str = "1.2.826.0.1.3680043.10.1337", explicit: bool = False) -> None:
    """beams: dicts like read_ion_plan returns. Every layer becomes a PAIR of control points (weights, then zeros),
    as treatment planning systems and generate_water_cube.py:328-363 write them."""
    ds = Dataset()
    sop = uid_root + ".5"
    ds["SpecificCharacterSet"] = "ISO_IR 100"; ds["SOPClassUID"] = RT_ION_PLAN_STORAGE; ds["SOPInstanceUID"] = sop
    ds["StudyDate"] = "20191011"; ds["StudyTime"] = "150000"; ds["Modality"] = "RTPLAN"
    ds["PatientName"] = "Cube^Water"; ds["PatientID"] = "WaterCube50"
    ds["StudyInstanceUID"] = uid_root + ".1"; ds["SeriesInstanceUID"] = uid_root + ".6"; ds["FrameOfReferenceUID"] = uid_root + ".4"
    ds["RTPlanLabel"] = "Test"; ds["RTPlanName"] = "Pristine test"; ds["RTPlanGeometry"] = "PATIENT"
    fg = Dataset(); fg["FractionGroupNumber"] = 1; fg["NumberOfFractionsPlanned"] = 1; fg["NumberOfBeams"] = len(beams)
    refs = []
    for i, b in enumerate(beams):
        r = Dataset(); r["ReferencedBeamNumber"] = i + 1
        r["BeamMeterset"] = float(sum(float(l["w"].sum()) for l in b["layers"]))
        refs.append(r)
    fg["ReferencedBeamSequence"] = refs
    ds["FractionGroupSequence"] = [fg]
    pss = Dataset(); pss["PatientPosition"] = "HFS"; pss["PatientSetupNumber"] = 1
    ds["PatientSetupSequence"] = [pss]
    seq = []
    for i, b in enumerate(beams):
        be = Dataset()
        be["TreatmentMachineName"] = "Gantry"; be["PrimaryDosimeterUnit"] = "NP"; be["BeamNumber"] = i + 1; be["BeamName"] = b["name"]
        be["BeamType"] = "STATIC"; be["RadiationType"] = "PROTON"; be["TreatmentDeliveryType"] = "TREATMENT"
        be["ScanMode"] = "MODULATED"; be["VirtualSourceAxisDistances"] = [float(b["sad"][0]), float(b["sad"][1])]
        be["NumberOfRangeShifters"] = 0; be["NumberOfControlPoints"] = 2 * len(b["layers"])
        total = float(sum(float(l["w"].sum()) for l in b["layers"]))
        be["FinalCumulativeMetersetWeight"] = total
        cps, cw = [], 0.0
        for li, l in enumerate(b["layers"]):
            for j in range(2):
                cp = Dataset()
                cp["NominalBeamEnergyUnit"] = "MEV"; cp["ControlPointIndex"] = 2 * li + j; cp["NominalBeamEnergy"] = float(l["energy"])
                if li == 0 and j == 0:
                    cp["GantryAngle"] = float(b["gantry_angle"]); cp["GantryRotationDirection"] = "NONE"
                    cp["BeamLimitingDeviceAngle"] = float(b.get("collimator_angle", 0.0)); cp["PatientSupportAngle"] = float(b.get("couch_angle", 0.0))
                    cp["IsocenterPosition"] = [float(v) for v in b["isocenter"]]; cp["SnoutPosition"] = 200.0
                cp["CumulativeMetersetWeight"] = cw
                cp["ScanSpotTuneID"] = "Tune"; cp["NumberOfScanSpotPositions"] = int(l["x"].size)
                xy = np.empty(2 * l["x"].size, np.float32); xy[0::2] = l["x"]; xy[1::2] = l["y"]
                cp["ScanSpotPositionMap"] = [float(v) for v in xy]
                w = l["w"] if j == 0 else np.zeros_like(l["w"])
                cp["ScanSpotMetersetWeights"] = [float(v) for v in w]
                cp["ScanningSpotSize"] = [float(l["fwhm"][0]), float(l["fwhm"][1])]; cp["NumberOfPaintings"] = 1
                cw += float(w.sum())
                cps.append(cp)
        be["IonControlPointSequence"] = cps
        seq.append(be)
    ds["IonBeamSequence"] = seq
    write_file(path, ds, RT_ION_PLAN_STORAGE, sop, EXPLICIT_LE if explicit else IMPLICIT_LE)


def generate_water_cube(outdir: str, n: int = 512, margin: int = 6, gantry_angle: float = 0.0, beam_name: str = "G000",
                        sad=(2000.0, 1800.0), energies=None, fwhm: float = 6.0, weight: float = 1.0e9) -> None:
    """The README example data set (generate_water_cube.py:96-105, 180, 24-25, 328-363): n^3 voxels @1 mm, -1000 HU rim of
    ``margin`` px around a 0 HU cube, slices written with descending z; image origin such that (0,0,0) is the centre of
    the anterior water surface; one beam, one spot (0,0) of ``weight`` per energy 90..230 MeV step 5, FWHM 6 mm,
    layers in descending energy, isocentre (0,0,0)."""
    cube = np.full((n, n, n), -1000, np.int16)
    cube[margin:-margin, margin:-margin, margin:-margin] = 0
    origin = ((-n / 2 + 0.5) * 1.0, (-margin + 0.5) * 1.0, (n / 2 - 0.5) * 1.0)
    write_ct_volume(os.path.join(outdir, "ct"), cube, origin, (1.0, 1.0, -1.0))
    if energies is None:
        energies = [90 + 5 * i for i in range(29)]
    layers = [{"energy": float(e), "fwhm": (fwhm, fwhm), "x": np.zeros(1, np.float32), "y": np.zeros(1, np.float32),
               "w": np.full(1, weight, np.float32)} for e in list(energies)[::-1]]
    write_ion_plan(os.path.join(outdir, "rtplan.dcm"),
                   [{"name": beam_name, "gantry_angle": gantry_angle, "isocenter": (0.0, 0.0, 0.0), "sad": sad, "layers": layers}])
