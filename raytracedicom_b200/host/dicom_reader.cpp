// See dicom_reader.h. Written from the DICOM standard (PS3.5 data-set encoding, PS3.10 file format) and from what the
// reference's readers extract; no ITK / GDCM / rti code is involved.
#include "dicom_reader.h"

#include <dirent.h>
#include <sys/stat.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <iostream>
#include <stdexcept>

namespace dcm {
namespace {

// Implicit-VR dictionary: only the attributes this driver interprets (everything else is kept as raw bytes).
const char* vrOf(unsigned t) {
    switch (t) {
        case tag(0x0008, 0x0060): return "CS";                                  // Modality
        case tag(0x0020, 0x000E): return "UI";                                  // SeriesInstanceUID
        case tag(0x0020, 0x0032): case tag(0x0020, 0x0037): case tag(0x0028, 0x0030): case tag(0x0018, 0x0050):
        case tag(0x0028, 0x1052): case tag(0x0028, 0x1053): return "DS";        // IPP, IOP, PixelSpacing, thickness, rescale
        case tag(0x0028, 0x0010): case tag(0x0028, 0x0011): case tag(0x0028, 0x0100): case tag(0x0028, 0x0103): return "US";
        case tag(0x7FE0, 0x0010): return "OW";
        case tag(0x300A, 0x03A2): case tag(0x300A, 0x03A8): case tag(0x300A, 0x0070): case tag(0x300C, 0x0004):
        case tag(0x300A, 0x0180): return "SQ";
        case tag(0x300A, 0x00C2): return "LO";                                  // BeamName
        case tag(0x300A, 0x030A): case tag(0x300A, 0x0394): case tag(0x300A, 0x0396): case tag(0x300A, 0x0398): return "FL";
        case tag(0x300A, 0x0114): case tag(0x300A, 0x011E): case tag(0x300A, 0x0120): case tag(0x300A, 0x0122):
        case tag(0x300A, 0x012C): return "DS";
        case tag(0x300A, 0x0392): return "IS";
        default: return "";
    }
}
bool longVr(const std::string& vr) {
    return vr == "OB" || vr == "OW" || vr == "OF" || vr == "SQ" || vr == "UT" || vr == "UN" || vr == "OD" || vr == "OL" || vr == "UC" || vr == "UR";
}
struct Buf {
    const std::vector<unsigned char>& b;
    uint16_t u16(size_t p) const { if (p + 2 > b.size()) throw std::runtime_error("DICOM: truncated file"); return uint16_t(b[p] | (b[p + 1] << 8)); }
    uint32_t u32(size_t p) const { if (p + 4 > b.size()) throw std::runtime_error("DICOM: truncated file"); return uint32_t(b[p]) | (uint32_t(b[p + 1]) << 8) | (uint32_t(b[p + 2]) << 16) | (uint32_t(b[p + 3]) << 24); }
};

// Parses elements in [pos, end) (or up to an item delimiter when the item has undefined length); returns the new position.
constexpr int kMaxDepth = 32;   // nesting of sequences: a corrupt file must not overflow the stack
size_t parse(const Buf& f, size_t pos, size_t end, bool explicitVr, DataSet& out, int depth = 0) {
    if (depth > kMaxDepth) throw std::runtime_error("DICOM: sequences nested deeper than 32 levels");
    end = std::min(end, f.b.size());       // lengths come from the file: never trust them beyond its size
    while (pos + 8 <= end) {
        const unsigned g = f.u16(pos), e = f.u16(pos + 2);
        if (g == 0xFFFE && (e == 0xE00D || e == 0xE0DD)) return pos + 8;      // item / sequence delimitation
        std::string vr; uint32_t len;
        if (explicitVr) {
            if (pos + 6 > f.b.size()) throw std::runtime_error("DICOM: truncated file");
            vr.assign(reinterpret_cast<const char*>(&f.b[pos + 4]), 2);
            if (longVr(vr)) { len = f.u32(pos + 8); pos += 12; } else { len = f.u16(pos + 6); pos += 8; }
        } else {
            len = f.u32(pos + 4); pos += 8;
            vr = vrOf(tag(g, e));
            if (vr.empty() && len == 0xFFFFFFFFu) vr = "SQ";
        }
        Element el; el.vr = vr;
        if (vr == "SQ" || (vr == "UN" && len == 0xFFFFFFFFu)) {
            const size_t seqEnd = (len == 0xFFFFFFFFu) ? end : std::min(pos + size_t(len), end);
            while (pos + 8 <= seqEnd) {
                const unsigned ig = f.u16(pos), ie = f.u16(pos + 2); const uint32_t ilen = f.u32(pos + 4);
                pos += 8;
                if (ig == 0xFFFE && ie == 0xE0DD) break;
                if (!(ig == 0xFFFE && ie == 0xE000)) throw std::runtime_error("DICOM: item tag expected inside a sequence");
                DataSet item;
                if (ilen == 0xFFFFFFFFu) pos = parse(f, pos, seqEnd, explicitVr, item, depth + 1);
                else { if (pos + ilen > f.b.size()) throw std::runtime_error("DICOM: truncated item"); parse(f, pos, pos + ilen, explicitVr, item, depth + 1); pos += ilen; }
                el.items.push_back(std::move(item));
            }
        } else {
            if (len == 0xFFFFFFFFu) throw std::runtime_error("DICOM: encapsulated (compressed) pixel data is not supported");
            if (pos + len > f.b.size()) throw std::runtime_error("DICOM: truncated element");
            el.raw.assign(f.b.begin() + pos, f.b.begin() + pos + len);
            pos += len;
        }
        out[tag(g, e)] = std::move(el);
    }
    return pos;
}

}  // namespace

DataSet readFile(const std::string& path) {
    std::ifstream in(path, std::ios::binary);
    if (!in) throw std::runtime_error("cannot open " + path);
    std::vector<unsigned char> bytes((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
    if (bytes.size() < 144 || memcmp(&bytes[128], "DICM", 4) != 0) throw std::runtime_error(path + ": not a DICOM Part-10 file");
    const Buf f{bytes};
    if (!(f.u16(132) == 0x0002 && f.u16(134) == 0x0000)) throw std::runtime_error(path + ": file meta group length missing");
    const size_t metaEnd = 144 + size_t(f.u32(140));
    if (metaEnd > bytes.size()) throw std::runtime_error(path + ": file meta group longer than the file");
    DataSet meta, ds;
    parse(f, 144, metaEnd, true, meta);
    const std::string ts = str(meta, tag(0x0002, 0x0010));
    bool explicitVr;
    if (ts == "1.2.840.10008.1.2") explicitVr = false;
    else if (ts == "1.2.840.10008.1.2.1") explicitVr = true;
    else throw std::runtime_error(path + ": transfer syntax " + ts + " not supported (uncompressed little endian only)");
    parse(f, metaEnd, bytes.size(), explicitVr, ds);
    return ds;
}

bool has(const DataSet& ds, unsigned t) { return ds.find(t) != ds.end(); }

std::string str(const DataSet& ds, unsigned t) {
    auto it = ds.find(t);
    if (it == ds.end()) return "";
    std::string s(it->second.raw.begin(), it->second.raw.end());
    while (!s.empty() && (s.back() == ' ' || s.back() == '\0')) s.pop_back();
    size_t a = 0; while (a < s.size() && s[a] == ' ') ++a;
    return s.substr(a);
}

std::vector<double> numbers(const DataSet& ds, unsigned t) {
    std::vector<double> out;
    auto it = ds.find(t);
    if (it == ds.end()) return out;
    const Element& el = it->second;
    std::string vr = el.vr.empty() ? vrOf(t) : el.vr;
    const unsigned char* p = el.raw.data(); const size_t n = el.raw.size();
    auto bin = [&](size_t w, auto conv) { for (size_t i = 0; i + w <= n; i += w) out.push_back(conv(p + i)); };
    if (vr == "FL") bin(4, [](const unsigned char* q) { float v; memcpy(&v, q, 4); return double(v); });
    else if (vr == "FD") bin(8, [](const unsigned char* q) { double v; memcpy(&v, q, 8); return v; });
    else if (vr == "US") bin(2, [](const unsigned char* q) { uint16_t v; memcpy(&v, q, 2); return double(v); });
    else if (vr == "SS") bin(2, [](const unsigned char* q) { int16_t v; memcpy(&v, q, 2); return double(v); });
    else if (vr == "UL") bin(4, [](const unsigned char* q) { uint32_t v; memcpy(&v, q, 4); return double(v); });
    else if (vr == "SL") bin(4, [](const unsigned char* q) { int32_t v; memcpy(&v, q, 4); return double(v); });
    else {   // DS / IS / unknown: backslash-separated decimal strings
        std::string s(el.raw.begin(), el.raw.end());
        size_t a = 0;
        while (a <= s.size()) {
            size_t b = s.find('\\', a); if (b == std::string::npos) b = s.size();
            const std::string tok = s.substr(a, b - a);
            if (tok.find_first_not_of(" \0", 0, 2) != std::string::npos) out.push_back(std::strtod(tok.c_str(), nullptr));
            a = b + 1;
        }
    }
    return out;
}

const std::vector<DataSet>& seq(const DataSet& ds, unsigned t) {
    auto it = ds.find(t);
    if (it == ds.end()) throw std::runtime_error("DICOM: required sequence missing");
    return it->second.items;
}

}  // namespace dcm

// ---------------------------------------------------------------------------------------------------------
namespace {

struct D3 { double x, y, z; };
D3 cross(D3 a, D3 b) { return D3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
double dot(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

struct M33 { double a[3][3]; };
M33 mul(const M33& p, const M33& q) {
    M33 r;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { r.a[i][j] = 0; for (int k = 0; k < 3; ++k) r.a[i][j] += p.a[i][k] * q.a[k][j]; }
    return r;
}
M33 inverse(const M33& m) {
    const double (*a)[3] = m.a;
    const double det = a[0][0] * (a[1][1] * a[2][2] - a[1][2] * a[2][1]) - a[0][1] * (a[1][0] * a[2][2] - a[1][2] * a[2][0]) +
                       a[0][2] * (a[1][0] * a[2][1] - a[1][1] * a[2][0]);
    if (det == 0.0) throw std::runtime_error("singular image orientation matrix");
    M33 r;
    r.a[0][0] = (a[1][1] * a[2][2] - a[1][2] * a[2][1]) / det; r.a[0][1] = (a[0][2] * a[2][1] - a[0][1] * a[2][2]) / det; r.a[0][2] = (a[0][1] * a[1][2] - a[0][2] * a[1][1]) / det;
    r.a[1][0] = (a[1][2] * a[2][0] - a[1][0] * a[2][2]) / det; r.a[1][1] = (a[0][0] * a[2][2] - a[0][2] * a[2][0]) / det; r.a[1][2] = (a[0][2] * a[1][0] - a[0][0] * a[1][2]) / det;
    r.a[2][0] = (a[1][0] * a[2][1] - a[1][1] * a[2][0]) / det; r.a[2][1] = (a[0][1] * a[2][0] - a[0][0] * a[2][1]) / det; r.a[2][2] = (a[0][0] * a[1][1] - a[0][1] * a[1][0]) / det;
    return r;
}
M33 fromAffine(const Float3AffineTransform& t) {
    const Matrix3x3 m = t.getMatrix();
    const float3 r0 = m.row0(), r1 = m.row1(), r2 = m.row2();
    return M33{{{r0.x, r0.y, r0.z}, {r1.x, r1.y, r1.z}, {r2.x, r2.y, r2.z}}};
}

struct Slice { dcm::DataSet ds; double loc; };

}  // namespace

Float3AffineTransform itk_reader(const std::string& imagePath, std::vector<float>& imageData, unsigned int& N, uint3& dim, clock_t& t) {
    using namespace dcm;
    t = clock();
    std::vector<std::string> names;
    if (DIR* d = opendir(imagePath.c_str())) {
        while (dirent* e = readdir(d)) names.push_back(e->d_name);
        closedir(d);
    } else {
        throw std::runtime_error("cannot open directory " + imagePath);
    }
    std::sort(names.begin(), names.end());
    std::vector<Slice> slices;
    std::string series;
    for (const std::string& n : names) {
        const std::string p = imagePath + "/" + n;
        struct stat sb;
        if (stat(p.c_str(), &sb) != 0 || !S_ISREG(sb.st_mode)) continue;
        DataSet ds;
        try { ds = readFile(p); } catch (const std::exception&) { continue; }
        if (str(ds, tag(0x0008, 0x0060)) != "CT" || !has(ds, tag(0x7FE0, 0x0010))) continue;
        slices.push_back(Slice{std::move(ds), 0.0});
    }
    if (slices.empty()) {
        std::cout << "The directory " << imagePath << " contains no DICOM Series:\n";
        throw std::runtime_error("Abort: no DICOM series in " + imagePath);
    }
    // the reference reads the first series UID GDCM reports (dicom_reader.cpp:57-58): take the smallest UID present
    series = str(slices[0].ds, tag(0x0020, 0x000E));
    for (const Slice& s : slices) series = std::min(series, str(s.ds, tag(0x0020, 0x000E)));
    slices.erase(std::remove_if(slices.begin(), slices.end(), [&](const Slice& s) { return str(s.ds, tag(0x0020, 0x000E)) != series; }), slices.end());
    std::cout << "The directory " << imagePath << " contains the following DICOM Series:\n" << series << "\n\nPlease wait while reading series:\n" << series << std::endl;

    const std::vector<double> iop = numbers(slices[0].ds, tag(0x0020, 0x0037));
    if (iop.size() != 6) throw std::runtime_error("ImageOrientationPatient missing");
    const D3 row{iop[0], iop[1], iop[2]}, col{iop[3], iop[4], iop[5]}, normal = cross(row, col);
    for (Slice& s : slices) {
        const std::vector<double> ipp = numbers(s.ds, tag(0x0020, 0x0032));
        if (ipp.size() != 3) throw std::runtime_error("ImagePositionPatient missing");
        s.loc = dot(normal, D3{ipp[0], ipp[1], ipp[2]});
    }
    std::stable_sort(slices.begin(), slices.end(), [](const Slice& a, const Slice& b) { return a.loc < b.loc; });

    const unsigned ny = unsigned(numbers(slices[0].ds, tag(0x0028, 0x0010)).at(0)), nx = unsigned(numbers(slices[0].ds, tag(0x0028, 0x0011)).at(0));
    const std::vector<double> ps = numbers(slices[0].ds, tag(0x0028, 0x0030));   // (row spacing, column spacing)
    if (ps.size() != 2) throw std::runtime_error("PixelSpacing missing");
    dim = make_uint3(nx, ny, unsigned(slices.size()));
    N = dim.x * dim.y * dim.z;
    imageData.resize(N);
    constexpr short HUOFFSET = 1000;   // dicom_reader.cpp:24, 104
    for (size_t k = 0; k < slices.size(); ++k) {
        const DataSet& ds = slices[k].ds;
        if (unsigned(numbers(ds, tag(0x0028, 0x0010)).at(0)) != ny || unsigned(numbers(ds, tag(0x0028, 0x0011)).at(0)) != nx ||
            int(numbers(ds, tag(0x0028, 0x0100)).at(0)) != 16)
            throw std::runtime_error("inconsistent slice geometry / only 16-bit pixels are supported");
        const std::vector<double> pr = numbers(ds, tag(0x0028, 0x0103)), sl = numbers(ds, tag(0x0028, 0x1053)), ic = numbers(ds, tag(0x0028, 0x1052));
        const bool isSigned = pr.empty() || pr[0] == 1.0;
        const double slope = sl.empty() ? 1.0 : sl[0], icpt = ic.empty() ? 0.0 : ic[0];
        const std::vector<unsigned char>& px = ds.at(tag(0x7FE0, 0x0010)).raw;
        if (px.size() < size_t(nx) * ny * 2) throw std::runtime_error("PixelData shorter than Rows x Columns");
        float* dst = imageData.data() + k * size_t(nx) * ny;
        for (size_t i = 0; i < size_t(nx) * ny; ++i) {
            const uint16_t u = uint16_t(px[2 * i] | (px[2 * i + 1] << 8));
            const double stored = isSigned ? double(int16_t(u)) : double(u);
            const short hu = short(stored * slope + icpt);          // ITK pixel type short (dicom_reader.cpp:17)
            dst[i] = float(hu + HUOFFSET);
        }
    }
    const std::vector<double> p0 = numbers(slices[0].ds, tag(0x0020, 0x0032));
    double dz;
    if (slices.size() > 1) dz = slices[1].loc - slices[0].loc;
    else { const std::vector<double> th = numbers(slices[0].ds, tag(0x0018, 0x0050)); dz = th.empty() ? 1.0 : th[0]; }
    t = clock() - t;
    std::cout << "Done!\n\nRead image: " << float(t) / CLOCKS_PER_SEC << " seconds\n\n";
    // imDir * imSpacing (dicom_reader.cpp:117-128): columns of imDir = row direction, column direction, slice normal
    const Matrix3x3 m(float(row.x * ps[1]), float(col.x * ps[0]), float(normal.x * dz),
                      float(row.y * ps[1]), float(col.y * ps[0]), float(normal.y * dz),
                      float(row.z * ps[1]), float(col.z * ps[0]), float(normal.z * dz));
    return Float3AffineTransform(m, make_float3(float(p0[0]), float(p0[1]), float(p0[2])));
}

std::vector<PlanBeam> readIonPlan(const std::string& path) {
    using namespace dcm;
    const DataSet ds = readFile(path);
    if (str(ds, tag(0x0008, 0x0060)) != "RTPLAN" || !has(ds, tag(0x300A, 0x03A2))) throw std::runtime_error(path + ": not an RT Ion Plan");
    // beamFromPlan's gantry / couch chain is the head-first-supine one (IEC 61217 as rti_coordinate_transform.hpp:48-58 applies it):
    // any other patient position would be dose-calculated mirrored or rotated, silently. Refuse it.
    if (has(ds, tag(0x300A, 0x0180)))
        for (const DataSet& ps : seq(ds, tag(0x300A, 0x0180))) {
            const std::string pp = str(ps, tag(0x0018, 0x5100));
            if (!pp.empty() && pp != "HFS") throw std::runtime_error(path + ": PatientPosition " + pp + " is not supported (head-first supine, HFS, only)");
        }
    std::vector<PlanBeam> beams;
    for (const DataSet& be : seq(ds, tag(0x300A, 0x03A2))) {
        PlanBeam pb;
        pb.name = str(be, tag(0x300A, 0x00C2));
        const std::vector<double> sad = numbers(be, tag(0x300A, 0x030A));
        if (sad.size() != 2) throw std::runtime_error("VirtualSourceAxisDistances missing in beam " + pb.name);
        pb.sad = make_float2(float(sad[0]), float(sad[1]));
        const std::vector<DataSet>& cps = seq(be, tag(0x300A, 0x03A8));
        if (cps.empty()) throw std::runtime_error("beam " + pb.name + " has no control points");
        auto first = [&](unsigned t, double def) { const std::vector<double> v = numbers(cps[0], t); return v.empty() ? def : v[0]; };
        pb.collimatorAngle = float(first(tag(0x300A, 0x0120), 0.0));
        pb.gantryAngle = float(first(tag(0x300A, 0x011E), 0.0));
        pb.couchAngle = float(first(tag(0x300A, 0x0122), 0.0));
        const std::vector<double> iso = numbers(cps[0], tag(0x300A, 0x012C));
        if (iso.size() != 3) throw std::runtime_error("IsocenterPosition missing in beam " + pb.name);
        pb.isocenter = make_float3(float(iso[0]), float(iso[1]), float(iso[2]));
        for (size_t i = 0; i < cps.size(); i += 2) {     // odd control points carry zero weights (rti_beam_module_ion.hpp:86-90)
            const DataSet& cp = cps[i];
            const std::vector<double> e = numbers(cp, tag(0x300A, 0x0114)), np = numbers(cp, tag(0x300A, 0x0392)), xy = numbers(cp, tag(0x300A, 0x0394)),
                                      w = numbers(cp, tag(0x300A, 0x0396)), fw = numbers(cp, tag(0x300A, 0x0398));
            if (e.empty() || np.empty() || fw.size() != 2) throw std::runtime_error("incomplete ion control point in beam " + pb.name);
            if (!(e[0] > 0.0 && e[0] < 1.0e4) || !(fw[0] > 0.0 && fw[0] < 1.0e4) || !(fw[1] > 0.0 && fw[1] < 1.0e4) || !(np[0] >= 0.0 && np[0] <= 1.0e7))
                throw std::runtime_error("ion control point with an energy, spot size or spot count out of range in beam " + pb.name);
            const size_t n = size_t(np[0]);
            if (xy.size() < 2 * n || w.size() < n) throw std::runtime_error("scan spot map shorter than NumberOfScanSpotPositions in beam " + pb.name);
            PlanLayer l; l.energy = float(e[0]); l.fwhmX = float(fw[0]); l.fwhmY = float(fw[1]);
            for (size_t j = 0; j < n; ++j) { l.x.push_back(float(xy[2 * j])); l.y.push_back(float(xy[2 * j + 1])); l.w.push_back(float(w[j])); }
            pb.layers.push_back(std::move(l));
        }
        beams.push_back(std::move(pb));
    }
    return beams;
}

// ---------------------------------------------------------------------------------------------------------
// Plan beam -> BeamSettings. Same rules, in the same order and precision (double, rounded to float at the end), as
// raytracedicom_b200/plan.py, which documents the conventions and their sources in the reference.
namespace {

constexpr double kFwhmToSigma = 2.355;   // rbe_1p1.hpp:90
constexpr double kLatticeTol = 1.0e-3;   // mm
constexpr double kMaxPosition = 1.0e4;   // mm: anything farther from the axis than 10 m is a corrupt file
constexpr double kMaxLatticeNodes = 4096.0;

void lattice(const std::vector<double>& values, double defaultPitch, int minCount, double& x0, double& pitch, int& n) {
    std::vector<double> u;
    for (double v : values) {
        if (!std::isfinite(v) || std::fabs(v) > kMaxPosition) throw std::runtime_error("scan spot position is not a finite distance in mm");
        u.push_back(std::nearbyint(v / kLatticeTol) * kLatticeTol);
    }
    std::sort(u.begin(), u.end());
    u.erase(std::unique(u.begin(), u.end()), u.end());
    if (u.size() == 1) pitch = defaultPitch;
    else {
        pitch = u[1] - u[0];
        for (size_t i = 2; i < u.size(); ++i) pitch = std::min(pitch, u[i] - u[i - 1]);
        for (double v : u) {
            const double k = (v - u[0]) / pitch;
            if (std::fabs(k - std::nearbyint(k)) * pitch > 10 * kLatticeTol) throw std::runtime_error("scan spot positions do not lie on a regular lattice");
        }
    }
    const double nodes = std::nearbyint((u.back() - u.front()) / pitch) + 1.0;
    if (!(nodes <= kMaxLatticeNodes)) throw std::runtime_error("scan spot lattice has too many nodes");
    n = int(nodes);
    x0 = u.front();
    for (int side = 0; n < minCount; side ^= 1) { if (side == 0) x0 -= pitch; ++n; }
}

}  // namespace

std::unique_ptr<PlanBeamSettings> beamFromPlan(const PlanBeam& pb, const Float3AffineTransform& imIdxToWorld, uint3 dim) {
    if (pb.layers.empty()) throw std::runtime_error("beam " + pb.name + " has no energy layers");
    std::vector<double> allx, ally;
    for (const PlanLayer& l : pb.layers) { allx.insert(allx.end(), l.x.begin(), l.x.end()); ally.insert(ally.end(), l.y.begin(), l.y.end()); }
    if (allx.empty()) throw std::runtime_error("beam " + pb.name + " has no scan spots");
    double x0, dx, y0, dy; int sx, sy;
    lattice(allx, 5.0, 3, x0, dx, sx);
    lattice(ally, 5.0, 3, y0, dy, sy);
    const size_t L = pb.layers.size();
    std::unique_ptr<PlanBeamSettings> out(new PlanBeamSettings());
    out->weights.assign(L * size_t(sy) * sx, 0.0f);
    std::vector<float> energies(L); std::vector<float2> sigmas(L);
    for (size_t li = 0; li < L; ++li) {
        const PlanLayer& l = pb.layers[li];
        for (size_t j = 0; j < l.x.size(); ++j) {
            const int ix = int(std::nearbyint((double(l.x[j]) - x0) / dx)), iy = int(std::nearbyint((double(l.y[j]) - y0) / dy));
            if (ix < 0 || ix >= sx || iy < 0 || iy >= sy || !std::isfinite(l.w[j])) throw std::runtime_error("scan spot outside its lattice or non-finite weight in beam " + pb.name);
            out->weights[(li * sy + iy) * sx + ix] += l.w[j];       // repaintings of one position add up
        }
        energies[li] = l.energy;
        sigmas[li] = make_float2(float(double(l.fwhmX) / kFwhmToSigma), float(double(l.fwhmY) / kFwhmToSigma));
    }
    // IEC gantry -> DICOM patient (HFS): T Rz(-couch) Ry(gantry), T: (x, y, z)_fixed -> (x, -z, y)_patient
    const double g = double(pb.gantryAngle) * M_PI / 180.0, s = -double(pb.couchAngle) * M_PI / 180.0;
    const M33 ry{{{std::cos(g), 0.0, std::sin(g)}, {0.0, 1.0, 0.0}, {-std::sin(g), 0.0, std::cos(g)}}};
    const M33 rz{{{std::cos(s), -std::sin(s), 0.0}, {std::sin(s), std::cos(s), 0.0}, {0.0, 0.0, 1.0}}};
    const M33 tm{{{1.0, 0.0, 0.0}, {0.0, 0.0, -1.0}, {0.0, 1.0, 0.0}}};
    const M33 rot = mul(tm, mul(rz, ry));
    const M33 aInv = inverse(fromAffine(imIdxToWorld));
    const M33 md = mul(aInv, rot);
    const float3 org = imIdxToWorld.getOffset();
    const double d[3] = {double(pb.isocenter.x) - org.x, double(pb.isocenter.y) - org.y, double(pb.isocenter.z) - org.z};
    float mf[3][3], vf[3];
    for (int i = 0; i < 3; ++i) {
        vf[i] = float(aInv.a[i][0] * d[0] + aInv.a[i][1] * d[1] + aInv.a[i][2] * d[2]);
        for (int j = 0; j < 3; ++j) mf[i][j] = float(md.a[i][j]);
    }
    // tracer start plane / number of steps: the CT box seen along the beam axis (+-2 mm), cases.steps_and_start
    M33 mfd; for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) mfd.a[i][j] = mf[i][j];
    const M33 minv = inverse(mfd);
    double zmin = 1e300, zmax = -1e300;
    const double nd[3] = {double(dim.x), double(dim.y), double(dim.z)};
    for (int c = 0; c < 8; ++c) {
        const double idx[3] = {((c & 1) ? nd[0] : 0.0) - 0.5 - vf[0], ((c & 2) ? nd[1] : 0.0) - 0.5 - vf[1], ((c & 4) ? nd[2] : 0.0) - 0.5 - vf[2]};
        const double z = minv.a[2][0] * idx[0] + minv.a[2][1] * idx[1] + minv.a[2][2] * idx[2];
        zmin = std::min(zmin, z); zmax = std::max(zmax, z);
    }
    const double zStart = std::ceil(zmax + 2.0), zEnd = std::floor(zmin - 2.0);
    const unsigned steps = unsigned(std::ceil((zStart - zEnd) / 1.0));
    out->weightImage.reset(new HostPinnedImage3D<float>(out->weights.data(), make_uint3(unsigned(sx), unsigned(sy), unsigned(L))));
    const Float3AffineTransform gantryToImIdx(Matrix3x3(mf[0][0], mf[0][1], mf[0][2], mf[1][0], mf[1][1], mf[1][2], mf[2][0], mf[2][1], mf[2][2]),
                                              make_float3(vf[0], vf[1], vf[2]));
    out->beam.reset(new BeamSettings(out->weightImage.get(), energies, sigmas, make_float2(1.0f, 1.0f), steps, pb.sad,
                                     Float3IdxTransform(make_float3(float(dx), float(dy), -1.0f), make_float3(float(x0), float(y0), float(zStart))),
                                     gantryToImIdx, gantryToImIdx));
    return out;
}
