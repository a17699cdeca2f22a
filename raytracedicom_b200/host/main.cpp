// Host driver of the B200 dose engine: the role of the reference's main() (src/main.cu:20-252) on top of the C ABI.
//   rtdose_main --water_cube_test [--phys_data DIR/] [--output_directory DIR]      (reference: -DWATER_CUBE_TEST build)
//   rtdose_main --case plan.rtdcase [--output_directory DIR]                        (raw CT + BeamSettings, any number of beams)
//   rtdose_main --ct_dir DIR --rtplan FILE [--beams A,B] --phys_data DIR/            (the reference's DICOM input, src/main.cu:105-197)
// The DICOM path reads the CT series and the RT Ion Plan with host/dicom_reader.cpp (no ITK/GDCM/rti) and - unlike the
// reference, whose main() prints the plan's spots but hands all-zero weights / energies / sigmas to the dose engine
// (src/main.cu:185-197) and stops at one beam (main.cu:117-120) - maps every requested beam to a BeamSettings.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <limits>
#include <memory>
#include <stdexcept>

#include "../../include/rtdose_shim.hpp"
#include "config.h"
#include "dicom_reader.h"
#include "energy_reader.h"
#include "host_types.h"

namespace {

// Ordered-list helpers with the semantics of src/vector_find.h:130-145 and src/vector_interpolate.h:19-30.
float findDecimalOrdered(const std::vector<float>& v, float x) {
    if (x >= v.back()) return float(v.size() - 1);
    if (x < v.front()) return 0.0f;
    const size_t hi = size_t(std::upper_bound(v.begin(), v.end(), x) - v.begin());   // first element > x
    const size_t lo = hi - 1;
    return float(lo) + (x - v[lo]) / (v[hi] - v[lo]);
}
float vectorInterpolate(const std::vector<float>& v, float idx) {
    if (idx <= 0.0f) return v.front();
    if (idx >= float(v.size() - 1)) return v.back();
    float ip; const float frac = std::modf(idx, &ip);
    const size_t i = size_t(ip);
    return v[i] + (v[i + 1] - v[i]) * frac;
}

struct Plan {   // everything main() hands to cudaWrapperProtons, with owned storage
    uint3 dim{0, 0, 0}, doseDim{0, 0, 0};
    std::vector<float> imageData;
    std::vector<std::vector<float>> weightData;
    std::vector<std::unique_ptr<HostPinnedImage3D<float>>> weightImages;
    std::vector<BeamSettings> beams;
    std::vector<std::unique_ptr<PlanBeamSettings>> planBeams;
    EnergyStruct luts; bool haveLuts = false;
    int nuclearCorr = 0;   // model the nuclear tables in luts belong to (0 = none)
};

// DICOM input: CT series + RT Ion Plan -> plan (src/main.cu:105-197 completed).
void readDicom(const Config& config, Plan& p) {
    unsigned int N = 0; clock_t t;
    const Float3AffineTransform imIdxToWorld = itk_reader(config.ct_dir, p.imageData, N, p.dim, t);
    p.doseDim = p.dim;
    const std::vector<PlanBeam> all = readIonPlan(config.rtplan);
    std::vector<std::string> names = config.beams;
    if (names.empty()) for (const PlanBeam& b : all) names.push_back(b.name);
    for (size_t i = 0; i < names.size(); ++i) {
        std::cout << "Loading field " << i << " corresponding to beamname " << names[i] << std::endl;
        const PlanBeam* pb = nullptr;
        for (const PlanBeam& b : all) if (b.name == names[i]) pb = &b;
        if (!pb) throw std::runtime_error("beam " + names[i] + " not found in " + config.rtplan);
        std::cout << "Angles: " << pb->collimatorAngle << " " << pb->gantryAngle << " " << pb->couchAngle << " deg\n"
                  << "IsoCenter: " << pb->isocenter.x << " " << pb->isocenter.y << " " << pb->isocenter.z << " mm\n"
                  << "ImgPixels: " << p.dim.x << " " << p.dim.y << " " << p.dim.z << "\n";
        size_t nSpots = 0;
        for (const PlanLayer& l : pb->layers) nSpots += l.x.size();
        std::cout << "Layers: " << pb->layers.size() << ", spots: " << nSpots << "\n\n";
        p.planBeams.push_back(beamFromPlan(*pb, imIdxToWorld, p.dim));
        p.beams.push_back(*p.planBeams.back()->beam);
    }
}

// RTDCASE1 writer (layout: oracle/ref_harness/case_io.h; reader: readCaseFile below, raytracedicom_b200/cases.py).
void writeCaseFile(const std::string& path, Plan& p) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    auto wr = [&](const void* src, size_t n) { if (n && fwrite(src, 1, n, f) != n) { fclose(f); throw std::runtime_error("write failed: " + path); } };
    auto wri = [&](int v) { wr(&v, 4); };
    auto wrf = [&](float v) { wr(&v, 4); };
    auto wrv = [&](const std::vector<float>& v) { wr(v.data(), v.size() * 4); };
    wr("RTDCASE1", 8);
    wri(int(p.dim.x)); wri(int(p.dim.y)); wri(int(p.dim.z)); wrv(p.imageData);
    wri(int(p.doseDim.x)); wri(int(p.doseDim.y)); wri(int(p.doseDim.z));
    const EnergyStruct& l = p.luts;
    wri(l.nEnergySamples); wri(l.nEnergies); wrv(l.energiesPerU); wrv(l.peakDepths); wrv(l.scaleFacts); wrv(l.ciddMatrix);
    wri(l.nDensitySamples); wrf(l.densityScaleFact); wrv(l.densityVector);
    wri(l.nSpSamples); wrf(l.spScaleFact); wrv(l.spVector);
    wri(l.nRRlSamples); wrf(l.rRlScaleFact); wrv(l.rRlVector);
    wri(int(p.beams.size()));
    for (BeamSettings& b : p.beams) {
        const uint3 wd = b.getWeights()->getDims();
        wri(int(wd.x)); wri(int(wd.y)); wri(int(wd.z)); wri(int(b.getSteps()));
        wr(b.getWeights()->getImData(), size_t(wd.x) * wd.y * wd.z * 4);
        wrv(b.getEnergies());
        for (const float2& s : b.getSpotSigmas()) { wrf(s.x); wrf(s.y); }
        wrf(b.getRaySpacing().x); wrf(b.getRaySpacing().y); wrf(b.getSourceDist().x); wrf(b.getSourceDist().y);
        const float3 de = b.getSpotIdxToGantry().getDelta(), of = b.getSpotIdxToGantry().getOffset();
        wrf(de.x); wrf(de.y); wrf(de.z); wrf(of.x); wrf(of.y); wrf(of.z);
        for (const Float3AffineTransform& t : {b.getGantryToImIdx(), b.getGantryToDoseIdx()}) {
            const Matrix3x3 m = t.getMatrix();
            for (const float3& r : {m.row0(), m.row1(), m.row2()}) { wrf(r.x); wrf(r.y); wrf(r.z); }
            wrf(t.getOffset().x); wrf(t.getOffset().y); wrf(t.getOffset().z);
        }
    }
    if (!l.nucWeightMatrix.empty() && p.nuclearCorr) {   // optional trailer: the LUTs of a NUCLEAR_CORR build
        wr("RTDNUCL1", 8); wri(p.nuclearCorr); wrv(l.nucWeightMatrix); wrv(l.nucSqSigmaMatrix);
    }
    fclose(f);
}

// The WATER_CUBE_TEST configuration (src/main.cu:39-43, 61-62, 74-99, 192-197).
void makeWaterCubeTest(Plan& p, const EnergyStruct& ciddData) {
    p.dim = p.doseDim = make_uint3(256, 256, 256);
    p.imageData.assign(size_t(256) * 256 * 256, 1000.0f);
    // imIdxToWorld = identity matrix, offset (-128,-128,-106); gantry == world  ->  gantryToImIdx = its inverse
    const Float3AffineTransform gantryToImIdx(Matrix3x3(), make_float3(128.0f, 128.0f, 106.0f));
    const float inf = std::numeric_limits<float>::infinity();
    const Float3IdxTransform fanIdxToFan(make_float3(3.0f, 3.0f, -1.0f), make_float3(-48.0f, -48.0f, 128.0f));
    const unsigned nLayers = 20;
    const uint3 beamDim = make_uint3(33, 33, nLayers);
    p.weightData.emplace_back(size_t(33) * 33 * nLayers);
    for (float& w : p.weightData.back()) w = 90.0f + 10.0f * float(rand()) / float(RAND_MAX);
    p.weightImages.emplace_back(new HostPinnedImage3D<float>(p.weightData.back().data(), beamDim));
    std::vector<float> energies(nLayers); std::vector<float2> sigmas(nLayers);
    float e = 118.12f; const float step = (172.51f - 118.12f) / float(nLayers - 1);
    for (unsigned i = 0; i < nLayers; ++i, e += step) {
        energies[i] = e;
        const float peak = vectorInterpolate(ciddData.peakDepths, findDecimalOrdered(ciddData.energiesPerU, e));
        sigmas[i].x = sigmas[i].y = 2.3f + 290.0f / (peak + 15.0f);
    }
    p.beams.emplace_back(p.weightImages.back().get(), energies, sigmas, make_float2(1.0f, 1.0f), 512u, make_float2(inf, inf), fanIdxToFan,
                         gantryToImIdx, gantryToImIdx);
}

// RTDCASE1 reader (layout: oracle/ref_harness/case_io.h, writer: raytracedicom_b200/cases.py Case.write).
void readCaseFile(const std::string& path, Plan& p) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("cannot open " + path);
    auto rd = [&](void* dst, size_t n) { if (n && fread(dst, 1, n, f) != n) { fclose(f); throw std::runtime_error("case file truncated: " + path); } };
    auto rdi = [&]() { int v; rd(&v, 4); return v; };
    auto rdf = [&]() { float v; rd(&v, 4); return v; };
    auto rdv = [&](std::vector<float>& v, size_t n) { v.resize(n); rd(v.data(), n * 4); };
    char magic[8]; rd(magic, 8);
    if (memcmp(magic, "RTDCASE1", 8) != 0) { fclose(f); throw std::runtime_error("bad magic in " + path); }
    int d[3]; rd(d, 12); p.dim = make_uint3(d[0], d[1], d[2]);
    rdv(p.imageData, size_t(d[0]) * d[1] * d[2]);
    rd(d, 12); p.doseDim = make_uint3(d[0], d[1], d[2]);
    EnergyStruct& l = p.luts;
    l.nEnergySamples = rdi(); l.nEnergies = rdi();
    rdv(l.energiesPerU, l.nEnergies); rdv(l.peakDepths, l.nEnergies); rdv(l.scaleFacts, l.nEnergies);
    rdv(l.ciddMatrix, size_t(l.nEnergies) * l.nEnergySamples);
    l.nDensitySamples = rdi(); l.densityScaleFact = rdf(); rdv(l.densityVector, l.nDensitySamples);
    l.nSpSamples = rdi(); l.spScaleFact = rdf(); rdv(l.spVector, l.nSpSamples);
    l.nRRlSamples = rdi(); l.rRlScaleFact = rdf(); rdv(l.rRlVector, l.nRRlSamples);
    p.haveLuts = true;
    const int nb = rdi();
    for (int b = 0; b < nb; ++b) {
        const int sx = rdi(), sy = rdi(), L = rdi(), steps = rdi();
        p.weightData.emplace_back(); rdv(p.weightData.back(), size_t(sx) * sy * L);
        p.weightImages.emplace_back(new HostPinnedImage3D<float>(p.weightData.back().data(), make_uint3(sx, sy, L)));
        std::vector<float> en, sg; rdv(en, L); rdv(sg, 2 * size_t(L));
        std::vector<float2> sig(L);
        for (int i = 0; i < L; ++i) sig[i] = make_float2(sg[2 * i], sg[2 * i + 1]);
        float rs[2], sd[2], de[3], of[3], im[9], iv[3], dm[9], dv[3];
        rd(rs, 8); rd(sd, 8); rd(de, 12); rd(of, 12); rd(im, 36); rd(iv, 12); rd(dm, 36); rd(dv, 12);
        p.beams.emplace_back(p.weightImages.back().get(), en, sig, make_float2(rs[0], rs[1]), unsigned(steps), make_float2(sd[0], sd[1]),
                             Float3IdxTransform(make_float3(de[0], de[1], de[2]), make_float3(of[0], of[1], of[2])),
                             Float3AffineTransform(Matrix3x3(im[0], im[1], im[2], im[3], im[4], im[5], im[6], im[7], im[8]), make_float3(iv[0], iv[1], iv[2])),
                             Float3AffineTransform(Matrix3x3(dm[0], dm[1], dm[2], dm[3], dm[4], dm[5], dm[6], dm[7], dm[8]), make_float3(dv[0], dv[1], dv[2])));
    }
    char tail[8];
    if (fread(tail, 1, 8, f) == 8 && memcmp(tail, "RTDNUCL1", 8) == 0) {
        p.nuclearCorr = rdi();
        rdv(l.nucWeightMatrix, size_t(l.nEnergies) * l.nEnergySamples); rdv(l.nucSqSigmaMatrix, size_t(l.nEnergies) * l.nEnergySamples);
    }
    fclose(f);
}

}  // namespace

int main(int argc, char** argv) {
    Config config(argc, argv);
    if (config.exit) return 1;
    try {
        Plan plan;
        const auto t0 = std::chrono::steady_clock::now();
        if (!config.case_file.empty()) {
            readCaseFile(config.case_file, plan);
        } else if (config.water_cube_test) {
            if (config.phys_data_directory.empty()) throw std::runtime_error("--water_cube_test needs --phys_data DIR/ (the reference's PHYS_DATA_DIRECTORY)");
            plan.luts = energyReader(config.phys_data_directory, true, config.nuclear_corr);
            plan.haveLuts = true;
            makeWaterCubeTest(plan, plan.luts);
        } else if (!config.ct_dir.empty() && !config.rtplan.empty()) {
            if (config.phys_data_directory.empty()) throw std::runtime_error("DICOM input needs --phys_data DIR/ (the reference's PHYS_DATA_DIRECTORY)");
            plan.luts = energyReader(config.phys_data_directory, false, config.nuclear_corr);
            plan.haveLuts = true;
            readDicom(config, plan);
        } else {
            throw std::runtime_error("no input: give --ct_dir DIR --rtplan FILE [--beams ...], --case FILE or --water_cube_test");
        }
        if (config.nuclear_corr && !plan.luts.nucWeightMatrix.empty() && !plan.nuclearCorr) plan.nuclearCorr = config.nuclear_corr;
        if (!config.dump_case.empty()) {
            writeCaseFile(config.dump_case, plan);
            std::cout << "Written " << config.dump_case << " (" << plan.beams.size() << " beam(s)); stopping before the dose calculation\n";
            return 0;
        }
        const double tRead = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        std::cout << "Read cumulative energy matrix: " << tRead << " seconds.\n\n";

        std::vector<float> doseData(size_t(plan.doseDim.x) * plan.doseDim.y * plan.doseDim.z, 0.0f);
        HostPinnedImage3D<float> doseVol(doseData.data(), plan.doseDim);
        HostPinnedImage3D<float> imVol(plan.imageData.data(), plan.dim);

        std::cout << "Executing code on GPU...\n\n";
        rtdose::ShimOptions so; so.device = config.gpu_id;
        if (config.nuclear_corr) {
            if (plan.luts.nucWeightMatrix.empty() || (plan.nuclearCorr && plan.nuclearCorr != config.nuclear_corr)) {      // --case input: the LUTs came with the case file, the nuclear tables come from --phys_data
                if (config.phys_data_directory.empty()) throw std::runtime_error("--nuclear_corr needs --phys_data DIR/ for the model's look-up table");
                const EnergyStruct n = energyReader(config.phys_data_directory, false, config.nuclear_corr);
                if (n.nEnergies != plan.luts.nEnergies || n.nEnergySamples != plan.luts.nEnergySamples) throw std::runtime_error("nuclear look-up table does not match the plan's IDD grid");
                plan.luts.nucWeightMatrix = n.nucWeightMatrix; plan.luts.nucSqSigmaMatrix = n.nucSqSigmaMatrix;
            }
            so.opts.nuclearCorr = config.nuclear_corr;
        }
        rtd_timing tm; so.timing = &tm;
        rtdose::cudaWrapperProtons(&imVol, &doseVol, plan.beams, plan.luts, std::cout, so);
        if (config.timing)
            std::cout << "    Trace " << tm.msTrace << " ms, setup " << tm.msSetup << " ms, spot->ray " << tm.msConv << " ms, IDD+sigma " << tm.msIddSigma
                      << " ms, radius " << tm.msRadius << " ms, superposition " << tm.msSuperp << " ms, transfer " << tm.msTransfer << " ms; upload "
                      << tm.msUpload << " ms, download " << tm.msDownload << " ms; " << tm.updates << " updates, " << tm.kernelLaunches << " launches, "
                      << tm.hostSyncs << " host syncs\n\n";
        std::cout << "Done!\n\n";

        const std::string out = config.output_directory + "/dose.dat";   // raw float32, x fastest (src/main.cu:211-214)
        std::ofstream fout(out.c_str(), std::ios::out | std::ios::binary);
        fout.write(reinterpret_cast<const char*>(doseData.data()), std::streamsize(doseData.size() * sizeof(float)));
        fout.close();
        std::cout << "Written " << out << " with size " << plan.doseDim.x << "x" << plan.doseDim.y << "x" << plan.doseDim.z << "\n\n";
        std::cout << "Max:" << *std::max_element(doseData.begin(), doseData.end()) << std::endl;
    } catch (const std::exception& e) {
        std::cerr << "Error: " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
