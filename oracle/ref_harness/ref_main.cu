// End-to-end harness for the UNMODIFIED reference dose entry point
// cudaWrapperProtons() (/root/reference/src/kernel_wrapper.cu:381, declared kernel_wrapper.cuh:161).
// It plays the role of the reference's main() (/root/reference/src/main.cu:20-252) but fills the
// BeamSettings from an RTDCASE1 file instead of the WATER_CUBE_TEST constants / the stub DICOM path.
// TEST + BASELINE INFRASTRUCTURE ONLY: built into oracle/_ref/ref_dose by oracle/ref_harness/Makefile.
//
// usage: ref_dose <case.rtdcase> <out_dose.f32> [repeat]
// Prints one line "REF_TIMING total_ms=<reference's own global timer> wall_ms=<std::chrono around the call>".
#include <chrono>
#include <iostream>
#include <sstream>
#include "kernel_wrapper.cuh"
#include "float3_affine_transform.cuh"
#include "float3_idx_transform.cuh"
#include "vector_functions.hpp"
#include "case_io.h"

static EnergyStruct toEnergyStruct(const CaseLuts& l) {
    EnergyStruct e;
    e.nEnergySamples = l.nEnergySamples; e.nEnergies = l.nEnergies;
    e.energiesPerU = l.energies; e.peakDepths = l.peakDepths; e.scaleFacts = l.scaleFacts; e.ciddMatrix = l.cidd;
    e.nDensitySamples = l.nDens; e.densityScaleFact = l.densScale; e.densityVector = l.dens;
    e.nSpSamples = l.nSp; e.spScaleFact = l.spScale; e.spVector = l.sp;
    e.nRRlSamples = l.nRrl; e.rRlScaleFact = l.rrlScale; e.rRlVector = l.rrl;
#ifdef NUCLEAR_CORR
    if (l.nuclearModel != NUCLEAR_CORR + 1) throw std::runtime_error("case file does not carry the nuclear LUTs of this build's NUCLEAR_CORR model");
    e.nucWeightMatrix = l.nucWeight; e.nucSqSigmaMatrix = l.nucSqSigma;
#endif
    return e;
}

int main(int argc, char** argv) {
    if (argc < 3) { std::cerr << "usage: ref_dose case out_dose.f32\n"; return 2; }
    try {
        CaseFile c = readCase(argv[1]);
        EnergyStruct es = toEnergyStruct(c.luts);
        const uint3 ctDim = make_uint3(c.ctDims[0], c.ctDims[1], c.ctDims[2]);
        const uint3 doseDim = make_uint3(c.doseDims[0], c.doseDims[1], c.doseDims[2]);
        std::vector<float> dose(size_t(doseDim.x) * doseDim.y * doseDim.z, 0.0f);
        cudaFree(0);
        HostPinnedImage3D<float>* doseVol = new HostPinnedImage3D<float>(dose.data(), doseDim);
        HostPinnedImage3D<float>* imVol = new HostPinnedImage3D<float>(c.ct.data(), ctDim);
        std::vector<BeamSettings> beams;
        for (auto& b : c.beams) {
            HostPinnedImage3D<float>* w = new HostPinnedImage3D<float>(b.weights.data(), make_uint3(b.Sx, b.Sy, b.L));
            std::vector<float2> sig(b.L);
            for (int i = 0; i < b.L; ++i) sig[i] = make_float2(b.sigmas[2 * i], b.sigmas[2 * i + 1]);
            Float3IdxTransform sitg(make_float3(b.spotDelta[0], b.spotDelta[1], b.spotDelta[2]),
                                    make_float3(b.spotOffset[0], b.spotOffset[1], b.spotOffset[2]));
            Float3AffineTransform gtii(Matrix3x3(b.imM[0], b.imM[1], b.imM[2], b.imM[3], b.imM[4], b.imM[5], b.imM[6], b.imM[7], b.imM[8]),
                                       make_float3(b.imV[0], b.imV[1], b.imV[2]));
            Float3AffineTransform gtdi(Matrix3x3(b.doseM[0], b.doseM[1], b.doseM[2], b.doseM[3], b.doseM[4], b.doseM[5], b.doseM[6], b.doseM[7], b.doseM[8]),
                                       make_float3(b.doseV[0], b.doseV[1], b.doseV[2]));
            beams.push_back(BeamSettings(w, b.energies, sig, make_float2(b.raySpacing[0], b.raySpacing[1]), (unsigned)b.steps,
                                         make_float2(b.sourceDist[0], b.sourceDist[1]), sitg, gtii, gtdi));
        }
        std::ostringstream log;
        auto t0 = std::chrono::steady_clock::now();
        cudaWrapperProtons(imVol, doseVol, beams, es, log);   // deletes imVol, doseVol, weights; resets the device
        auto t1 = std::chrono::steady_clock::now();
        const std::string s = log.str();
        std::cout << s;
        double totalMs = -1.0;
        size_t p = s.find("Total global execution time");
        if (p != std::string::npos) {
            size_t q = s.find(':', p);
            if (q != std::string::npos) totalMs = atof(s.c_str() + q + 1);
        }
        std::cout << "REF_TIMING total_ms=" << totalMs
                  << " wall_ms=" << std::chrono::duration<double, std::milli>(t1 - t0).count() << std::endl;
        writeRaw(argv[2], dose.data(), dose.size() * sizeof(float));
        float mx = 0; for (float v : dose) if (v > mx) mx = v;
        std::cout << "REF_MAX " << mx << std::endl;
    } catch (const std::exception& e) {
        std::cerr << "ref_dose failed: " << e.what() << std::endl; return 1;
    } catch (const char* e) {
        std::cerr << "ref_dose failed: " << e << std::endl; return 1;
    }
    return 0;
}
