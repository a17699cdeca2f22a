// rtdose_shim.hpp - C++ drop-in for the reference's hot-path entry point on top of the C ABI (rtdose.h).
//
// The reference declares (src/kernel_wrapper.cuh:161, implemented src/kernel_wrapper.cu:381-1369)
//     void cudaWrapperProtons(HostPinnedImage3D<float>* const imVol, HostPinnedImage3D<float>* const doseVol,
//                             const std::vector<BeamSettings> beams, const EnergyStruct iddData, std::ostream& outStream);
// A maintainer replaces the call in src/main.cu:206 by
//     rtdose::cudaWrapperProtons(imVol, doseVol, beams, ciddData, std::cout);
// and links librtdose.so instead of kernel_wrapper.o / gpu_convolution_2d.o and the parameter classes.
//
// The function is a template over the reference's own value types, so this header needs none of the reference's
// headers and contains none of its code; it only uses their public accessors:
//     HostPinnedImage3D<float>::getImData(), getDims()                         (src/host_image_3d.cuh:38,44)
//     BeamSettings::getWeights(), getEnergies(), getSpotSigmas(), getRaySpacing(), getSteps(), getSourceDist(),
//                   getSpotIdxToGantry(), getGantryToImIdx(), getGantryToDoseIdx()        (src/beam_settings.h:32-90)
//     Float3IdxTransform::getDelta(), getOffset(); Float3AffineTransform::getMatrix(), getOffset(); Matrix3x3::row0..2()
//     EnergyStruct's public fields                                                        (src/energy_struct.h:13-38)
// Behaviour kept from the reference: dose is accumulated onto doseVol's content (cu:542); failures surface as
// std::runtime_error (cuda_errchk.cu:12-22; the reference's throw("...") of a const char* at cu:965 becomes a
// runtime_error with the same text); the log line "Total global execution time ..." (cu:1357-1360) is written.
// Behaviour NOT kept unless RTDOSE_SHIM_REFERENCE_OWNERSHIP is defined: the reference deletes imVol, doseVol and every
// beam's weight image and then calls cudaDeviceReset() (cu:856, 1366-1368).
#ifndef RTDOSE_SHIM_HPP
#define RTDOSE_SHIM_HPP

#include <ostream>
#include <stdexcept>
#include <string>
#include <vector>

#include "rtdose.h"

namespace rtdose {

namespace detail {
template <class Affine>
inline void packAffine(const Affine& a, float m[9], float v[3]) {
    const auto mat = a.getMatrix();
    const auto r0 = mat.row0(), r1 = mat.row1(), r2 = mat.row2();
    m[0] = r0.x; m[1] = r0.y; m[2] = r0.z; m[3] = r1.x; m[4] = r1.y; m[5] = r1.z; m[6] = r2.x; m[7] = r2.y; m[8] = r2.z;
    const auto o = a.getOffset();
    v[0] = o.x; v[1] = o.y; v[2] = o.z;
}

// EnergyStruct has nucWeightMatrix / nucSqSigmaMatrix only in the reference's NUCLEAR_CORR builds (src/energy_struct.h:33-36)
template <class E> inline auto nucWeight(const E& e, int) -> decltype(e.nucWeightMatrix.data()) { return e.nucWeightMatrix.empty() ? nullptr : e.nucWeightMatrix.data(); }
template <class E> inline const float* nucWeight(const E&, long) { return nullptr; }
template <class E> inline auto nucSqSigma(const E& e, int) -> decltype(e.nucSqSigmaMatrix.data()) { return e.nucSqSigmaMatrix.empty() ? nullptr : e.nucSqSigmaMatrix.data(); }
template <class E> inline const float* nucSqSigma(const E&, long) { return nullptr; }

struct Context {   // RAII: the context is destroyed on every exit path
    rtd_ctx* c = nullptr;
    explicit Context(int device) {
        const rtd_status s = rtd_create(&c, device);
        if (s == RTD_ERR_NO_DEVICE) throw std::runtime_error("cudaAssert: no CUDA-capable device is detected (rtdose has no CPU fallback)");
        if (s != RTD_OK) throw std::runtime_error("rtd_create failed");
    }
    ~Context() { rtd_destroy(c); }
    void check(rtd_status s) const { if (s != RTD_OK) throw std::runtime_error(rtd_last_error(c)); }
};
}  // namespace detail

// Options beyond the reference's signature (its --gpu_id is parsed but never used, src/config.cpp:13-15).
struct ShimOptions {
    int device = 0;
    rtd_opts opts;
    rtd_timing* timing = nullptr;   // optional: stage report, the counterpart of FINE_GRAINED_TIMING (cu:1291-1311)
    ShimOptions() {
        rtd_default_opts(&opts);
#ifdef NUCLEAR_CORR            // the reference's build option (CMakeLists.txt:57-79: SOUKUP 0, FLUKA 1, GAUSS_FIT 2) selects the run-time model
        opts.nuclearCorr = NUCLEAR_CORR + 1;
#endif
    }
};

template <class Image, class Beam, class Energy>
void cudaWrapperProtons(Image* const imVol, Image* const doseVol, std::vector<Beam> beams, const Energy& iddData, std::ostream& outStream,
                        const ShimOptions& so = ShimOptions()) {
    detail::Context ctx(so.device);

    rtd_luts l;
    l.nEnergySamples = iddData.nEnergySamples; l.nEnergies = iddData.nEnergies;
    l.energiesPerU = iddData.energiesPerU.data(); l.peakDepths = iddData.peakDepths.data();
    l.scaleFacts = iddData.scaleFacts.data(); l.ciddMatrix = iddData.ciddMatrix.data();
    l.nDensitySamples = iddData.nDensitySamples; l.densityScaleFact = iddData.densityScaleFact; l.densityVector = iddData.densityVector.data();
    l.nSpSamples = iddData.nSpSamples; l.spScaleFact = iddData.spScaleFact; l.spVector = iddData.spVector.data();
    l.nRRlSamples = iddData.nRRlSamples; l.rRlScaleFact = iddData.rRlScaleFact; l.rRlVector = iddData.rRlVector.data();
    l.nucWeightMatrix = detail::nucWeight(iddData, 0); l.nucSqSigmaMatrix = detail::nucSqSigma(iddData, 0);
    ctx.check(rtd_set_luts(ctx.c, &l));

    const auto id = imVol->getDims();
    const auto dd = doseVol->getDims();
    ctx.check(rtd_set_dose_grid(ctx.c, (int)dd.x, (int)dd.y, (int)dd.z));

    std::vector<rtd_beam> rb(beams.size());
    std::vector<std::vector<float>> sig(beams.size());
    for (size_t i = 0; i < beams.size(); ++i) {
        Beam& b = beams[i];
        rtd_beam& o = rb[i];
        const auto wd = b.getWeights()->getDims();
        o.weights = b.getWeights()->getImData();
        o.sx = (int)wd.x; o.sy = (int)wd.y; o.nLayers = (int)wd.z;
        o.energies = b.getEnergies().data();
        sig[i].resize(2 * b.getSpotSigmas().size());
        for (size_t k = 0; k < b.getSpotSigmas().size(); ++k) { sig[i][2 * k] = b.getSpotSigmas()[k].x; sig[i][2 * k + 1] = b.getSpotSigmas()[k].y; }
        o.sigmas = sig[i].data();
        const auto rs = b.getRaySpacing(); o.raySpacing[0] = rs.x; o.raySpacing[1] = rs.y;
        o.steps = b.getSteps();
        const auto sd = b.getSourceDist(); o.sourceDist[0] = sd.x; o.sourceDist[1] = sd.y;
        const auto t = b.getSpotIdxToGantry();
        const auto de = t.getDelta(), of = t.getOffset();
        o.spotDelta[0] = de.x; o.spotDelta[1] = de.y; o.spotDelta[2] = de.z;
        o.spotOffset[0] = of.x; o.spotOffset[1] = of.y; o.spotOffset[2] = of.z;
        detail::packAffine(b.getGantryToImIdx(), o.gantryToImIdxM, o.gantryToImIdxV);
        detail::packAffine(b.getGantryToDoseIdx(), o.gantryToDoseIdxM, o.gantryToDoseIdxV);
        o.layerBegin = 0; o.layerEnd = 0;
    }

    rtd_timing tm;
    // one call: CT and dose sub-volumes travel inside (the reference uploads / downloads both volumes whole, cu:420-451, 542, 1357)
    const rtd_status s = rtd_plan_dose(ctx.c, imVol->getImData(), (int)id.x, (int)id.y, (int)id.z, rb.data(), (int)rb.size(),
                                       doseVol->getImData(), &so.opts, &tm);
    if (s == RTD_ERR_RADIUS) throw std::runtime_error("Found larger than allowed kernel superposition radius");   // cu:965
    ctx.check(s);
    if (so.timing) *so.timing = tm;
    outStream << "    Total global execution time (excluding GPU initialisation): " << tm.msTotal << " ms.\n\n";   // cu:1357-1360

#ifdef RTDOSE_SHIM_REFERENCE_OWNERSHIP
    for (auto& b : beams) delete b.getWeights();   // cu:856
    delete imVol; delete doseVol;                  // cu:1366-1367 (cudaDeviceReset() is deliberately not reproduced)
#endif
}

}  // namespace rtdose
#endif  // RTDOSE_SHIM_HPP
