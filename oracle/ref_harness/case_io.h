// Reader for the "RTDCASE1" binary case files written by raytracedicom_b200/cases.py.
// TEST INFRASTRUCTURE ONLY (oracle / reference harness); the product library never includes this.
//
// Layout (little endian, no padding):
//   char[8]  "RTDCASE1"
//   int32[3] ct dims (nx,ny,nz)        float32[nx*ny*nz] CT as HU+1000, x fastest
//   int32[3] dose dims
//   LUTs: int32 nEnergySamples, nEnergies; float32 energies[nE], peakDepths[nE], scaleFacts[nE], cidd[nE*nS];
//         int32 nDens; float32 densScale; float32 dens[nDens];   same for SP; same for 1/X0
//   int32 nBeams; per beam:
//     int32 Sx,Sy,L,steps; float32 weights[L*Sy*Sx]; energies[L]; sigmas[2L] (x,y interleaved);
//     raySpacing[2]; sourceDist[2]; spotDelta[3]; spotOffset[3]; gantryToImIdx m[9], v[3]; gantryToDoseIdx m[9], v[3]
//   optional trailer (LUTs of a NUCLEAR_CORR build): char[8] "RTDNUCL1"; int32 model (1 SOUKUP, 2 FLUKA, 3 GAUSS_FIT);
//     float32 nucWeight[nE*nS]; float32 nucSqSigma[nE*nS]
#ifndef RTD_CASE_IO_H
#define RTD_CASE_IO_H
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <stdexcept>

struct CaseLuts {
    int nEnergySamples = 0, nEnergies = 0;
    std::vector<float> energies, peakDepths, scaleFacts, cidd;
    int nDens = 0; float densScale = 0; std::vector<float> dens;
    int nSp = 0; float spScale = 0; std::vector<float> sp;
    int nRrl = 0; float rrlScale = 0; std::vector<float> rrl;
    int nuclearModel = 0; std::vector<float> nucWeight, nucSqSigma;
};

struct CaseBeam {
    int Sx = 0, Sy = 0, L = 0, steps = 0;
    std::vector<float> weights, energies, sigmas;
    float raySpacing[2], sourceDist[2], spotDelta[3], spotOffset[3];
    float imM[9], imV[3], doseM[9], doseV[3];
};

struct CaseFile {
    int ctDims[3], doseDims[3];
    std::vector<float> ct;
    CaseLuts luts;
    std::vector<CaseBeam> beams;
};

namespace caseio {
inline void rd(FILE* f, void* p, size_t n) {
    if (n && fread(p, 1, n, f) != n) throw std::runtime_error("case file truncated");
}
inline int rdi(FILE* f) { int v; rd(f, &v, 4); return v; }
inline float rdf(FILE* f) { float v; rd(f, &v, 4); return v; }
inline void rdv(FILE* f, std::vector<float>& v, size_t n) { v.resize(n); rd(f, v.data(), n * 4); }
}

inline CaseFile readCase(const std::string& path) {
    using namespace caseio;
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("cannot open " + path);
    CaseFile c;
    char magic[8]; rd(f, magic, 8);
    if (memcmp(magic, "RTDCASE1", 8) != 0) throw std::runtime_error("bad magic in " + path);
    rd(f, c.ctDims, 12);
    rdv(f, c.ct, size_t(c.ctDims[0]) * c.ctDims[1] * c.ctDims[2]);
    rd(f, c.doseDims, 12);
    CaseLuts& l = c.luts;
    l.nEnergySamples = rdi(f); l.nEnergies = rdi(f);
    rdv(f, l.energies, l.nEnergies); rdv(f, l.peakDepths, l.nEnergies); rdv(f, l.scaleFacts, l.nEnergies);
    rdv(f, l.cidd, size_t(l.nEnergies) * l.nEnergySamples);
    l.nDens = rdi(f); l.densScale = rdf(f); rdv(f, l.dens, l.nDens);
    l.nSp = rdi(f); l.spScale = rdf(f); rdv(f, l.sp, l.nSp);
    l.nRrl = rdi(f); l.rrlScale = rdf(f); rdv(f, l.rrl, l.nRrl);
    int nb = rdi(f);
    c.beams.resize(nb);
    for (auto& b : c.beams) {
        b.Sx = rdi(f); b.Sy = rdi(f); b.L = rdi(f); b.steps = rdi(f);
        rdv(f, b.weights, size_t(b.L) * b.Sy * b.Sx);
        rdv(f, b.energies, b.L); rdv(f, b.sigmas, 2 * size_t(b.L));
        rd(f, b.raySpacing, 8); rd(f, b.sourceDist, 8); rd(f, b.spotDelta, 12); rd(f, b.spotOffset, 12);
        rd(f, b.imM, 36); rd(f, b.imV, 12); rd(f, b.doseM, 36); rd(f, b.doseV, 12);
    }
    char tail[8];
    if (fread(tail, 1, 8, f) == 8 && memcmp(tail, "RTDNUCL1", 8) == 0) {
        l.nuclearModel = rdi(f);
        rdv(f, l.nucWeight, size_t(l.nEnergies) * l.nEnergySamples); rdv(f, l.nucSqSigma, size_t(l.nEnergies) * l.nEnergySamples);
    }
    fclose(f);
    return c;
}

inline void writeRaw(const std::string& path, const void* p, size_t bytes) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    if (bytes && fwrite(p, 1, bytes, f) != bytes) { fclose(f); throw std::runtime_error("short write " + path); }
    fclose(f);
}
#endif
