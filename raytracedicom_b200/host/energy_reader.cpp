// energyReader: loads the physics look-up tables from whitespace-separated text files into an EnergyStruct.
// Same file names, header fields and error text as the reference's loader (src/energy_reader.cpp:11-101); the
// compile-time WATER_CUBE_TEST switch of the reference (cpp:77-84) is a run-time argument here.
#include "energy_reader.h"

#include <cmath>
#include <fstream>
#include <stdexcept>

namespace {
std::ifstream openTable(const std::string& dir, const char* name) {
    std::ifstream f(dir + name);
    if (!f) throw std::runtime_error("Failed to open " + dir + name);
    return f;
}
void readValues(std::ifstream& f, std::vector<float>& dst, size_t n, const char* what) {
    dst.resize(n);
    for (float& v : dst)
        if (!(f >> v)) throw std::runtime_error(std::string("Truncated look-up table: ") + what);
}
// "<count> <scale>" header followed by <count> samples
void readScaledVector(const std::string& dir, const char* name, int& n, float& scale, std::vector<float>& v) {
    std::ifstream f = openTable(dir, name);
    if (!(f >> n >> scale) || n <= 0) throw std::runtime_error(std::string("Bad header in ") + name);
    readValues(f, v, size_t(n), name);
}
}  // namespace

EnergyStruct energyReader(const std::string& dataPath, bool waterCubeTest, int nuclearCorr) {
    EnergyStruct e;
    {
        const char* name = "proton_cumul_ddd_data.txt";
        std::ifstream f = openTable(dataPath, name);
        if (!(f >> e.nEnergySamples >> e.nEnergies) || e.nEnergySamples <= 0 || e.nEnergies <= 0)
            throw std::runtime_error(std::string("Bad header in ") + name);
        const size_t nE = size_t(e.nEnergies);
        readValues(f, e.energiesPerU, nE, name);
        readValues(f, e.peakDepths, nE, name);
        readValues(f, e.scaleFacts, nE, name);
        readValues(f, e.ciddMatrix, nE * size_t(e.nEnergySamples), name);
    }
    readScaledVector(dataPath, "density_Schneider2000_adj.txt", e.nDensitySamples, e.densityScaleFact, e.densityVector);
    readScaledVector(dataPath, "HU_to_SP_H&N_adj.txt", e.nSpSamples, e.spScaleFact, e.spVector);
    readScaledVector(dataPath, waterCubeTest ? "radiation_length_inc_water.txt" : "radiation_length.txt", e.nRRlSamples, e.rRlScaleFact, e.rRlVector);
    if (nuclearCorr) {
        // energy_reader.cpp:103-162: the header repeats the IDD file's grid and must agree with it
        if (nuclearCorr < 1 || nuclearCorr > 3) throw std::runtime_error("nuclearCorr must be 1 (Soukup), 2 (Fluka) or 3 (fit)");
        const char* names[] = {"nuclear_weights_and_sigmas_Soukup.txt", "nuclear_weights_and_sigmas_Fluka.txt", "nuclear_weights_and_sigmas_fit.txt"};
        const std::string name = names[nuclearCorr - 1];
        std::ifstream f = openTable(dataPath, name.c_str());
        int nSamples = 0, nEnergies = 0;
        if (!(f >> nSamples >> nEnergies) || nSamples != e.nEnergySamples || nEnergies != e.nEnergies)
            throw std::runtime_error("Number of samples or energies in " + name + " different from proton_cumul_ddd_data.txt");
        auto sameAs = [&](const std::vector<float>& ref, const char* what) {
            for (float r : ref) { float v; if (!(f >> v) || std::fabs(r - v) > 0.01f) throw std::runtime_error(std::string(what) + " in " + name + " different from proton_cumul_ddd_data.txt"); }
        };
        sameAs(e.energiesPerU, "Energies"); sameAs(e.peakDepths, "Peak depths"); sameAs(e.scaleFacts, "Scale facts");
        const size_t n = size_t(e.nEnergies) * size_t(e.nEnergySamples);
        readValues(f, e.nucWeightMatrix, n, name.c_str());
        readValues(f, e.nucSqSigmaMatrix, n, name.c_str());
    }
    return e;
}
