// energyReader: loads the physics look-up tables from whitespace-separated text files into an EnergyStruct.
// Same file names, header fields and error text as the reference's loader (src/energy_reader.cpp:11-101); the
// compile-time WATER_CUBE_TEST switch of the reference (cpp:77-84) is a run-time argument here.
#include "energy_reader.h"

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

EnergyStruct energyReader(const std::string& dataPath, bool waterCubeTest) {
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
    return e;
}
