#ifndef RTD_ENERGY_READER_H
#define RTD_ENERGY_READER_H
#include <string>
#include "host_types.h"
// dataPath must end with a path separator, as in the reference (PHYS_DATA_DIRECTORY, CMakeLists.txt:75-79).
// nuclearCorr: 0 = none; 1 SOUKUP, 2 FLUKA, 3 GAUSS_FIT also load that model's nuclear_weights_and_sigmas_*.txt (energy_reader.cpp:103-162).
EnergyStruct energyReader(const std::string& dataPath, bool waterCubeTest = false, int nuclearCorr = 0);
#endif
