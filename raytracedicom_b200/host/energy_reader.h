#ifndef RTD_ENERGY_READER_H
#define RTD_ENERGY_READER_H
#include <string>
#include "host_types.h"
// dataPath must end with a path separator, as in the reference (PHYS_DATA_DIRECTORY, CMakeLists.txt:75-79).
EnergyStruct energyReader(const std::string& dataPath, bool waterCubeTest = false);
#endif
