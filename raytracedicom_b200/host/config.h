#ifndef RTD_CONFIG_H
#define RTD_CONFIG_H
#include <string>
#include <vector>
// The reference's run-time options (src/config.h:20-32, src/config.cpp:13-51), parsed without CLI11, plus the
// switches that are compile-time -D flags in the reference (CMakeLists.txt:36-79) and two raw input modes beside
// the DICOM input (--ct_dir / --rtplan / --beams, read by host/dicom_reader.cpp without ITK/GDCM).
struct Config {
    Config(int argc, char** argv);
    unsigned short gpu_id{0};
    std::string ct_dir, rtplan, output_directory{"."}, config_file;
    std::vector<std::string> beams;
    bool exit{false};
    // additions
    bool water_cube_test{false};      // -DWATER_CUBE_TEST of the reference, at run time
    std::string case_file;            // RTDCASE1 file (raytracedicom_b200/cases.py): raw CT + BeamSettings
    std::string phys_data_directory;  // PHYS_DATA_DIRECTORY
    bool timing{false};               // FINE_GRAINED_TIMING-style stage report
    int nuclear_corr{0};              // NUCLEAR_CORR of the reference at run time: 0 off, 1 soukup, 2 fluka, 3 gauss_fit
    std::string dump_case;            // write the assembled plan (CT + BeamSettings + LUTs) as an RTDCASE1 file and stop: no GPU needed
};
#endif
