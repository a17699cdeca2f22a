// Command-line / config-file options with the reference's names (src/config.cpp:13-51).
#include "config.h"

#include <fstream>
#include <iostream>
#include <map>
#include <sstream>

namespace {
const char* kUsage =
    "RayTraceDicom (B200 engine)\n"
    "  --gpu_id N               GPU to use (default 0)\n"
    "  --ct_dir DIR             directory with the CT DICOM series\n"
    "  --rtplan FILE            RT-Ion plan\n"
    "  --beams A,B,...          beam names to include\n"
    "  --output_directory DIR   where dose.dat is written (default .)\n"
    "  --config_file FILE       key = value lines with the option names above\n"
    "  --water_cube_test        the reference's WATER_CUBE_TEST configuration\n"
    "  --case FILE              raw case file (RTDCASE1) instead of DICOM input\n"
    "  --phys_data DIR/         directory of the physics look-up tables\n"
    "  --timing                 print the stage timing report\n"
    "  --nuclear_corr MODEL     the reference's NUCLEAR_CORR build option: soukup | fluka | gauss_fit (needs --phys_data)\n"
    "  --dump_case FILE         write the assembled plan as an RTDCASE1 file and stop (no GPU needed)\n";

void splitList(const std::string& s, std::vector<std::string>& out) {
    std::stringstream ss(s);
    for (std::string tok; std::getline(ss, tok, ',');)
        if (!tok.empty()) out.push_back(tok);
}
}  // namespace

Config::Config(int argc, char** argv) {
    std::map<std::string, std::string> kv;
    auto set = [&](const std::string& key, const std::string& val) {
        if (key == "gpu_id") gpu_id = static_cast<unsigned short>(std::stoi(val));
        else if (key == "ct_dir") ct_dir = val;
        else if (key == "rtplan") rtplan = val;
        else if (key == "beams") splitList(val, beams);
        else if (key == "output_directory") output_directory = val;
        else if (key == "case") case_file = val;
        else if (key == "phys_data") phys_data_directory = val;
        else if (key == "dump_case") dump_case = val;
        else if (key == "nuclear_corr") {
            nuclear_corr = val == "soukup" ? 1 : val == "fluka" ? 2 : (val == "gauss_fit" || val == "fit") ? 3 : (val == "off" || val == "0") ? 0 : -1;
            if (nuclear_corr < 0) { std::cerr << "Unknown nuclear model: " << val << "\n" << kUsage; exit = true; }
        }
        else { std::cerr << "Unknown option: " << key << "\n" << kUsage; exit = true; }
    };
    for (int i = 1; i < argc && !exit; ++i) {
        std::string a = argv[i];
        if (a == "-h" || a == "--help") { std::cout << kUsage; exit = true; break; }
        if (a.rfind("--", 0) != 0) { std::cerr << "Unexpected argument: " << a << "\n" << kUsage; exit = true; break; }
        a = a.substr(2);
        if (a == "water_cube_test") { water_cube_test = true; continue; }
        if (a == "timing") { timing = true; continue; }
        std::string val;
        const size_t eq = a.find('=');
        if (eq != std::string::npos) { val = a.substr(eq + 1); a = a.substr(0, eq); }
        else if (i + 1 < argc) val = argv[++i];
        else { std::cerr << "Missing value for --" << a << "\n"; exit = true; break; }
        if (a == "config_file") config_file = val; else set(a, val);
    }
    if (!exit && !config_file.empty()) {
        std::ifstream f(config_file);
        if (!f) { std::cerr << "Cannot open config file " << config_file << "\n"; exit = true; }
        for (std::string line; !exit && std::getline(f, line);) {
            const size_t hash = line.find('#'); if (hash != std::string::npos) line.erase(hash);
            const size_t eq = line.find('='); if (eq == std::string::npos) continue;
            auto trim = [](std::string s) { const char* ws = " \t\r\""; s.erase(0, s.find_first_not_of(ws)); s.erase(s.find_last_not_of(ws) + 1); return s; };
            set(trim(line.substr(0, eq)), trim(line.substr(eq + 1)));
        }
    }
    if (exit) return;
    std::cout << "Configuration:\n  gpu_id: " << gpu_id << "\n  ct_dir: " << ct_dir << "\n  rtplan: " << rtplan << "\n  output_directory: " << output_directory
              << "\n  beams:";
    for (const auto& b : beams) std::cout << ' ' << b;
    std::cout << "\n";
}
