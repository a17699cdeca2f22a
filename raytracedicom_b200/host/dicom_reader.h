// DICOM input of the host driver without ITK / GDCM / rti (none exist in this image):
//   itk_reader()      - same signature and results as the reference's ITK series reader (src/dicom_reader.h:24,
//                       src/dicom_reader.cpp:15-129): slices of the first series ordered along the slice normal, short
//                       pixels + 1000 as float, x fastest, returns imIdxToWorld = direction * spacing with the origin.
//   readIonPlan()     - the RT Ion Plan attributes rti touches (rti_beam_module_ion.hpp:86-104,
//                       rti_treatment_machine_ion.hpp:54-93): per beam the angles, isocentre, virtual SADs and the spots of
//                       every EVEN control point (odd ones carry zero weights and are skipped).
//   beamFromPlan()    - plan beam + CT geometry -> BeamSettings; the Python mirror with the full
//                       description of the conventions is raytracedicom_b200/plan.py.
// Parser scope: Part-10 files, Implicit or Explicit VR Little Endian, nested sequences of defined / undefined length,
// uncompressed 16-bit pixel data (what generate_water_cube.py and treatment planning systems export for CT / plans).
#ifndef RTD_DICOM_READER_H
#define RTD_DICOM_READER_H
#include <ctime>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "host_types.h"

namespace dcm {

struct Element;
using DataSet = std::map<unsigned, Element>;   // key = group << 16 | element
struct Element {
    std::string vr;                 // "" when unknown (implicit VR, tag not in the dictionary)
    std::vector<unsigned char> raw; // value bytes (empty for sequences)
    std::vector<DataSet> items;     // sequence items
};
constexpr unsigned tag(unsigned g, unsigned e) { return (g << 16) | e; }

DataSet readFile(const std::string& path);                       // throws std::runtime_error
bool has(const DataSet& ds, unsigned t);
std::string str(const DataSet& ds, unsigned t);                  // trimmed string value ("" if absent)
std::vector<double> numbers(const DataSet& ds, unsigned t);      // DS / IS / FL / FD / US / SS / UL / SL as doubles
const std::vector<DataSet>& seq(const DataSet& ds, unsigned t);  // throws if absent

}  // namespace dcm

Float3AffineTransform itk_reader(const std::string& imagePath, std::vector<float>& imageData, unsigned int& N, uint3& dim, clock_t& t);

struct PlanLayer { float energy; float fwhmX, fwhmY; std::vector<float> x, y, w; };
struct PlanBeam {
    std::string name;
    float gantryAngle = 0, collimatorAngle = 0, couchAngle = 0;
    float3 isocenter{0, 0, 0};
    float2 sad{0, 0};
    std::vector<PlanLayer> layers;
};
std::vector<PlanBeam> readIonPlan(const std::string& path);

// Owned storage of one BeamSettings built from a plan beam.
struct PlanBeamSettings {
    std::vector<float> weights;                       // [L][sy][sx]
    std::unique_ptr<HostPinnedImage3D<float>> weightImage;
    std::unique_ptr<BeamSettings> beam;
};
// imIdxToWorld: what itk_reader returned; dim: CT dims. Ray spacing (1,1) mm, 1 mm steps (src/main.cu:192, 57-64).
std::unique_ptr<PlanBeamSettings> beamFromPlan(const PlanBeam& pb, const Float3AffineTransform& imIdxToWorld, uint3 dim);

#endif
