// Host-side value types of the dose path, written from scratch with the reference's names and accessors so that
// main.cpp reads like the reference's main.cu and include/rtdose_shim.hpp accepts them unchanged:
//   HostPinnedImage3D<T>   <- src/host_image_3d.cuh:16-48   (non-owning view; pinning is left to the library)
//   Matrix3x3, Float3AffineTransform, Float3IdxTransform <- src/matrix_3x3.cuh, float3_affine_transform.cuh, float3_idx_transform.cuh
//   BeamSettings           <- src/beam_settings.h:17-112
//   EnergyStruct           <- src/energy_struct.h:13-38
// Plain C++17, no CUDA headers: the device work is behind the C ABI (include/rtdose.h).
#ifndef RTD_HOST_TYPES_H
#define RTD_HOST_TYPES_H
#include <vector>

struct float2 { float x, y; };
struct float3 { float x, y, z; };
struct uint3 { unsigned int x, y, z; };
inline float2 make_float2(float x, float y) { return float2{x, y}; }
inline float3 make_float3(float x, float y, float z) { return float3{x, y, z}; }
inline uint3 make_uint3(unsigned x, unsigned y, unsigned z) { return uint3{x, y, z}; }

template <typename T>
class HostPinnedImage3D {
public:
    HostPinnedImage3D(T* const imagePtr, const uint3 dimensions) : imPtr(imagePtr), dims(dimensions) {}
    T* getImData() const { return imPtr; }
    uint3 getDims() const { return dims; }
    size_t size() const { return size_t(dims.x) * dims.y * dims.z; }
private:
    T* const imPtr;
    const uint3 dims;
};

class Matrix3x3 {
public:
    Matrix3x3() : r0{1, 0, 0}, r1{0, 1, 0}, r2{0, 0, 1} {}
    Matrix3x3(float a00, float a01, float a02, float a10, float a11, float a12, float a20, float a21, float a22)
        : r0{a00, a01, a02}, r1{a10, a11, a12}, r2{a20, a21, a22} {}
    float3 row0() const { return r0; }
    float3 row1() const { return r1; }
    float3 row2() const { return r2; }
private:
    float3 r0, r1, r2;
};

class Float3AffineTransform {   // p -> m*p + v
public:
    Float3AffineTransform() : v{0, 0, 0} {}
    Float3AffineTransform(const Matrix3x3& matrix, const float3 offset) : m(matrix), v(offset) {}
    Matrix3x3 getMatrix() const { return m; }
    float3 getOffset() const { return v; }
private:
    Matrix3x3 m;
    float3 v;
};

class Float3IdxTransform {      // idx -> idx*delta + offset
public:
    Float3IdxTransform() : d{1, 1, 1}, o{0, 0, 0} {}
    Float3IdxTransform(const float3 delta, const float3 offset) : d(delta), o(offset) {}
    float3 getDelta() const { return d; }
    float3 getOffset() const { return o; }
private:
    float3 d, o;
};

class BeamSettings {
public:
    BeamSettings(HostPinnedImage3D<float>* const spotWeights, const std::vector<float>& beamEnergies, const std::vector<float2>& spotSigmas,
                 const float2 raySpacing, const unsigned int tracerSteps, const float2 sourceDist, const Float3IdxTransform spotIdxToGantry,
                 const Float3AffineTransform& gantryToImIdx, const Float3AffineTransform& gantryToDoseIdx)
        : weights_(spotWeights), layerEnergies_(beamEnergies), layerSigmas_(spotSigmas), rayPitch_(raySpacing), nSteps_(tracerSteps),
          sourceToIso_(sourceDist), spotToGantry_(spotIdxToGantry), toImage_(gantryToImIdx), toDose_(gantryToDoseIdx) {}
    HostPinnedImage3D<float>* getWeights() { return weights_; }
    std::vector<float>& getEnergies() { return layerEnergies_; }
    std::vector<float2>& getSpotSigmas() { return layerSigmas_; }
    float2 getRaySpacing() const { return rayPitch_; }
    unsigned int getSteps() const { return nSteps_; }
    float2 getSourceDist() const { return sourceToIso_; }
    Float3IdxTransform getSpotIdxToGantry() const { return spotToGantry_; }
    Float3AffineTransform getGantryToImIdx() const { return toImage_; }
    Float3AffineTransform getGantryToDoseIdx() const { return toDose_; }
private:
    HostPinnedImage3D<float>* weights_;
    std::vector<float> layerEnergies_;
    std::vector<float2> layerSigmas_;
    float2 rayPitch_;
    unsigned int nSteps_;
    float2 sourceToIso_;
    Float3IdxTransform spotToGantry_;
    Float3AffineTransform toImage_, toDose_;
};

struct EnergyStruct {
    int nEnergySamples = 0, nEnergies = 0;
    std::vector<float> energiesPerU, peakDepths, scaleFacts, ciddMatrix;
    int nDensitySamples = 0; float densityScaleFact = 0; std::vector<float> densityVector;
    int nSpSamples = 0; float spScaleFact = 0; std::vector<float> spVector;
    int nRRlSamples = 0; float rRlScaleFact = 0; std::vector<float> rRlVector;
    std::vector<float> nucWeightMatrix, nucSqSigmaMatrix;   // the reference's NUCLEAR_CORR builds only (energy_struct.h:33-36); empty otherwise
};

#endif
