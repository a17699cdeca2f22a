// Host-side geometry and LUT-index arithmetic of the dose path, in IEEE fp32 with the reference's operation
// order so that every integer it feeds (ray-grid size, step cut-offs, transfer bounding box) is bit-identical.
// Compile with -ffp-contract=off (the reference's host code is built for baseline x86-64: no FMA contraction).
//
// Restated from (all under /root/reference/src):
//   matrix_3x3.cu:20-53, float3_affine_transform.cu:16-24, float3_idx_transform.cu:17-21,
//   float3_from_fan_transform.cu:22-42, density_and_sp_tracer_params.cu:19-30,
//   fill_idd_and_sigma_params.cu:12-26,74-83, transfer_param_struct_div3.cu:9-20,
//   vector_find.h:62-145, vector_interpolate.h:19-30, kernel_wrapper.cu:612-663,792-849,1185-1210
#pragma once
#include <cmath>
#include <vector>
#include <algorithm>

namespace rtd {

#ifdef __CUDACC__
#define RTD_HD __host__ __device__
#else
#define RTD_HD
#endif
struct V3 { float x, y, z; };
inline RTD_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
inline V3 add(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 sub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 mul(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 mul(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
inline V3 div(V3 a, V3 b) { return v3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline V3 div(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

struct M3 { V3 r0, r1, r2; };
inline M3 m3(const float* p) { M3 m; m.r0 = v3(p[0], p[1], p[2]); m.r1 = v3(p[3], p[4], p[5]); m.r2 = v3(p[6], p[7], p[8]); return m; }
inline V3 mulMV(const M3& m, V3 a) { return v3(dot(m.r0, a), dot(m.r1, a), dot(m.r2, a)); }
inline M3 transpose(const M3& m) { M3 t; t.r0 = v3(m.r0.x, m.r1.x, m.r2.x); t.r1 = v3(m.r0.y, m.r1.y, m.r2.y); t.r2 = v3(m.r0.z, m.r1.z, m.r2.z); return t; }
inline float det(const M3& m) {
    return (m.r0.x * (m.r1.y * m.r2.z - m.r1.z * m.r2.y) - m.r0.y * (m.r1.x * m.r2.z - m.r1.z * m.r2.x) + m.r0.z * (m.r1.x * m.r2.y - m.r1.y * m.r2.x));
}
// matrix_3x3.cu:48-53: adjugate * diag(1/det); the diagonal product adds exact zeros (x*s + y*0 + z*0).
inline M3 inverse(const M3& m) {
    const float s = float(1. / det(m));
    const V3 a0 = v3(m.r1.y * m.r2.z - m.r1.z * m.r2.y, m.r0.z * m.r2.y - m.r0.y * m.r2.z, m.r0.y * m.r1.z - m.r0.z * m.r1.y);
    const V3 a1 = v3(m.r1.z * m.r2.x - m.r1.x * m.r2.z, m.r0.x * m.r2.z - m.r0.z * m.r2.x, m.r0.z * m.r1.x - m.r0.x * m.r1.z);
    const V3 a2 = v3(m.r1.x * m.r2.y - m.r1.y * m.r2.x, m.r0.y * m.r2.x - m.r0.x * m.r2.y, m.r0.x * m.r1.y - m.r0.y * m.r1.x);
    const V3 c0 = v3(s, 0.f, 0.f), c1 = v3(0.f, s, 0.f), c2 = v3(0.f, 0.f, s);   // columns of diag(s)
    M3 r;
    r.r0 = v3(dot(a0, c0), dot(a0, c1), dot(a0, c2));
    r.r1 = v3(dot(a1, c0), dot(a1, c1), dot(a1, c2));
    r.r2 = v3(dot(a2, c0), dot(a2, c1), dot(a2, c2));
    return r;
}

struct Affine { M3 m; V3 v; };
inline V3 transformPoint(const Affine& a, V3 p) { return add(mulMV(a.m, p), a.v); }
inline Affine inverse(const Affine& a) { Affine r; r.m = inverse(a.m); r.v = mulMV(inverse(a.m), mul(a.v, -1.f)); return r; }

struct IdxT { V3 delta, offset; };
inline IdxT inverse(const IdxT& t) { IdxT r; r.delta = div(v3(1.f, 1.f, 1.f), t.delta); r.offset = div(mul(t.offset, -1.f), t.delta); return r; }

// float3_from_fan_transform.cu:37-42
inline V3 fromFan(const IdxT& fitf, const float dist[2], const Affine& gtii, V3 fanIdx) {
    V3 p = add(mul(fanIdx, fitf.delta), fitf.offset);
    p.x *= 1.0f - p.z / dist[0];
    p.y *= 1.0f - p.z / dist[1];
    return transformPoint(gtii, p);
}

// ---- ordered-list helpers ------------------------------------------------------------------------------
inline int findFirstLargerOrdered(const float* list, int n, float value) {     // vector_find.h:62-82
    int upper = n - 1, lower = 0;
    if (list[n - 1] <= value) return upper;
    if (list[0] > value) return 0;
    while (upper - lower > 1) { int pivot = (upper + lower) / 2; if (list[pivot] <= value) lower = pivot; else upper = pivot; }
    return lower + 1;
}
inline int findLastSmallerOrEqOrdered(const float* list, int n, float value) {  // vector_find.h:96-116
    int upper = n - 1, lower = 0;
    if (list[n - 1] <= value) return upper;
    if (list[0] > value) return -1;
    while (upper - lower > 1) { int pivot = (upper + lower) / 2; if (list[pivot] <= value) lower = pivot; else upper = pivot; }
    return lower;
}
inline float findDecimalOrdered(const float* list, int n, float value) {        // vector_find.h:130-145
    if (value >= list[n - 1]) return float(n - 1);
    if (value < list[0]) return 0.0f;
    unsigned f = (unsigned)findLastSmallerOrEqOrdered(list, n, value);
    float corr = (value - list[f]) / (list[f + 1] - list[f]);
    return float(f) + corr;
}
inline float vectorInterpolate(const float* list, int n, float idx) {          // vector_interpolate.h:19-30
    if (idx <= 0.0f) return list[0];
    if (idx >= float(n - 1)) return list[n - 1];
    float ip; float dec = std::modf(idx, &ip);
    unsigned f = (unsigned)ip;
    float corr = (list[f + 1] - list[f]) * dec;
    return list[f] + corr;
}
inline int roundTo(int val, int multiple) { return ((val + multiple - 1) / multiple) * multiple; }   // kernel_wrapper.cu:45-48

// ---- device parameter blocks (plain floats; consumed by the kernels in kernels.cuh) ---------------------
struct TraceParams {          // DensityAndSpTracerParams, density_and_sp_tracer_params.cuh:70-79
    float densityScale, spScale; unsigned steps;
    V3 coefOffset, coefIdxI, coefIdxJ, transl, corner, delta; float distX, distY;
};
inline TraceParams makeTraceParams(float densScale, float spScale, unsigned steps, const IdxT& fitf, const float dist[2], const Affine& gtii) {
    TraceParams p; p.densityScale = densScale; p.spScale = spScale; p.steps = steps;
    p.distX = dist[0]; p.distY = dist[1]; p.corner = fitf.offset; p.delta = fitf.delta;
    const M3 t = transpose(gtii.m);
    p.coefOffset = sub(sub(t.r2, div(mul(t.r0, p.corner.x), p.distX)), div(mul(t.r1, p.corner.y), p.distY));
    p.coefIdxI = mul(t.r0, p.delta.x);
    p.coefIdxJ = mul(t.r1, p.delta.y);
    p.transl = add(add(add(gtii.v, mul(t.r2, p.corner.z)), mul(mul(t.r0, p.corner.x), 1.0f - p.corner.z / p.distX)),
                   mul(mul(t.r1, p.corner.y), 1.0f - p.corner.z / p.distY));
    return p;
}

struct FillGeom {             // geometry part of FillIddAndSigmaParams, fill_idd_and_sigma_params.cu:12-26
    V3 corner, delta; float distX, distY, volConst, volLin, volSq;
};
inline FillGeom makeFillGeom(const IdxT& fitf, const float dist[2]) {
    FillGeom g; g.corner = fitf.offset; g.delta = fitf.delta; g.distX = dist[0]; g.distY = dist[1];
    const V3 c = g.corner, d = g.delta;
    const float a = std::fabs(d.x * d.y * d.z);
    g.volConst = a * (1.0f - c.z / g.distX - c.z / g.distY + (c.z * c.z + d.z * d.z / 12.0f) / (g.distX * g.distY));
    g.volLin = a * d.z * (-1.0f / g.distX - 1.0f / g.distY + 2.0f * c.z / (g.distX * g.distY));
    g.volSq = a * d.z * d.z / (g.distX * g.distY);
    return g;
}
// fill_idd_and_sigma_params.cu:74-83 evaluated on the HOST (kernel_wrapper.cu:838): IEEE division.
inline void sigmaSqAirCoefsHost(float r0, bool nozzle, float& a, float& b) {
    if (nozzle) { a = 0.00270f / (r0 - 4.50f); b = -4.39f / (r0 - 3.86f); } else { a = 0.0f; b = 0.0f; }
}

struct TransferParams {       // TransferParamStructDiv3, transfer_param_struct_div3.cuh:39-45
    V3 globalOffset, coefOffset, coefIdxI, coefIdxJ, inc; float normDistX, normDistY;
};
// rayIdxToDoseIdx.invertAndShift(shift) -> TransferParamStructDiv3 (from_fan.cu:27-30, transfer_param_struct_div3.cu:9-20)
inline TransferParams makeTransferParams(const IdxT& fitf, const float dist[2], const Affine& gtdi, V3 shift) {
    IdxT ftfi = inverse(fitf); ftfi.offset = add(ftfi.offset, shift);
    const Affine iitg = inverse(gtdi);
    const M3 t = transpose(iitg.m);
    TransferParams p;
    p.coefIdxI = mul(t.r0, ftfi.delta);
    p.coefIdxJ = mul(t.r1, ftfi.delta);
    p.coefOffset = mul(iitg.v, ftfi.delta);
    p.globalOffset = ftfi.offset;
    p.inc = mul(t.r2, ftfi.delta);
    p.normDistX = ftfi.delta.z * dist[0];
    p.normDistY = ftfi.delta.z * dist[1];
    return p;
}

struct RayGrid { int lSteps, bSteps, rSteps, tSteps, dimX, dimY; IdxT rayIdxToGantry; };
// kernel_wrapper.cu:612-663
inline RayGrid makeRayGrid(const float spotDelta[3], const float spotOffset[3], int sx, int sy, const float raySpacing[2],
                           float maxSigX, float maxSigY, float convCutoff) {
    RayGrid g;
    const V3 res = v3(raySpacing[0], raySpacing[1], spotDelta[2]);
    g.lSteps = int(std::ceil((spotOffset[0] - (convCutoff * maxSigX + 0.5f * res.x)) / res.x));
    g.bSteps = int(std::ceil((spotOffset[1] - (convCutoff * maxSigY + 0.5f * res.y)) / res.y));
    g.rSteps = int(std::floor(((unsigned(sx) - 1) * spotDelta[0] + spotOffset[0] + (convCutoff * maxSigX + 0.5f * res.x)) / res.x));
    g.tSteps = int(std::floor(((unsigned(sy) - 1) * spotDelta[1] + spotOffset[1] + (convCutoff * maxSigY + 0.5f * res.y)) / res.y));
    g.rayIdxToGantry.delta = res;
    g.rayIdxToGantry.offset = v3(res.x * g.lSteps, res.y * g.bSteps, spotOffset[2]);
    g.dimX = roundTo(g.rSteps - g.lSteps + 1, 32);
    g.dimY = roundTo(g.tSteps - g.bSteps + 1, 8);
    return g;
}

}  // namespace rtd
