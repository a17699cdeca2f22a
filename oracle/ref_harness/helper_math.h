// Minimal stand-in for the CUDA-samples header "helper_math.h", which the reference includes
// (/root/reference/src/kernel_wrapper.cu:11 and the float3_* / matrix_3x3 sources) but which is
// not installed in this image. Only the float3 operators the reference actually uses are provided,
// written as the obvious component-wise expressions. TEST INFRASTRUCTURE ONLY: used to rebuild the
// unmodified reference into oracle/_ref/ (see oracle/ref_harness/Makefile); never linked into the product.
#ifndef RTD_REF_HELPER_MATH_SHIM_H
#define RTD_REF_HELPER_MATH_SHIM_H
#include <cuda_runtime.h>
#include <cmath>
#include <algorithm>
#ifndef __CUDACC__
using std::min;
using std::max;
#endif
#define HM_FN inline __host__ __device__
HM_FN float3 make_float3(float s) { return make_float3(s, s, s); }
HM_FN float3 operator-(const float3 &a) { return make_float3(-a.x, -a.y, -a.z); }
HM_FN float3 operator+(float3 a, float3 b) { return make_float3(a.x + b.x, a.y + b.y, a.z + b.z); }
HM_FN float3 operator+(float3 a, float b) { return make_float3(a.x + b, a.y + b, a.z + b); }
HM_FN float3 operator+(float b, float3 a) { return make_float3(a.x + b, a.y + b, a.z + b); }
HM_FN void operator+=(float3 &a, float3 b) { a.x += b.x; a.y += b.y; a.z += b.z; }
HM_FN void operator+=(float3 &a, float b) { a.x += b; a.y += b; a.z += b; }
HM_FN float3 operator-(float3 a, float3 b) { return make_float3(a.x - b.x, a.y - b.y, a.z - b.z); }
HM_FN float3 operator-(float3 a, float b) { return make_float3(a.x - b, a.y - b, a.z - b); }
HM_FN float3 operator-(float b, float3 a) { return make_float3(b - a.x, b - a.y, b - a.z); }
HM_FN void operator-=(float3 &a, float3 b) { a.x -= b.x; a.y -= b.y; a.z -= b.z; }
HM_FN void operator-=(float3 &a, float b) { a.x -= b; a.y -= b; a.z -= b; }
HM_FN float3 operator*(float3 a, float3 b) { return make_float3(a.x * b.x, a.y * b.y, a.z * b.z); }
HM_FN float3 operator*(float3 a, float b) { return make_float3(a.x * b, a.y * b, a.z * b); }
HM_FN float3 operator*(float b, float3 a) { return make_float3(b * a.x, b * a.y, b * a.z); }
HM_FN void operator*=(float3 &a, float3 b) { a.x *= b.x; a.y *= b.y; a.z *= b.z; }
HM_FN void operator*=(float3 &a, float b) { a.x *= b; a.y *= b; a.z *= b; }
HM_FN float3 operator/(float3 a, float3 b) { return make_float3(a.x / b.x, a.y / b.y, a.z / b.z); }
HM_FN float3 operator/(float3 a, float b) { return make_float3(a.x / b, a.y / b, a.z / b); }
HM_FN float3 operator/(float b, float3 a) { return make_float3(b / a.x, b / a.y, b / a.z); }
HM_FN void operator/=(float3 &a, float3 b) { a.x /= b.x; a.y /= b.y; a.z /= b.z; }
HM_FN void operator/=(float3 &a, float b) { a.x /= b; a.y /= b; a.z /= b; }
HM_FN float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
#undef HM_FN
#endif
