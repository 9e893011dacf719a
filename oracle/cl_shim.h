// cl_shim.h -- TEST INFRASTRUCTURE.  The minimum of OpenCL C that the reference's kernel source
// (RayTrace/resources/shaders/rayTracer.cl) needs to compile as C++ on the host: the float4 vector
// type with the operators the kernel uses, dot(), the address-space / kernel qualifiers and
// get_global_id().  oracle/build_ref.py compiles the .cl file where it lies against this header into
// oracle/_ref/libraytracer_cl_ref.so (no OpenCL ICD, CL headers or clew exist in this image: "OpenCL
// ICD unavailable on this box").  Choices OpenCL leaves to the implementation, fixed here and stated
// in DESIGN.md: no FMA contraction (built -ffp-contract=off), dot() summed left to right, sqrt
// correctly rounded, float -> int conversion by truncation (C semantics).
#ifndef RTB_ORACLE_CL_SHIM_H_
#define RTB_ORACLE_CL_SHIM_H_

#include <cmath>
#include <cstddef>

struct float4 {
    float x, y, z, w;
};
inline float4 make_float4(float x, float y, float z, float w) { float4 r = {x, y, z, w}; return r; }
inline float4 operator-(const float4& a, const float4& b) { return make_float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
inline float4 operator*(float s, const float4& a) { return make_float4(s * a.x, s * a.y, s * a.z, s * a.w); }
inline float dot(const float4& a, const float4& b) { return ((a.x * b.x + a.y * b.y) + a.z * b.z) + a.w * b.w; }
inline float cl_sqrtf(float x) { return std::sqrt(x); }

#define __kernel
#define __global
#define sqrt(x) cl_sqrtf(x)

extern thread_local size_t cl_shim_global_id;
inline size_t get_global_id(int) { return cl_shim_global_id; }

#endif
