// TEST INFRASTRUCTURE -- not part of the product path.
//
// OpenCL-C-as-C++17 shim: the minimum vocabulary needed to compile the
// reference's own kernel sources (shade_and_reflect_kernel.cl, shade_kernel.cl,
// hittest_kernel.cl) for the host CPU, unchanged apart from three mechanical
// `sed` rewrites applied at build time to a temporary copy (see oracle/Makefile):
//     \.xyz\b      -> .xyz()
//     (float4)(    -> mk_float4(
//     (float3)(    -> mk_float3(
// The result (oracle/_ref/libsar_ref.so) IS the reference's code for the path and
// serves as the parity oracle and as the timed CPU baseline.
//
// Choices the OpenCL spec leaves to the implementation, fixed here and mirrored
// by the CUDA path:  dot() sums left to right without contraction;
// normalize(v) = v / sqrt(dot(v,v));  sqrt and '/' are IEEE correctly rounded;
// pow is the C library's powf.  Build with -ffp-contract=off.
#pragma once

#include <cmath>
#include <cstdint>

typedef unsigned int uint;
typedef unsigned long ulong;

struct float3;

struct float4 {
    float x, y, z, w;
    inline float3 xyz() const;
    float4& operator+=(const float4& o) {
        x += o.x; y += o.y; z += o.z; w += o.w;
        return *this;
    }
};

// OpenCL float3 occupies 16 bytes (OpenCL 1.2 s6.1.5); pixelData has 16-byte stride.
struct alignas(16) float3 {
    float x, y, z;
    float3 xyz() const { return *this; }
    float3& operator+=(const float3& o) {
        x += o.x; y += o.y; z += o.z;
        return *this;
    }
};

inline float3 float4::xyz() const { return float3{x, y, z}; }

struct alignas(64) float16 {
    float s0, s1, s2, s3, s4, s5, s6, s7, s8, s9, sA, sB, sC, sD, sE, sF;
};

inline float4 mk_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
inline float4 mk_float4(const float3& v, float w) { return float4{v.x, v.y, v.z, w}; }
inline float3 mk_float3(float x, float y, float z) { return float3{x, y, z}; }

inline float3 operator+(const float3& a, const float3& b) { return float3{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline float3 operator-(const float3& a, const float3& b) { return float3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline float3 operator-(const float3& a) { return float3{-a.x, -a.y, -a.z}; }
inline float3 operator*(const float3& a, float s) { return float3{a.x * s, a.y * s, a.z * s}; }
inline float3 operator*(float s, const float3& a) { return float3{s * a.x, s * a.y, s * a.z}; }

inline float4 operator+(const float4& a, const float4& b) { return float4{a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
inline float4 operator-(const float4& a, const float4& b) { return float4{a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w}; }
inline float4 operator*(const float4& a, float s) { return float4{a.x * s, a.y * s, a.z * s, a.w * s}; }
inline float4 operator*(float s, const float4& a) { return float4{s * a.x, s * a.y, s * a.z, s * a.w}; }

inline float dot(const float3& a, const float3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float dot(const float4& a, const float4& b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
inline float3 normalize(const float3& v) {
    const float len = std::sqrt(dot(v, v));
    return float3{v.x / len, v.y / len, v.z / len};
}
inline float4 normalize(const float4& v) {
    const float len = std::sqrt(dot(v, v));
    return float4{v.x / len, v.y / len, v.z / len, v.w / len};
}

using std::copysign;
using std::fmax;
using std::fmin;
using std::pow;
using std::sqrt;

#define __kernel
#define __global

// One work-item at a time per host thread.
extern thread_local long cl_shim_global_id;
inline long get_global_id(int) { return cl_shim_global_id; }
