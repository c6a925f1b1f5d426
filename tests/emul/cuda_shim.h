// TEST INFRASTRUCTURE -- not part of the product path.
//
// Lets the device headers of the CUDA path (rt_device.cuh, rt_pixel.cuh) compile as plain host C++ so that
// the kernels' logic -- conservative filters, BVH traversal, the per-pixel state machine -- can be checked
// against the oracle in the CPU-only test tier, without a GPU.  Built with -ffp-contract=off so that the
// __f*_rn stand-ins below are unfused, exactly like the intrinsics they replace.  The product library
// (librt_b200.so) never includes this file.
#pragma once
#define RT_HOST_EMULATION 1

#include <math.h>
#include <stdint.h>
#include <string.h>

#define __device__
#define __host__
#define __global__
#define __forceinline__ inline

struct float2 {
    float x, y;
};
struct alignas(16) float4 {
    float x, y, z, w;
};
static inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }

template <typename T>
static inline T __ldg(const T* p) {
    return *p;
}
static inline float __fmul_rn(float a, float b) { return a * b; }
static inline float __fadd_rn(float a, float b) { return a + b; }
static inline float __fsub_rn(float a, float b) { return a - b; }
static inline float __fdiv_rn(float a, float b) { return a / b; }
static inline float __fsqrt_rn(float a) { return sqrtf(a); }
static inline int __float_as_int(float f) {
    int i;
    memcpy(&i, &f, 4);
    return i;
}
static inline unsigned __float_as_uint(float f) {
    unsigned i;
    memcpy(&i, &f, 4);
    return i;
}
static inline float __int_as_float(int i) {
    float f;
    memcpy(&f, &i, 4);
    return f;
}

typedef int cudaError_t;
typedef void* cudaStream_t;
